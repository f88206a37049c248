// libqmmx_mc.so — C ABI (include/qmmx_mc.h) over the sm_100a kernels in kernels.cuh.
// One ctx = one GPU = one compute stream (+ one copy stream for the host-bar pipeline).  No CPU fallback.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "scan_lean.cuh"
#include "scan_tile.cuh"
#include "tables_host.h"

using namespace qmmx;

struct qmmx_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr, own_stream = nullptr, copy_stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_copied[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
    std::string err;
    uint64_t launches = 0;
    // host-side configuration
    int L = 0;
    double price[kMaxLevels];
    uint8_t solid[kMaxLevels];
    std::vector<qmmx_params> params;
    std::vector<double> lr; int n_lr = 0;
    std::vector<double> pol; int n_pol = 0;
    double hist_lo = -8.0, hist_hi = 8.0;
    double z_go = 0.0, z_skip = 0.0;
    bool dirty = true;
    // derived device state
    DevTables* d_tables = nullptr; int n_tables = 0; int max_ncell = 1;
    struct CellGrid { int32_t obase, shift, ncell; float guard[4]; int32_t ncls; int32_t n_out; float grid_lo, grid_hi, out_lo[4], out_hi[4]; uint32_t out_cls[4]; };
    std::vector<CellGrid> grids;
    std::vector<DevCombo> h_combos;
    uint32_t* d_mlmask = nullptr; double* d_mlprob = nullptr; double* d_pol = nullptr;
    unsigned long long* d_flat = nullptr; int flat_combos = 0;
    unsigned int* d_counters = nullptr; int counter_next = 0;      // scan_tile.cuh tile counters, one per launch of a run
    bool any_hl = false, any_tiny = false;
    bool no_sweep = false;     // QMMX_NO_SWEEP=1: one launch per combo even for sweep-shaped parameter sets
    bool lean_ok = false;      // scan_lean.cuh's sign-encoded gate table is valid for these levels
    bool force_v1 = false;     // QMMX_SCAN_V1=1: run the first-generation lockstep kernel (A/B measurements)
    bool no_tile = false;      // QMMX_NO_TILE=1: keep close-only scans on the lockstep lean kernel (A/B measurements)
    bool no_tma = false;       // QMMX_NO_TMA=1: stage the tile kernel's rows with cp.async instead of TMA (A/B measurements)
    void* encode_tiled = nullptr;   // cuTensorMapEncodeTiled, through cudaGetDriverEntryPoint (no libcuda link)
    // resident bars
    float *d_close = nullptr, *d_high = nullptr, *d_low = nullptr;
    float* d_vol = nullptr;                 // per-bar volume of the resident bars (library-owned), or null
    uint8_t* d_regime = nullptr;            // AccumulationBreakoutGuard regime per tick (regime_kernel), lives with d_vol
    std::vector<float> boot_vol;            // window volumes for the next qmmx_gen_bootstrap
    bool own_bars = false;
    int64_t paths = 0, pitch = 0; int32_t steps = 0;
    // staging for qmmx_run_scan_host
    float* d_stage[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
    int64_t stage_elems = 0;
    double* d_polw = nullptr; int64_t polw_pitch = 0, polw_paths = 0;   // QMMX_F_ONLINE_RETRAIN: learnt heads [14][pitch]
    float* d_ratios = nullptr; size_t ratios_cap = 0;     // bootstrap return ratios (qmmx_gen_bootstrap)
    bool ran = false;          // some kernel of this ctx has run (qmmx_last_timing)
    bool agg_ran = false;      // d_flat holds the aggregates of a scan / fused run ...
    int agg_combos = 0;        // ... of this many combos (qmmx_fetch_agg, qmmx_agg_device, qmmx_merge_flat check it)
    // NCCL through dlopen (qmmx_allreduce)
    void* nccl_comm = nullptr;             // communicator made by qmmx_nccl_init (owned)
    unsigned long long* d_gather = nullptr; size_t gather_slots = 0;
    // qmmx_run_last_bars_static: device address of g_static_scratch on this ctx's device (resolved in qmmx_create)
    uint8_t* static_scratch = nullptr;
    cudaError_t static_scratch_err = cudaErrorUnknown;
};

namespace {

int fail(qmmx_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg;
    return code;
}

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess)                                                                       \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? QMMX_E_NOMEM : QMMX_E_CUDA,           \
                        std::string(#call) + ": " + cudaGetErrorString(e_));                         \
    } while (0)

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

void free_volume(qmmx_ctx* c) {
    if (c->d_vol) cudaFree(c->d_vol);
    if (c->d_regime) cudaFree(c->d_regime);
    c->d_vol = nullptr; c->d_regime = nullptr;
}

void free_bars(qmmx_ctx* c) {
    free_volume(c);
    if (c->own_bars) { cudaFree(c->d_close); cudaFree(c->d_high); cudaFree(c->d_low); }
    c->d_close = c->d_high = c->d_low = nullptr;
    c->own_bars = false; c->paths = 0; c->pitch = 0; c->steps = 0;
}

int alloc_bars(qmmx_ctx* ctx, int64_t paths, int32_t steps, bool hl) {
    // same shape as the resident buffers: reuse them (a generator called once per step must not pay cudaMalloc/cudaFree)
    if (ctx->own_bars && ctx->d_close && ctx->paths == paths && ctx->steps == steps && (ctx->d_high != nullptr) == hl) return QMMX_OK;
    free_bars(ctx);
    const int64_t pitch = round_up(std::max<int64_t>(paths, 1), 32);
    const size_t bytes = (size_t)pitch * (size_t)std::max(steps, 1) * sizeof(float);
    CK(cudaMalloc(&ctx->d_close, bytes));
    ctx->own_bars = true;
    if (hl) { CK(cudaMalloc(&ctx->d_high, bytes)); CK(cudaMalloc(&ctx->d_low, bytes)); }
    ctx->paths = paths; ctx->pitch = pitch; ctx->steps = steps;
    return QMMX_OK;
}

template <class T>
int upload_vec(qmmx_ctx* ctx, T** dptr, const T* host, size_t n) {
    if (*dptr) { cudaFree(*dptr); *dptr = nullptr; }
    if (n == 0) return QMMX_OK;
    CK(cudaMalloc(dptr, n * sizeof(T)));
    CK(cudaMemcpyAsync(*dptr, host, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return QMMX_OK;
}

size_t scan_smem_bytes(const qmmx_ctx* c) {      // dynamic part only: the grid-cell bytes
    return (size_t)round_up(c->max_ncell, 16);
}

// (re)build every derived table after a configuration change
int prepare(qmmx_ctx* ctx) {
    if (!ctx->dirty) return QMMX_OK;
    if (ctx->params.empty()) return fail(ctx, QMMX_E_STATE, "qmmx_set_params has not been called");
    const int nc = (int)ctx->params.size();
    std::vector<DevTables> tabs;
    std::vector<std::pair<double, double>> keys;
    std::vector<DevCombo> combos(nc);
    ctx->any_hl = false; ctx->any_tiny = false;
    bool any_prob = false;
    for (int c = 0; c < nc; ++c) {
        const qmmx_params& P = ctx->params[c];
        int t = -1;
        for (size_t i = 0; i < keys.size(); ++i)
            if (std::memcmp(&keys[i].first, &P.contact_prox, 8) == 0 && std::memcmp(&keys[i].second, &P.q_min_prob, 8) == 0) t = (int)i;
        if (t < 0) {
            if ((int)keys.size() >= kMaxTables)
                return fail(ctx, QMMX_E_UNSUPPORTED, "more than 8 distinct (contact_prox, q_min_prob) pairs in one run");
            tabs.emplace_back();
            std::string e;
            if (build_tables(ctx->L, ctx->price, ctx->solid, P.contact_prox, P.q_min_prob, &tabs.back(), &e) != 0)
                return fail(ctx, QMMX_E_UNSUPPORTED, e);
            keys.emplace_back(P.contact_prox, P.q_min_prob);
            t = (int)keys.size() - 1;
        }
        DevCombo& D = combos[c];
        std::memset(&D, 0, sizeof(D));
        D.sp = P.stop_padding; D.tp = P.tp_padding; D.qmin = P.q_min_prob; D.prox = P.contact_prox;
        D.pmax = 0.0001 > P.contact_prox ? 0.0001 : P.contact_prox;
        D.vstrong = P.veto_vol_strong; D.vprox = P.veto_prox;
        double wr = P.w_rules, wm = P.w_ml, s = wr + wm;                // M:1714-1719
        if (s <= 0) { wr = 1.0; wm = 0.0; s = 1.0; }
        D.wr = wr / s; D.wm = wm / s;
        // first step after an exit at step i that may evaluate an entry: i + ceil(cooldown_s / 60), and never the exit
        // tick itself (evaluate_entry ran before the exit, M:2936 / M:2966), so a zero cooldown counts as one step
        D.cool_steps = P.cooldown_s <= 0 ? 1 : (P.cooldown_s + 59) / 60;
        D.overtouch = P.overtouch_limit; D.spd = P.steps_per_day; D.table = t;
        D.flags = P.flags & 0xffffu;
        const bool ml_on = !(P.flags & QMMX_F_DISABLE_ML_GATE) && ctx->n_lr > 0;
        if (ml_on) D.flags |= kCbHasMl;
        if (ml_on && (P.flags & QMMX_F_USE_BLEND)) { D.flags |= kCbHasProb; any_prob = true; }
        if (P.flags & QMMX_F_EXIT_ON_HL) ctx->any_hl = true;
        if (P.flags & QMMX_F_TINY_PRICES) ctx->any_tiny = true;
    }
    for (int c = 1; c < nc; ++c)
        if (((ctx->params[c].flags ^ ctx->params[0].flags) & (QMMX_F_EXIT_ON_HL | QMMX_F_TINY_PRICES)) != 0)
            return fail(ctx, QMMX_E_UNSUPPORTED, "EXIT_ON_HL / TINY_PRICES must be the same for all combos of a run");
    ctx->n_tables = (int)tabs.size();
    ctx->max_ncell = 1;
    ctx->grids.clear();
    for (auto& t : tabs) { ctx->max_ncell = std::max(ctx->max_ncell, t.ncell); qmmx_ctx::CellGrid G{t.obase, t.shift, t.ncell, {t.guard[0], t.guard[1], t.guard[2], t.guard[3]}, t.ncls, t.n_out, t.grid_lo, t.grid_hi, {}, {}, {}};
        for (int j = 0; j < kMaxOut; ++j) { G.out_lo[j] = t.out_lo[j]; G.out_hi[j] = t.out_hi[j]; G.out_cls[j] = t.out_cls[j]; }
        ctx->grids.push_back(G); }
    ctx->lean_ok = true;       // every non-empty confidence interval must have a positive upper bound (sign encoding)
    for (auto& t : tabs) if (t.ncls < 0) ctx->lean_ok = false;      // and the class partition must fit one byte (tables.h)
    for (auto& t : tabs)
        for (int k = 0; k < t.L; ++k)
            for (int b = 0; b < 3; ++b)
                if (t.lev[k].cpass[b][0] <= t.lev[k].cpass[b][1] && !(t.lev[k].cpass[b][1] > 0.0f)) ctx->lean_ok = false;
    int rc;
    if ((rc = upload_vec(ctx, &ctx->d_tables, tabs.data(), tabs.size())) != 0) return rc;
    ctx->h_combos = combos;
    // logistic gates
    if (ctx->d_mlmask) { cudaFree(ctx->d_mlmask); ctx->d_mlmask = nullptr; }
    if (ctx->d_mlprob) { cudaFree(ctx->d_mlprob); ctx->d_mlprob = nullptr; }
    if (ctx->n_lr > 0) {
        const size_t per_mask = (size_t)ctx->n_lr * kMaxLevels, per_prob = per_mask * 32;
        if (any_prob && per_prob * nc * sizeof(double) > (size_t)512 << 20)
            return fail(ctx, QMMX_E_UNSUPPORTED, "blend-mode probability table exceeds 512 MiB (combos x days)");
        std::vector<uint32_t> mask(per_mask * nc, 0u);
        std::vector<double> prob(any_prob ? per_prob * nc : 0);
        for (int c = 0; c < nc; ++c) {
            if (!(combos[c].flags & kCbHasMl)) continue;
            build_ml_tables(ctx->L, ctx->price, ctx->params[c].stop_padding, ctx->params[c].q_min_prob, ctx->lr.data(),
                            ctx->n_lr, mask.data() + per_mask * c, any_prob ? prob.data() + per_prob * c : nullptr);
        }
        if ((rc = upload_vec(ctx, &ctx->d_mlmask, mask.data(), mask.size())) != 0) return rc;
        if (any_prob && (rc = upload_vec(ctx, &ctx->d_mlprob, prob.data(), prob.size())) != 0) return rc;
    }
    {   // policy table [n_pol][3][7], followed by the sum of |w| per (day, head): the error bound of the gate's fast path (engine.cuh)
        std::vector<double> pol(ctx->pol);
        for (int d = 0; d < ctx->n_pol; ++d)
            for (int h = 0; h < QMMX_POLICY_HEADS; ++h) {
                double sa = 0.0;
                for (int q = 0; q < QMMX_POLICY_DIM; ++q) sa += std::fabs(ctx->pol[((size_t)d * QMMX_POLICY_HEADS + h) * QMMX_POLICY_DIM + q]);
                pol.push_back(sa);
            }
        if ((rc = upload_vec(ctx, &ctx->d_pol, pol.data(), pol.size())) != 0) return rc;
    }
    if (ctx->flat_combos < nc) {
        if (ctx->d_flat) cudaFree(ctx->d_flat);
        ctx->d_flat = nullptr;
        CK(cudaMalloc(&ctx->d_flat, (size_t)nc * QMMX_AGG_SLOTS * 8));
        ctx->flat_combos = nc;
    }
    ctx->dirty = false;
    return QMMX_OK;
}

// kernel parameters of one launch: combo `c` over the given bars (parity outputs are set by the caller)
void fill_args(qmmx_ctx* ctx, ScanArgs& A, int c, const float* cl, const float* h, const float* l, int64_t pitch,
               int64_t paths, int32_t steps) {
    std::memset(&A, 0, sizeof(A));
    const DevCombo& D = ctx->h_combos[c];
    A.close = cl; A.high = h; A.low = l;
    A.vol = (cl && cl == ctx->d_close && (D.flags & QMMX_F_ENABLE_VETO)) ? ctx->d_vol : nullptr;
    A.regime = (cl && cl == ctx->d_close && ctx->d_vol) ? ctx->d_regime : nullptr;
    A.pitch = pitch; A.paths = paths; A.steps = steps; A.combo_index = c;
    A.tables = ctx->d_tables + D.table;
    if (ctx->d_counters && ctx->counter_next == kTileCounters) {      // a run with more launches than counters: recycle them, in stream order
        cudaMemsetAsync(ctx->d_counters, 0, kTileCounters * sizeof(unsigned int), ctx->stream);
        ctx->counter_next = 0;
    }
    A.tile_counter = ctx->d_counters ? ctx->d_counters + ctx->counter_next++ : nullptr;
    A.flat = ctx->d_flat + (size_t)c * QMMX_AGG_SLOTS;
    const size_t per_mask = (size_t)ctx->n_lr * kMaxLevels;
    A.mlmask = (ctx->d_mlmask && (D.flags & kCbHasMl)) ? ctx->d_mlmask + per_mask * c : nullptr;
    A.mlprob = (ctx->d_mlprob && (D.flags & kCbHasProb)) ? ctx->d_mlprob + per_mask * 32 * c : nullptr;
    A.pol = ctx->d_pol; A.n_pol = ctx->n_pol; A.n_days = ctx->n_lr;
    A.z_go = ctx->z_go; A.z_skip = ctx->z_skip;
    A.hist_lo = ctx->hist_lo; A.hist_inv_w = (double)QMMX_PNL_HIST_BINS / (ctx->hist_hi - ctx->hist_lo);
    A.n_levels = ctx->L;
    A.cell_obase = ctx->grids[D.table].obase; A.cell_shift = ctx->grids[D.table].shift; A.cell_ncell = ctx->grids[D.table].ncell;
    A.cell_lim = ((uint32_t)A.cell_ncell << A.cell_shift) - 1u;
    A.cell_bytes = (int32_t)scan_smem_bytes(ctx);
    for (int g = 0; g < 4; ++g) A.guard[g] = ctx->grids[D.table].guard[g];
    A.mix_base = (uint32_t)std::max(0, ctx->grids[D.table].ncls);
    {
        const qmmx_ctx::CellGrid& G = ctx->grids[D.table];
        A.n_out = G.n_out; A.grid_lo = G.grid_lo; A.grid_hi = G.grid_hi;
        for (int j = 0; j < kMaxOut; ++j) { A.out_lo[j] = G.out_lo[j]; A.out_hi[j] = G.out_hi[j]; A.out_cls[j] = G.out_cls[j]; }
    }
    A.combo = D;
}

dim3 scan_grid(const qmmx_ctx* ctx, int64_t paths) {
    const int64_t ntiles = (paths + 31) / 32;
    const int64_t want = (ntiles + kWarps - 1) / kWarps;
    const int64_t cap = (int64_t)ctx->sm_count * kBlocksPerSM;
    return dim3((unsigned)std::max<int64_t>(1, std::min(want, cap)), 1, 1);
}

template <class K>
int set_smem_attr(qmmx_ctx* ctx, K kernel, size_t bytes) {
    CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return QMMX_OK;
}

// the regime plane follows the resident close + volume planes (the reference's guard is always on; without volume it
// can never leave {unknown, accumulation}, which block nothing)
int compute_regime(qmmx_ctx* ctx) {
    if (!ctx->d_vol || !ctx->d_close || ctx->paths == 0 || ctx->steps == 0) return QMMX_OK;
    if (!ctx->d_regime) CK(cudaMalloc(&ctx->d_regime, (size_t)ctx->pitch * (size_t)ctx->steps));
    const int64_t blocks = std::min<int64_t>((ctx->paths + kThreads - 1) / kThreads, (int64_t)ctx->sm_count * 8);
    regime_kernel<<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(ctx->d_close, ctx->d_vol, ctx->d_regime, ctx->pitch, ctx->paths, ctx->steps);
    ctx->launches++;
    CK(cudaGetLastError());
    return QMMX_OK;
}

int init_flat(qmmx_ctx* ctx) {
    const int nc = (int)ctx->params.size();
    if (!ctx->d_counters) CK(cudaMalloc(&ctx->d_counters, kTileCounters * sizeof(unsigned int)));
    ctx->counter_next = 0;
    init_flat_kernel<<<std::max(1, (nc * QMMX_AGG_SLOTS + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_flat, nc, ctx->d_counters);
    ctx->agg_ran = true; ctx->agg_combos = nc;      // every scan / fused entry point starts here
    ctx->launches++;
    CK(cudaGetLastError());
    return QMMX_OK;
}

// tensor map of one fp32 plane [steps][paths] (row pitch in elements) for the tile kernel's TMA staging
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
bool make_bar_tmap(const qmmx_ctx* ctx, CUtensorMap* map, const float* plane, int64_t pitch, int64_t paths, int32_t steps) {
    if (!ctx->encode_tiled || ctx->no_tma || (pitch * 4) % 16 != 0 || ((uintptr_t)plane & 15) != 0 || paths < 1 || steps < 1) return false;
    const cuuint64_t dim[2] = {(cuuint64_t)paths, (cuuint64_t)steps};
    const cuuint64_t stride[1] = {(cuuint64_t)pitch * 4};
    const cuuint32_t box[2] = {32u, (cuuint32_t)kChunk}, estr[2] = {1u, 1u};
    return reinterpret_cast<EncodeTiledFn>(ctx->encode_tiled)(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(plane), dim,
                                                              stride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// ---- NCCL, loaded at run time -------------------------------------------------------------------------------------
struct NcclId { char internal[128]; };
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*CommCount)(void*, int*) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;

int load_nccl(qmmx_ctx* ctx) {
    if (g_nccl.lib) return QMMX_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);      // the copy the process already uses (torch's)
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW);
    if (!h) return fail(ctx, QMMX_E_NCCL, std::string("dlopen libnccl.so.2: ") + dlerror());
    NcclApi a;
    a.lib = h;
    a.GetUniqueId = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<int (*)(void**, int, NcclId, int)>(dlsym(h, "ncclCommInitRank"));
    a.CommCount = reinterpret_cast<int (*)(void*, int*)>(dlsym(h, "ncclCommCount"));
    a.AllGather = reinterpret_cast<int (*)(const void*, void*, size_t, int, void*, cudaStream_t)>(dlsym(h, "ncclAllGather"));
    a.CommDestroy = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclCommDestroy"));
    a.GetErrorString = reinterpret_cast<const char* (*)(int)>(dlsym(h, "ncclGetErrorString"));
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommCount || !a.AllGather || !a.CommDestroy || !a.GetErrorString)
        return fail(ctx, QMMX_E_NCCL, "libnccl.so.2 lacks a required symbol");
    g_nccl = a;
    return QMMX_OK;
}

int nccl_fail(qmmx_ctx* ctx, const char* what, int rc) {
    return fail(ctx, QMMX_E_NCCL, std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?"));
}

// aggregates of a scan / fused run of the CURRENT parameter set are on the device
int check_agg(qmmx_ctx* ctx) {
    if (!ctx->agg_ran || !ctx->d_flat) return fail(ctx, QMMX_E_STATE, "no scan or fused run has produced aggregates");
    if (ctx->dirty || ctx->agg_combos != (int)ctx->params.size())
        return fail(ctx, QMMX_E_STATE, "levels / params / models changed since the last run: run again before reading aggregates");
    return QMMX_OK;
}

struct ParityOut { qmmx_trade* trades = nullptr; int32_t* tcounts = nullptr; uint8_t* codes = nullptr; int32_t max_trades = 0; };

// one launch per parameter combo (run constants travel as kernel parameters); parity outputs belong to combo 0
int launch_scan(qmmx_ctx* ctx, const float* cl, const float* h, const float* l, int64_t pitch, int64_t paths,
                int32_t steps, const ParityOut* par) {
    const int nc = (int)ctx->params.size();
    const size_t smem = scan_smem_bytes(ctx);
    const dim3 grid = scan_grid(ctx, paths);
    const bool hl = ctx->any_hl, tiny = ctx->any_tiny;
    if (hl && (!h || !l)) return fail(ctx, QMMX_E_STATE, "EXIT_ON_HL needs high/low bars");
    if (paths <= 0 || steps <= 0) return QMMX_OK;
    if (steps > 65535) return fail(ctx, QMMX_E_BADARG, "steps per path must be <= 65535 (16-bit per-path tallies)");
    int rc = QMMX_OK;
#define LAUNCH2(MDv, HLv, PARv, TINYv, Wv)                                                    \
    do {                                                                                      \
        auto k = scan_kernel<MDv, HLv, PARv, TINYv, Wv>;                                      \
        if ((rc = set_smem_attr(ctx, k, smem)) != 0) return rc;                               \
        k<<<grid, kThreads, smem, ctx->stream>>>(A);                                          \
    } while (0)
#define LAUNCH(HLv, PARv, TINYv)                                                              \
    do {                                                                                      \
        if (multiday) { if (wide) LAUNCH2(true, HLv, PARv, TINYv, true); else LAUNCH2(true, HLv, PARv, TINYv, false); } \
        else { if (wide) LAUNCH2(false, HLv, PARv, TINYv, true); else LAUNCH2(false, HLv, PARv, TINYv, false); }        \
    } while (0)
    const bool multiday = ctx->n_lr > 1;
    const bool wide = ctx->L > 32;
    const bool lean = ctx->lean_ok && !tiny && !ctx->force_v1;
    const size_t lean_smem = lean_smem_bytes(ctx->L);
    const dim3 lean_grid((unsigned)std::max<int64_t>(1, std::min<int64_t>(((paths + 31) / 32 + kLeanWarps - 1) / kLeanWarps,
                                                                            (int64_t)ctx->sm_count * kLeanBlocksPerSM)), 1, 1);
#define LEAN3(MDv, HLv, PARv, Wv, SLv)                                                        \
    do {                                                                                      \
        auto k = lean_scan_kernel<MDv, HLv, PARv, Wv, SLv>;                                   \
        if ((rc = set_smem_attr(ctx, k, lean_smem)) != 0) return rc;                          \
        k<<<lean_grid, kLeanThreads, lean_smem, ctx->stream>>>(A);                            \
    } while (0)
#define LEAN3L(MDv, HLv, PARv, Wv)                                                            \
    do {                                                                                      \
        auto k = lean_scan_kernel<MDv, HLv, PARv, Wv, true, true>;                            \
        if ((rc = set_smem_attr(ctx, k, lean_smem)) != 0) return rc;                          \
        k<<<lean_grid, kLeanThreads, lean_smem, ctx->stream>>>(A);                            \
    } while (0)
#define LEAN2(MDv, HLv, PARv, Wv)                                                             \
    do {                                                                                      \
        if (A.polw) LEAN3L(MDv, HLv, PARv, Wv);                                               \
        else if (PARv || slow) LEAN3(MDv, HLv, PARv, Wv, true);                               \
        else LEAN3(MDv, HLv, false, Wv, false);                                               \
    } while (0)
#define LEAN(HLv, PARv)                                                                       \
    do {                                                                                      \
        if (multiday) { if (wide) LEAN2(true, HLv, PARv, true); else LEAN2(true, HLv, PARv, false); } \
        else { if (wide) LEAN2(false, HLv, PARv, true); else LEAN2(false, HLv, PARv, false); }        \
    } while (0)
    // parameter sweeps: combos that differ only in stop/target padding and cooldown run 32 to a warp over one path
    // (lean_sweep_kernel); groups too small to fill most of a warp keep the one-launch-per-combo kernels below
    std::vector<char> swept((size_t)nc, 0);
    if (lean && !par && !multiday && ctx->L > 0 && nc >= kSweepMinLanes && !ctx->no_sweep) {
        auto plain = [&](int c) {
            const DevCombo& D = ctx->h_combos[c];
            const bool on_resident = cl && cl == ctx->d_close;
            return !(D.flags & QMMX_F_USE_BLEND) && (D.flags & QMMX_F_DISABLE_POLICY_GATE) &&
                   !(on_resident && (D.flags & QMMX_F_ENABLE_VETO) && ctx->d_vol) && !(on_resident && ctx->d_vol && ctx->d_regime) &&
                   D.sp >= 0.0 && D.tp >= 0.0 && D.sp + D.tp > 0.0;
        };
        auto same_family = [&](const DevCombo& a, const DevCombo& b) {
            return a.table == b.table && a.qmin == b.qmin && a.prox == b.prox && a.pmax == b.pmax && a.wr == b.wr && a.wm == b.wm &&
                   a.vstrong == b.vstrong && a.vprox == b.vprox && a.overtouch == b.overtouch && a.spd == b.spd && a.flags == b.flags;
        };
        for (int c0 = 0; c0 < nc; ++c0) {
            if (swept[c0] || !plain(c0)) continue;
            std::vector<int> fam;
            for (int c = c0; c < nc; ++c)
                if (!swept[c] && plain(c) && same_family(ctx->h_combos[c0], ctx->h_combos[c])) fam.push_back(c);
            for (size_t at = 0; at < fam.size(); at += 32) {
                const int n = (int)std::min<size_t>(32, fam.size() - at);
                if (n < kSweepMinLanes) break;
                ScanArgs A;
                fill_args(ctx, A, fam[at], cl, h, l, pitch, paths, steps);
                SweepArgs W;
                std::memset(&W, 0, sizeof(W));
                for (int j = 0; j < n; ++j) {
                    const DevCombo& D = ctx->h_combos[fam[at + j]];
                    W.lane[j].sp = D.sp; W.lane[j].tp = D.tp; W.lane[j].cool_steps = D.cool_steps; W.lane[j].combo = fam[at + j];
                    swept[fam[at + j]] = 1;
                }
                W.n_lanes = n;
                W.mlmask = A.mlmask ? ctx->d_mlmask : nullptr;       // per-lane rows; the shared gate table carries no logistic sign
                W.mask_stride = (int64_t)ctx->n_lr * kMaxLevels;
                W.flat = ctx->d_flat;
                A.mlmask = nullptr;
                const size_t sw_smem = sweep_smem_bytes(ctx->L);
                const dim3 sw_grid((unsigned)std::max<int64_t>(1, std::min<int64_t>((paths + kLeanWarps - 1) / kLeanWarps, (int64_t)ctx->sm_count)), 1, 1);
#define SWEEP(HLv, Wv)                                                                        \
    do {                                                                                      \
        auto k = lean_sweep_kernel<HLv, Wv>;                                                  \
        if ((rc = set_smem_attr(ctx, k, sw_smem)) != 0) return rc;                            \
        k<<<sw_grid, kLeanThreads, sw_smem, ctx->stream>>>(A, W);                             \
    } while (0)
                if (hl) { if (wide) SWEEP(true, true); else SWEEP(true, false); }
                else { if (wide) SWEEP(false, true); else SWEEP(false, false); }
#undef SWEEP
                ctx->launches++;
                CK(cudaGetLastError());
            }
        }
    }
    for (int c = 0; c < nc; ++c) {
        if (swept[c]) continue;
        ScanArgs A;
        fill_args(ctx, A, c, cl, h, l, pitch, paths, steps);
        const bool parity = par && c == 0;
        if (parity) { A.trades = par->trades; A.tcounts = par->tcounts; A.codes = par->codes; A.max_trades = par->max_trades; }
        if ((A.combo.flags & QMMX_F_ONLINE_RETRAIN) && !(A.combo.flags & QMMX_F_DISABLE_POLICY_GATE)) {
            // the learnt heads are per path and per combo: one combo, the ctx's resident bars, the lean kernels
            const bool band = A.combo.sp >= 0.0 && A.combo.tp >= 0.0 && A.combo.sp + A.combo.tp > 0.0;
            if (nc != 1 || !cl || cl != ctx->d_close || !lean || !band || ctx->n_pol != 1)
                return fail(ctx, QMMX_E_UNSUPPORTED, "ONLINE_RETRAIN needs one combo, one policy table, resident bars and the lean scan kernel");
            if (!ctx->d_polw || ctx->polw_pitch != pitch) {
                if (ctx->d_polw) cudaFree(ctx->d_polw);
                ctx->d_polw = nullptr;
                CK(cudaMalloc(&ctx->d_polw, (size_t)14 * (size_t)pitch * sizeof(double)));
                ctx->polw_pitch = pitch;
            }
            ctx->polw_paths = paths;
            polw_init_kernel<<<(unsigned)std::min<int64_t>((14 * pitch + 255) / 256, 65535), 256, 0, ctx->stream>>>(ctx->d_polw, ctx->d_pol, pitch);
            ctx->launches++;
            CK(cudaGetLastError());
            A.polw = ctx->d_polw;
        }
        const bool slow = (A.combo.flags & QMMX_F_USE_BLEND) || !(A.combo.flags & QMMX_F_DISABLE_POLICY_GATE) || A.vol != nullptr || A.regime != nullptr;
        // the lean exit block assumes a single price cannot hit both thresholds: stop below / target above the level
        const bool band_ok = A.combo.sp >= 0.0 && A.combo.tp >= 0.0 && A.combo.sp + A.combo.tp > 0.0;
        if (lean && band_ok && !parity && !hl && !wide && !A.polw && !ctx->no_tile && pitch < ((int64_t)1 << 26)) {
            // close-only exits, <= 32 levels: the tile-event kernel (scan_tile.cuh)
            const size_t tsm = tile_smem_bytes();
#define TILE(MDv, SLv)                                                                        \
    do {                                                                                      \
        auto k = tile_scan_kernel<MDv, SLv>;                                                  \
        if ((rc = set_smem_attr(ctx, k, tsm)) != 0) return rc;                                \
        k<<<lean_grid, kLeanThreads, tsm, ctx->stream>>>(A);                                  \
    } while (0)
#define TILE_TMA(MDv, SLv)                                                                    \
    do {                                                                                      \
        auto k = tile_scan_tma_kernel<MDv, SLv>;                                              \
        if ((rc = set_smem_attr(ctx, k, tsm)) != 0) return rc;                                \
        k<<<lean_grid, kLeanThreads, tsm, ctx->stream>>>(A, tmap);                            \
    } while (0)
            CUtensorMap tmap;
            if (make_bar_tmap(ctx, &tmap, cl, pitch, paths, steps)) {
                if (multiday) { if (slow) TILE_TMA(true, true); else TILE_TMA(true, false); }
                else { if (slow) TILE_TMA(false, true); else TILE_TMA(false, false); }
            } else if (multiday) { if (slow) TILE(true, true); else TILE(true, false); }
            else { if (slow) TILE(false, true); else TILE(false, false); }
#undef TILE_TMA
#undef TILE
        } else if (lean && band_ok) {
            if (!parity) { if (hl) LEAN(true, false); else LEAN(false, false); }
            else { if (hl) LEAN(true, true); else LEAN(false, true); }
        } else if (!parity) {
            if (!hl && !tiny) LAUNCH(false, false, false);
            else if (hl && !tiny) LAUNCH(true, false, false);
            else if (!hl && tiny) LAUNCH(false, false, true);
            else LAUNCH(true, false, true);
        } else {
            if (!hl && !tiny) LAUNCH(false, true, false);
            else if (hl && !tiny) LAUNCH(true, true, false);
            else if (!hl && tiny) LAUNCH(false, true, true);
            else LAUNCH(true, true, true);
        }
        ctx->launches++;
        CK(cudaGetLastError());
    }
#undef LAUNCH
#undef LAUNCH2
#undef LEAN
#undef LEAN2
#undef LEAN3
#undef LEAN3L
    return QMMX_OK;
}

void unpack_one(const qmmx_ctx* ctx, const unsigned long long* f, int combo, qmmx_agg* a) {
    std::memset(a, 0, sizeof(*a));
    a->struct_size = sizeof(*a); a->combo = (uint32_t)combo;
    a->n_paths = f[U_NPATHS]; a->n_steps = f[U_NSTEPS]; a->n_trades = f[U_NTRADES]; a->n_wins = f[U_NWINS];
    a->n_open = f[U_NOPEN]; a->n_trade_overflow = f[U_NOVERFLOW];
    for (int c = 0; c < QMMX_N_ENTRY_CODES; ++c) a->entry_hist[c] = f[U_ENTRY + c];
    for (int c = 0; c < QMMX_N_EXIT_CODES; ++c) a->exit_hist[c] = f[U_EXIT + c];
    for (int c = 0; c < QMMX_PNL_HIST_BINS; ++c) a->pnl_hist[c] = f[U_HIST + c];
    double d[QMMX_AGG_ND + QMMX_AGG_NX];
    std::memcpy(d, f + QMMX_AGG_NU, sizeof(d));
    a->sum_pnl = d[D_SUM_PNL]; a->sum_pnl2 = d[D_SUM_PNL2]; a->sum_path_pnl2 = d[D_SUM_PATH2]; a->sum_open_mtm = d[D_SUM_MTM];
    a->max_path_pnl = d[QMMX_AGG_ND]; a->min_path_pnl = -d[QMMX_AGG_ND + 1];
    a->pnl_hist_lo = ctx ? ctx->hist_lo : 0.0; a->pnl_hist_hi = ctx ? ctx->hist_hi : 0.0;
}

GbmDev make_gbm(const qmmx_gbm* g, uint64_t seed, uint64_t path_offset) {
    GbmDev d;
    d.s0 = g->s0;
    d.drift = (float)(((double)g->mu_annual - 0.5 * (double)g->sigma_annual * (double)g->sigma_annual) *
                      (double)g->dt_years / (double)g->sub_steps);
    d.vol = (float)((double)g->sigma_annual * std::sqrt((double)g->dt_years / (double)g->sub_steps));
    d.drift2 = (float)((double)d.drift * 1.4426950408889634074); d.vol2 = (float)((double)d.vol * 1.4426950408889634074);
    d.sub_steps = g->sub_steps; d.cents = (g->flags & QMMX_GBM_ROUND_CENTS) ? 1 : 0;
    d.k0 = (uint32_t)seed; d.k1 = (uint32_t)(seed >> 32); d.path_offset = path_offset;
    for (uint32_t r = 0; r < 10; ++r) { d.rk[2 * r] = d.k0 + r * 0x9E3779B9u; d.rk[2 * r + 1] = d.k1 + r * 0xBB67AE85u; }
    return d;
}

int check_gbm(qmmx_ctx* ctx, const qmmx_gbm* g) {
    if (!g || g->struct_size != sizeof(qmmx_gbm)) return fail(ctx, QMMX_E_BADARG, "qmmx_gbm.struct_size mismatch");
    if (g->sub_steps != 1 && g->sub_steps != 4) return fail(ctx, QMMX_E_BADARG, "qmmx_gbm.sub_steps must be 1 or 4");
    return QMMX_OK;
}

}  // namespace

// ---- scratch of qmmx_run_last_bars_static: a file-scope device array, so that the entry point allocates nothing.
// Every GPU that loads the library holds these kStaticScratchBytes until the process exits.  In HOST code the name
// g_static_scratch is a host-side stand-in, not a device address: kernels get the pointer cudaGetSymbolAddress returns.
constexpr size_t kStaticScratchBytes = (size_t)1 << 20;
constexpr int kStaticMaxBars = QMMX_STATIC_MAX_BARS, kStaticMaxCand = QMMX_STATIC_MAX_CAND;
constexpr size_t static_last_bars_bytes(size_t nb, size_t nk) { return (2 * nb + 2 * nk) * 8 + 3 * nk * 4 + nk; }
static_assert(static_last_bars_bytes(kStaticMaxBars, kStaticMaxCand) <= kStaticScratchBytes, "static scratch too small for the stated maxima");
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static uint8_t* g_static_ptr[kStaticMaxDevices];
static cudaError_t g_static_err[kStaticMaxDevices];
static bool g_static_known[kStaticMaxDevices];

// once per device, with that device current (qmmx_create has just made it so)
static cudaError_t static_scratch_address(int device, uint8_t** out) {
    *out = nullptr;
    if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_known[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) { cudaGetLastError(); p = nullptr; }
        g_static_ptr[device] = static_cast<uint8_t*>(p);
        g_static_err[device] = e;
        g_static_known[device] = true;
    }
    *out = g_static_ptr[device];
    return g_static_err[device];
}

#define ENTER(c)                                                         \
    if (!(c)) return QMMX_E_BADARG;                                      \
    qmmx_ctx* ctx = (c);                                                 \
    CK(cudaSetDevice(ctx->device))

extern "C" {

int qmmx_abi_version(void) { return QMMX_ABI_VERSION; }

const char* qmmx_strerror(int status) {
    switch (status) {
        case QMMX_OK: return "ok";
        case QMMX_E_BADARG: return "bad argument";
        case QMMX_E_CUDA: return "CUDA error";
        case QMMX_E_NOMEM: return "out of device memory";
        case QMMX_E_STATE: return "call order: levels / params / bars not set";
        case QMMX_E_UNSUPPORTED: return "unsupported configuration";
        case QMMX_E_NCCL: return "NCCL unavailable or failed";
        default: return "unknown status";
    }
}

const char* qmmx_last_error(const qmmx_ctx* ctx) { return ctx ? ctx->err.c_str() : "null ctx"; }

int qmmx_create(int device, qmmx_ctx** out) {
    if (!out) return QMMX_E_BADARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return QMMX_E_CUDA;
    qmmx_ctx* ctx = new qmmx_ctx();
    ctx->device = device;
    auto bail = [&](int code) { delete ctx; return code; };
    if (cudaSetDevice(device) != cudaSuccess) return bail(QMMX_E_CUDA);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return bail(QMMX_E_CUDA);
    if (prop.major < 10) return bail(QMMX_E_UNSUPPORTED);   // sm_100a cubin only
    ctx->sm_count = prop.multiProcessorCount;
    { const char* v1 = std::getenv("QMMX_SCAN_V1"); ctx->force_v1 = v1 && v1[0] == '1'; }
    { const char* ns = std::getenv("QMMX_NO_SWEEP"); ctx->no_sweep = ns && ns[0] == '1'; }
    { const char* nt = std::getenv("QMMX_NO_TILE"); ctx->no_tile = nt && nt[0] == '1'; }
    { const char* nt = std::getenv("QMMX_NO_TMA"); ctx->no_tma = nt && nt[0] == '1'; }
    {
        cudaDriverEntryPointQueryResult qres;
        void* fn = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess) ctx->encode_tiled = fn;
        else cudaGetLastError();       // no encoder: the tile kernel stages with cp.async
    }
    if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess) return bail(QMMX_E_CUDA);
    if (cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) return bail(QMMX_E_CUDA);
    ctx->stream = ctx->own_stream;
    cudaEventCreate(&ctx->ev0); cudaEventCreate(&ctx->ev1);
    for (int i = 0; i < 2; ++i) {
        cudaEventCreateWithFlags(&ctx->ev_copied[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming);
    }
    ctx->z_go = sigmoid_threshold(0.60);     // M:3083-3085
    ctx->z_skip = sigmoid_threshold(0.55);
    ctx->static_scratch_err = static_scratch_address(device, &ctx->static_scratch);   // only qmmx_run_last_bars_static needs it
    *out = ctx;
    return QMMX_OK;
}

int qmmx_destroy(qmmx_ctx* ctx) {
    if (!ctx) return QMMX_E_BADARG;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    free_bars(ctx);
    for (int b = 0; b < 2; ++b) for (int f = 0; f < 3; ++f) cudaFree(ctx->d_stage[b][f]);
    cudaFree(ctx->d_tables); cudaFree(ctx->d_mlmask); cudaFree(ctx->d_mlprob);
    if (ctx->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(ctx->nccl_comm);
    cudaFree(ctx->d_pol); cudaFree(ctx->d_flat); cudaFree(ctx->d_ratios); cudaFree(ctx->d_polw); cudaFree(ctx->d_gather); cudaFree(ctx->d_counters);
    cudaEventDestroy(ctx->ev0); cudaEventDestroy(ctx->ev1);
    for (int i = 0; i < 2; ++i) { cudaEventDestroy(ctx->ev_copied[i]); cudaEventDestroy(ctx->ev_done[i]); }
    cudaStreamDestroy(ctx->own_stream); cudaStreamDestroy(ctx->copy_stream);
    delete ctx;
    return QMMX_OK;
}

int qmmx_set_stream(qmmx_ctx* c, void* cuda_stream) {
    ENTER(c);
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return QMMX_OK;
}

int qmmx_set_levels(qmmx_ctx* c, const double* price, const uint8_t* solid, int n_levels) {
    ENTER(c);
    if (n_levels < 0 || n_levels > QMMX_MAX_LEVELS || (n_levels > 0 && (!price || !solid)))
        return fail(ctx, QMMX_E_BADARG, "n_levels must be 0..64 with non-null arrays");
    for (int j = 0; j < n_levels; ++j) {
        if (!std::isfinite(price[j])) return fail(ctx, QMMX_E_BADARG, "level prices must be finite");
        ctx->price[j] = price[j]; ctx->solid[j] = solid[j] ? 1 : 0;
    }
    ctx->L = n_levels; ctx->dirty = true;
    return QMMX_OK;
}

int qmmx_set_params(qmmx_ctx* c, const qmmx_params* params, int n_combos) {
    ENTER(c);
    if (!params || n_combos < 1 || n_combos > QMMX_MAX_COMBOS) return fail(ctx, QMMX_E_BADARG, "n_combos must be 1..1024");
    for (int i = 0; i < n_combos; ++i) {
        const qmmx_params& P = params[i];
        if (P.struct_size != sizeof(qmmx_params)) return fail(ctx, QMMX_E_BADARG, "qmmx_params.struct_size mismatch");
        if (P.overtouch_limit < 1 || P.overtouch_limit > 15) return fail(ctx, QMMX_E_BADARG, "overtouch_limit must be 1..15");
        if (P.steps_per_day < 1) return fail(ctx, QMMX_E_BADARG, "steps_per_day must be >= 1");
        if (!(P.contact_prox == P.contact_prox) || !(P.q_min_prob == P.q_min_prob))
            return fail(ctx, QMMX_E_BADARG, "contact_prox / q_min_prob must not be NaN");
    }
    ctx->params.assign(params, params + n_combos);
    ctx->dirty = true;
    return QMMX_OK;
}

int qmmx_set_lr(qmmx_ctx* c, const double* coef_intercept, int n_days) {
    ENTER(c);
    if (n_days < 0 || n_days > kMaxDays || (n_days > 0 && !coef_intercept)) return fail(ctx, QMMX_E_BADARG, "n_days must be 0..64");
    ctx->lr.assign(coef_intercept, coef_intercept + (size_t)n_days * QMMX_LR_DIM);
    ctx->n_lr = n_days; ctx->dirty = true;
    return QMMX_OK;
}

int qmmx_set_policy(qmmx_ctx* c, const double* w, int n_days) {
    ENTER(c);
    if (n_days < 0 || n_days > kMaxDays || (n_days > 0 && !w)) return fail(ctx, QMMX_E_BADARG, "n_days must be 0..64");
    ctx->pol.assign(w, w + (size_t)n_days * QMMX_POLICY_HEADS * QMMX_POLICY_DIM);
    ctx->n_pol = n_days; ctx->dirty = true;
    return QMMX_OK;
}

int qmmx_set_pnl_hist_range(qmmx_ctx* c, double lo, double hi) {
    ENTER(c);
    if (!(hi > lo)) return fail(ctx, QMMX_E_BADARG, "pnl histogram range must satisfy hi > lo");
    ctx->hist_lo = lo; ctx->hist_hi = hi;
    return QMMX_OK;
}

int qmmx_upload_bars(qmmx_ctx* c, const float* close, const float* high, const float* low, int64_t paths, int32_t steps) {
    ENTER(c);
    if (!close || paths < 0 || steps < 0 || ((high == nullptr) != (low == nullptr)))
        return fail(ctx, QMMX_E_BADARG, "close must be non-null; high and low come together");
    int rc = alloc_bars(ctx, paths, steps, high != nullptr);
    if (rc != 0) return rc;
    free_volume(ctx);                       // new bars: a volume plane of the old ones would be stale
    if (paths == 0 || steps == 0) return QMMX_OK;
    const size_t w = (size_t)paths * 4, dp = (size_t)ctx->pitch * 4;
    CK(cudaMemcpy2DAsync(ctx->d_close, dp, close, w, w, steps, cudaMemcpyHostToDevice, ctx->stream));
    if (high) {
        CK(cudaMemcpy2DAsync(ctx->d_high, dp, high, w, w, steps, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpy2DAsync(ctx->d_low, dp, low, w, w, steps, cudaMemcpyHostToDevice, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return QMMX_OK;
}

int qmmx_bind_device_bars(qmmx_ctx* c, const float* d_close, const float* d_high, const float* d_low,
                          int64_t paths, int32_t steps, int64_t pitch) {
    ENTER(c);
    if (!d_close || paths < 0 || steps < 0 || pitch < paths || ((d_high == nullptr) != (d_low == nullptr)))
        return fail(ctx, QMMX_E_BADARG, "bad device bar description");
    free_bars(ctx);
    ctx->d_close = const_cast<float*>(d_close); ctx->d_high = const_cast<float*>(d_high); ctx->d_low = const_cast<float*>(d_low);
    ctx->own_bars = false; ctx->paths = paths; ctx->steps = steps; ctx->pitch = pitch;
    return QMMX_OK;
}

int qmmx_device_bars(qmmx_ctx* c, float** d_close, float** d_high, float** d_low, int64_t* pitch) {
    ENTER(c);
    if (d_close) *d_close = ctx->d_close;
    if (d_high) *d_high = ctx->d_high;
    if (d_low) *d_low = ctx->d_low;
    if (pitch) *pitch = ctx->pitch;
    return QMMX_OK;
}

int qmmx_pin_host(qmmx_ctx* c, void* ptr, size_t bytes) {
    ENTER(c);
    if (!ptr || bytes == 0) return fail(ctx, QMMX_E_BADARG, "null host buffer");
    CK(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
    return QMMX_OK;
}

int qmmx_unpin_host(qmmx_ctx* c, void* ptr) {
    ENTER(c);
    CK(cudaHostUnregister(ptr));
    return QMMX_OK;
}

int qmmx_download_bars(qmmx_ctx* c, float* close, float* high, float* low) {
    ENTER(c);
    if (!ctx->d_close) return fail(ctx, QMMX_E_STATE, "no resident bars");
    if (ctx->paths == 0 || ctx->steps == 0) return QMMX_OK;
    const size_t w = (size_t)ctx->paths * 4, dp = (size_t)ctx->pitch * 4;
    if (close) CK(cudaMemcpy2DAsync(close, w, ctx->d_close, dp, w, ctx->steps, cudaMemcpyDeviceToHost, ctx->stream));
    if (high) {
        if (!ctx->d_high) return fail(ctx, QMMX_E_STATE, "resident bars have no high/low");
        CK(cudaMemcpy2DAsync(high, w, ctx->d_high, dp, w, ctx->steps, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (low) {
        if (!ctx->d_low) return fail(ctx, QMMX_E_STATE, "resident bars have no high/low");
        CK(cudaMemcpy2DAsync(low, w, ctx->d_low, dp, w, ctx->steps, cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return QMMX_OK;
}

int qmmx_upload_volume(qmmx_ctx* c, const float* volume, int64_t paths, int32_t steps) {
    ENTER(c);
    if (!volume) { free_volume(ctx); return QMMX_OK; }
    if (!ctx->d_close) return fail(ctx, QMMX_E_STATE, "no resident bars (qmmx_upload_bars / qmmx_gen_* first)");
    if (!ctx->own_bars) return fail(ctx, QMMX_E_UNSUPPORTED, "volume needs library-owned bars (pitch is the library's)");
    if (paths != ctx->paths || steps != ctx->steps) return fail(ctx, QMMX_E_BADARG, "volume shape differs from the resident bars");
    if (paths == 0 || steps == 0) return QMMX_OK;
    if (!ctx->d_vol) CK(cudaMalloc(&ctx->d_vol, (size_t)ctx->pitch * (size_t)steps * sizeof(float)));
    const size_t w = (size_t)paths * 4, dp = (size_t)ctx->pitch * 4;
    CK(cudaMemcpy2DAsync(ctx->d_vol, dp, volume, w, w, steps, cudaMemcpyHostToDevice, ctx->stream));
    int rc = compute_regime(ctx);
    if (rc != 0) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return QMMX_OK;
}

int qmmx_download_volume(qmmx_ctx* c, float* volume) {
    ENTER(c);
    if (!ctx->d_vol) return fail(ctx, QMMX_E_STATE, "no resident volume");
    if (!volume) return fail(ctx, QMMX_E_BADARG, "null destination");
    if (ctx->paths == 0 || ctx->steps == 0) return QMMX_OK;
    const size_t w = (size_t)ctx->paths * 4, dp = (size_t)ctx->pitch * 4;
    CK(cudaMemcpy2DAsync(volume, w, ctx->d_vol, dp, w, ctx->steps, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return QMMX_OK;
}

int qmmx_set_bootstrap_volume(qmmx_ctx* c, const float* window_volume, int32_t window_len) {
    ENTER(c);
    if (!window_volume) { ctx->boot_vol.clear(); return QMMX_OK; }
    if (window_len < 2 || window_len > kMaxWindow + 1) return fail(ctx, QMMX_E_BADARG, "window_len must be 2..1025");
    ctx->boot_vol.assign(window_volume, window_volume + window_len);
    return QMMX_OK;
}

int qmmx_gen_gbm(qmmx_ctx* c, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps, const qmmx_gbm* gbm) {
    ENTER(c);
    int rc = check_gbm(ctx, gbm);
    if (rc != 0) return rc;
    if (paths < 0 || steps < 0) return fail(ctx, QMMX_E_BADARG, "negative size");
    if ((rc = alloc_bars(ctx, paths, steps, gbm->sub_steps == 4)) != 0) return rc;
    free_volume(ctx);
    if (paths == 0 || steps == 0) return QMMX_OK;
    const GbmDev g = make_gbm(gbm, seed, path_offset);
    const int64_t blocks = std::min<int64_t>((paths + kThreads - 1) / kThreads, (int64_t)ctx->sm_count * 8);
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    gen_gbm_kernel<<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(ctx->d_close, ctx->d_high, ctx->d_low, ctx->pitch, paths, steps, g);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_gen_bootstrap(qmmx_ctx* c, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                       const float* window_ohlc, int32_t window_len, int32_t block_len) {
    ENTER(c);
    const int nret = window_len - 1;
    if (!window_ohlc || nret < 1 || nret > kMaxWindow || block_len < 1 || paths < 0 || steps < 0)
        return fail(ctx, QMMX_E_BADARG, "window_len must be 2..1025, block_len >= 1");
    int rc = alloc_bars(ctx, paths, steps, true);
    if (rc != 0) return rc;
    if (paths == 0 || steps == 0) return QMMX_OK;
    const bool with_vol = (int32_t)ctx->boot_vol.size() == window_len;     // qmmx_set_bootstrap_volume
    std::vector<float> ratios((with_vol ? 4 : 3) * (size_t)nret);
    if (with_vol) for (int t = 1; t < window_len; ++t) ratios[3 * (size_t)nret + t - 1] = ctx->boot_vol[t];
    if (with_vol) { if (!ctx->d_vol) CK(cudaMalloc(&ctx->d_vol, (size_t)ctx->pitch * (size_t)steps * sizeof(float))); }
    else free_volume(ctx);
    for (int t = 1; t < window_len; ++t) {
        const float pc = window_ohlc[4 * (t - 1) + 3];
        ratios[t - 1] = window_ohlc[4 * t + 1] / pc;
        ratios[nret + t - 1] = window_ohlc[4 * t + 2] / pc;
        ratios[2 * nret + t - 1] = window_ohlc[4 * t + 3] / pc;
    }
    if (ctx->ratios_cap < ratios.size()) {
        if (ctx->d_ratios) cudaFree(ctx->d_ratios);
        ctx->d_ratios = nullptr; ctx->ratios_cap = 0;
        CK(cudaMalloc(&ctx->d_ratios, ratios.size() * 4));
        ctx->ratios_cap = ratios.size();
    }
    float* d_r = ctx->d_ratios;
    // pageable source: the runtime stages the copy before returning, so `ratios` may go out of scope
    cudaError_t e = cudaMemcpyAsync(d_r, ratios.data(), ratios.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        BootDev g;
        g.ratios = d_r; g.nret = nret; g.block_len = block_len; g.s0 = window_ohlc[4 * (window_len - 1) + 3];
        g.k0 = (uint32_t)seed; g.k1 = (uint32_t)(seed >> 32); g.path_offset = path_offset;
        const int64_t blocks = std::min<int64_t>((paths + kThreads - 1) / kThreads, (int64_t)ctx->sm_count * 8);
        cudaEventRecord(ctx->ev0, ctx->stream);
        gen_bootstrap_kernel<<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(ctx->d_close, ctx->d_high, ctx->d_low, with_vol ? ctx->d_vol : nullptr, ctx->pitch, paths, steps, g);
        ctx->launches++;
        e = cudaGetLastError();
        cudaEventRecord(ctx->ev1, ctx->stream);
    }
    if (e != cudaSuccess) return fail(ctx, QMMX_E_CUDA, std::string("gen_bootstrap: ") + cudaGetErrorString(e));
    if (with_vol && (rc = compute_regime(ctx)) != 0) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_run_scan_async(qmmx_ctx* c) {
    ENTER(c);
    if (!ctx->d_close) return fail(ctx, QMMX_E_STATE, "no resident bars (qmmx_upload_bars / qmmx_gen_* first)");
    int rc = prepare(ctx);
    if (rc != 0) return rc;
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    if ((rc = init_flat(ctx)) != 0) return rc;
    if ((rc = launch_scan(ctx, ctx->d_close, ctx->d_high, ctx->d_low, ctx->pitch, ctx->paths, ctx->steps, nullptr)) != 0) return rc;
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_fetch_agg(qmmx_ctx* c, qmmx_agg* agg_out) {
    ENTER(c);
    { int rc0 = check_agg(ctx); if (rc0 != 0) return rc0; }
    const int nc = (int)ctx->params.size();
    std::vector<unsigned long long> flat((size_t)nc * QMMX_AGG_SLOTS);
    CK(cudaMemcpyAsync(flat.data(), ctx->d_flat, flat.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (agg_out) for (int i = 0; i < nc; ++i) unpack_one(ctx, flat.data() + (size_t)i * QMMX_AGG_SLOTS, i, agg_out + i);
    return QMMX_OK;
}

int qmmx_run_scan(qmmx_ctx* c, qmmx_agg* agg_out, qmmx_trade* trades_out, int32_t* trade_counts_out,
                  int32_t max_trades_per_path, uint8_t* codes_out) {
    ENTER(c);
    const bool parity = trades_out || trade_counts_out || codes_out;
    if (!parity) {
        int rc = qmmx_run_scan_async(ctx);
        if (rc != 0) return rc;
        return qmmx_fetch_agg(ctx, agg_out);
    }
    if (!ctx->d_close) return fail(ctx, QMMX_E_STATE, "no resident bars (qmmx_upload_bars / qmmx_gen_* first)");
    if (trades_out && max_trades_per_path < 1) return fail(ctx, QMMX_E_BADARG, "max_trades_per_path must be >= 1");
    int rc = prepare(ctx);
    if (rc != 0) return rc;
    const int64_t P = ctx->paths; const int32_t T = ctx->steps;
    qmmx_trade* d_tr = nullptr; int32_t* d_cnt = nullptr; uint8_t* d_codes = nullptr;
    cudaError_t e = cudaSuccess;
    if (trades_out && P > 0) e = cudaMalloc(&d_tr, (size_t)P * max_trades_per_path * sizeof(qmmx_trade));
    if (e == cudaSuccess && trades_out && P > 0) e = cudaMemsetAsync(d_tr, 0, (size_t)P * max_trades_per_path * sizeof(qmmx_trade), ctx->stream);
    if (e == cudaSuccess && (trade_counts_out || trades_out) && P > 0) e = cudaMalloc(&d_cnt, (size_t)P * 4);
    if (e == cudaSuccess && codes_out && P > 0 && T > 0) e = cudaMalloc(&d_codes, (size_t)P * T);
    if (e == cudaSuccess) {
        ParityOut par;
        par.trades = d_tr; par.tcounts = d_cnt; par.max_trades = trades_out ? max_trades_per_path : 0; par.codes = d_codes;
        cudaEventRecord(ctx->ev0, ctx->stream);
        rc = init_flat(ctx);
        if (rc == 0) rc = launch_scan(ctx, ctx->d_close, ctx->d_high, ctx->d_low, ctx->pitch, P, T, &par);
        cudaEventRecord(ctx->ev1, ctx->stream);
        if (rc == 0) {
            if (trades_out && d_tr) e = cudaMemcpyAsync(trades_out, d_tr, (size_t)P * max_trades_per_path * sizeof(qmmx_trade), cudaMemcpyDeviceToHost, ctx->stream);
            if (e == cudaSuccess && trade_counts_out && d_cnt) e = cudaMemcpyAsync(trade_counts_out, d_cnt, (size_t)P * 4, cudaMemcpyDeviceToHost, ctx->stream);
            if (e == cudaSuccess && codes_out && d_codes) e = cudaMemcpyAsync(codes_out, d_codes, (size_t)P * T, cudaMemcpyDeviceToHost, ctx->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        }
    }
    cudaFree(d_tr); cudaFree(d_cnt); cudaFree(d_codes);
    if (rc != 0) return rc;
    if (e != cudaSuccess) return fail(ctx, e == cudaErrorMemoryAllocation ? QMMX_E_NOMEM : QMMX_E_CUDA, std::string("run_scan: ") + cudaGetErrorString(e));
    ctx->ran = true;
    return qmmx_fetch_agg(ctx, agg_out);
}

// QMMX_F_ONLINE_RETRAIN in the fused (register-only) kernels: the bars never exist, the per-path learnt heads do --
// [14][pitch] float64 in HBM, read and written once per closed trade (112 B), as in the resident-bar scan
static int setup_fused_retrain(qmmx_ctx* ctx, ScanArgs& A, int64_t paths, bool wide) {
    const bool band = A.combo.sp >= 0.0 && A.combo.tp >= 0.0 && A.combo.sp + A.combo.tp > 0.0;
    if (ctx->params.size() != 1 || !ctx->lean_ok || ctx->force_v1 || !band || ctx->n_pol != 1 || wide)
        return fail(ctx, QMMX_E_UNSUPPORTED, "ONLINE_RETRAIN (fused) needs one combo, one policy table, <= 32 levels and the lean engine");
    const int64_t pitch = round_up(std::max<int64_t>(paths, 1), 32);
    if (!ctx->d_polw || ctx->polw_pitch != pitch) {
        if (ctx->d_polw) cudaFree(ctx->d_polw);
        ctx->d_polw = nullptr;
        CK(cudaMalloc(&ctx->d_polw, (size_t)14 * (size_t)pitch * sizeof(double)));
        ctx->polw_pitch = pitch;
    }
    ctx->polw_paths = paths;
    polw_init_kernel<<<(unsigned)std::min<int64_t>((14 * pitch + 255) / 256, 65535), 256, 0, ctx->stream>>>(ctx->d_polw, ctx->d_pol, pitch);
    ctx->launches++;
    CK(cudaGetLastError());
    A.polw = ctx->d_polw; A.pitch = pitch;
    return QMMX_OK;
}

static bool any_retrain(const qmmx_ctx* ctx) {
    for (const DevCombo& D : ctx->h_combos)
        if ((D.flags & QMMX_F_ONLINE_RETRAIN) && !(D.flags & QMMX_F_DISABLE_POLICY_GATE)) return true;
    return false;
}

int qmmx_run_scan_host(qmmx_ctx* c, const float* close, const float* high, const float* low, int64_t paths,
                       int32_t steps, int64_t chunk_paths, qmmx_agg* agg_out) {
    ENTER(c);
    if (!close || paths < 0 || steps < 0 || ((high == nullptr) != (low == nullptr)))
        return fail(ctx, QMMX_E_BADARG, "close must be non-null; high and low come together");
    int rc = prepare(ctx);
    if (rc != 0) return rc;
    if (ctx->any_hl && !high) return fail(ctx, QMMX_E_STATE, "EXIT_ON_HL needs high/low bars");
    if (chunk_paths <= 0) chunk_paths = 1 << 17;
    chunk_paths = round_up(std::min<int64_t>(chunk_paths, std::max<int64_t>(paths, 1)), 32);
    const int nf = high ? 3 : 1;
    const int64_t need = chunk_paths * std::max(steps, 1);
    if (ctx->stage_elems < need || (nf == 3 && !ctx->d_stage[0][1])) {
        for (int b = 0; b < 2; ++b) for (int f = 0; f < 3; ++f) { cudaFree(ctx->d_stage[b][f]); ctx->d_stage[b][f] = nullptr; }
        ctx->stage_elems = 0;
        for (int b = 0; b < 2; ++b) for (int f = 0; f < nf; ++f) CK(cudaMalloc(&ctx->d_stage[b][f], (size_t)need * 4));
        ctx->stage_elems = need;
    }
    const float* src[3] = {close, high, low};
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    if ((rc = init_flat(ctx)) != 0) return rc;
    int n = 0;
    for (int64_t off = 0; off < paths; off += chunk_paths, ++n) {
        const int b = n & 1;
        const int64_t w = std::min(chunk_paths, paths - off);
        if (n >= 2) CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_done[b], 0));
        for (int f = 0; f < nf; ++f)
            CK(cudaMemcpy2DAsync(ctx->d_stage[b][f], (size_t)chunk_paths * 4, src[f] + off, (size_t)paths * 4,
                                 (size_t)w * 4, steps, cudaMemcpyHostToDevice, ctx->copy_stream));
        CK(cudaEventRecord(ctx->ev_copied[b], ctx->copy_stream));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_copied[b], 0));
        if ((rc = launch_scan(ctx, ctx->d_stage[b][0], ctx->d_stage[b][1], ctx->d_stage[b][2], chunk_paths, w, steps, nullptr)) != 0) return rc;
        CK(cudaEventRecord(ctx->ev_done[b], ctx->stream));
    }
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ran = true;
    return qmmx_fetch_agg(ctx, agg_out);
}

int qmmx_run_fused_gbm_async(qmmx_ctx* c, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps, const qmmx_gbm* gbm) {
    ENTER(c);
    int rc = check_gbm(ctx, gbm);
    if (rc != 0) return rc;
    if (paths < 0 || steps < 0) return fail(ctx, QMMX_E_BADARG, "negative size");
    if ((rc = prepare(ctx)) != 0) return rc;
    if (ctx->any_tiny) return fail(ctx, QMMX_E_UNSUPPORTED, "TINY_PRICES is not available in the fused GBM kernel");
    const bool retrain = any_retrain(ctx);
    if (ctx->any_hl && gbm->sub_steps != 4) return fail(ctx, QMMX_E_STATE, "EXIT_ON_HL needs sub_steps = 4 (OHLC bars)");
    const GbmDev g = make_gbm(gbm, seed, path_offset);
    const size_t smem = scan_smem_bytes(ctx);
    const dim3 grid = scan_grid(ctx, paths);
    if (steps > 65535) return fail(ctx, QMMX_E_BADARG, "steps per path must be <= 65535 (16-bit per-path tallies)");
    if (gbm->sub_steps == 4 && !ctx->any_hl)
        return fail(ctx, QMMX_E_UNSUPPORTED, "fused GBM with sub_steps = 4 requires EXIT_ON_HL (use sub_steps = 1 for close-only)");
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    if ((rc = init_flat(ctx)) != 0) return rc;
    const bool wide = ctx->L > 32, hl = gbm->sub_steps == 4;
    const size_t lean_smem = lean_smem_bytes(ctx->L);
    const dim3 lean_grid((unsigned)std::max<int64_t>(1, std::min<int64_t>(((paths + 31) / 32 + kLeanWarps - 1) / kLeanWarps,
                                                                            (int64_t)ctx->sm_count * kLeanBlocksPerSM)), 1, 1);
    for (int c = 0; c < (int)ctx->params.size() && paths > 0 && steps > 0; ++c) {
        ScanArgs A;
        fill_args(ctx, A, c, nullptr, nullptr, nullptr, 0, paths, steps);
#define FUSED(HLv, Wv)                                                                        \
        do {                                                                                  \
            if ((rc = set_smem_attr(ctx, fused_gbm_kernel<HLv, Wv>, smem)) != 0) return rc;   \
            fused_gbm_kernel<HLv, Wv><<<grid, kThreads, smem, ctx->stream>>>(A, g);           \
        } while (0)
#define LFUSED(MDv, HLv, Wv, SLv)                                                             \
        do {                                                                                  \
            if ((rc = set_smem_attr(ctx, lean_fused_gbm_kernel<MDv, HLv, Wv, SLv>, lean_smem)) != 0) return rc;   \
            lean_fused_gbm_kernel<MDv, HLv, Wv, SLv><<<lean_grid, kLeanThreads, lean_smem, ctx->stream>>>(A, g);  \
        } while (0)
#define LFUSED2(HLv, Wv)                                                                      \
        do {                                                                                  \
            if (multiday) { if (slow) LFUSED(true, HLv, Wv, true); else LFUSED(true, HLv, Wv, false); }   \
            else { if (slow) LFUSED(false, HLv, Wv, true); else LFUSED(false, HLv, Wv, false); }          \
        } while (0)
        const bool multiday = ctx->n_lr > 1;
        const bool slow = (A.combo.flags & QMMX_F_USE_BLEND) || !(A.combo.flags & QMMX_F_DISABLE_POLICY_GATE);
        const bool band_ok = A.combo.sp >= 0.0 && A.combo.tp >= 0.0 && A.combo.sp + A.combo.tp > 0.0;
        if (retrain) {              // per-path OnlinePolicy heads learn from the path's own trades
            if ((rc = setup_fused_retrain(ctx, A, paths, wide)) != 0) return rc;
#define LFUSED_L(MDv, HLv)                                                                    \
        do {                                                                                  \
            if ((rc = set_smem_attr(ctx, lean_fused_gbm_kernel<MDv, HLv, false, true, true>, lean_smem)) != 0) return rc;   \
            lean_fused_gbm_kernel<MDv, HLv, false, true, true><<<lean_grid, kLeanThreads, lean_smem, ctx->stream>>>(A, g);  \
        } while (0)
            if (multiday) { if (hl) LFUSED_L(true, true); else LFUSED_L(true, false); }
            else { if (hl) LFUSED_L(false, true); else LFUSED_L(false, false); }
#undef LFUSED_L
        } else if (ctx->lean_ok && !ctx->force_v1 && band_ok && !hl && !wide && !ctx->no_tile) {
            const size_t tsm = tile_smem_bytes();
#define TFUSED(MDv, SLv)                                                                      \
        do {                                                                                  \
            if ((rc = set_smem_attr(ctx, tile_fused_gbm_kernel<MDv, SLv>, tsm)) != 0) return rc;   \
            tile_fused_gbm_kernel<MDv, SLv><<<lean_grid, kLeanThreads, tsm, ctx->stream>>>(A, g);  \
        } while (0)
            if (multiday) { if (slow) TFUSED(true, true); else TFUSED(true, false); }
            else { if (slow) TFUSED(false, true); else TFUSED(false, false); }
#undef TFUSED
        } else if (ctx->lean_ok && !ctx->force_v1 && band_ok) {
            if (hl) { if (wide) LFUSED2(true, true); else LFUSED2(true, false); }
            else { if (wide) LFUSED2(false, true); else LFUSED2(false, false); }
        } else if (hl) { if (wide) FUSED(true, true); else FUSED(true, false); }
        else { if (wide) FUSED(false, true); else FUSED(false, false); }
#undef LFUSED2
#undef LFUSED
#undef FUSED
        ctx->launches++;
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_run_fused_bootstrap_async(qmmx_ctx* c, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                                   const float* window_ohlc, int32_t window_len, int32_t block_len) {
    ENTER(c);
    const int nret = window_len - 1;
    if (!window_ohlc || nret < 1 || nret > kMaxWindow || block_len < 1 || paths < 0 || steps < 0)
        return fail(ctx, QMMX_E_BADARG, "window_len must be 2..1025, block_len >= 1");
    int rc = prepare(ctx);
    if (rc != 0) return rc;
    if (ctx->any_tiny || !ctx->lean_ok) return fail(ctx, QMMX_E_UNSUPPORTED, "fused bootstrap needs the lean engine (no TINY_PRICES, positive level prices)");
    const bool retrain = any_retrain(ctx);
    if (steps > 65535) return fail(ctx, QMMX_E_BADARG, "steps per path must be <= 65535 (16-bit per-path tallies)");
    std::vector<float> ratios(3 * (size_t)nret);
    for (int t = 1; t < window_len; ++t) {
        const float pc = window_ohlc[4 * (t - 1) + 3];
        ratios[t - 1] = window_ohlc[4 * t + 1] / pc;
        ratios[nret + t - 1] = window_ohlc[4 * t + 2] / pc;
        ratios[2 * nret + t - 1] = window_ohlc[4 * t + 3] / pc;
    }
    if (ctx->ratios_cap < ratios.size()) {
        if (ctx->d_ratios) cudaFree(ctx->d_ratios);
        ctx->d_ratios = nullptr; ctx->ratios_cap = 0;
        CK(cudaMalloc(&ctx->d_ratios, ratios.size() * 4));
        ctx->ratios_cap = ratios.size();
    }
    CK(cudaMemcpyAsync(ctx->d_ratios, ratios.data(), ratios.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    BootDev g;
    g.ratios = ctx->d_ratios; g.nret = nret; g.block_len = block_len; g.s0 = window_ohlc[4 * (window_len - 1) + 3];
    g.k0 = (uint32_t)seed; g.k1 = (uint32_t)(seed >> 32); g.path_offset = path_offset;
    const bool wide = ctx->L > 32, hl = ctx->any_hl, multiday = ctx->n_lr > 1;
    const size_t smem = wide ? (size_t)LeanLayout<64>::kEndBoot : (size_t)LeanLayout<32>::kEndBoot;
    const dim3 grid((unsigned)std::max<int64_t>(1, std::min<int64_t>(((paths + 31) / 32 + kLeanWarps - 1) / kLeanWarps,
                                                                       (int64_t)ctx->sm_count * kLeanBlocksPerSM)), 1, 1);
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    if ((rc = init_flat(ctx)) != 0) return rc;
    for (int cc = 0; cc < (int)ctx->params.size() && paths > 0 && steps > 0; ++cc) {
        ScanArgs A;
        fill_args(ctx, A, cc, nullptr, nullptr, nullptr, 0, paths, steps);
        if (!(A.combo.sp >= 0.0 && A.combo.tp >= 0.0 && A.combo.sp + A.combo.tp > 0.0))
            return fail(ctx, QMMX_E_UNSUPPORTED, "fused bootstrap needs stop_padding, tp_padding >= 0 and not both 0");
        const bool slow = (A.combo.flags & QMMX_F_USE_BLEND) || !(A.combo.flags & QMMX_F_DISABLE_POLICY_GATE);
        if (retrain) {
            if ((rc = setup_fused_retrain(ctx, A, paths, wide)) != 0) return rc;
#define FB_L(MDv, HLv)                                                                        \
        do {                                                                                  \
            if ((rc = set_smem_attr(ctx, lean_fused_boot_kernel<MDv, HLv, false, true, true>, smem)) != 0) return rc;   \
            lean_fused_boot_kernel<MDv, HLv, false, true, true><<<grid, kLeanThreads, smem, ctx->stream>>>(A, g);       \
        } while (0)
            if (multiday) { if (hl) FB_L(true, true); else FB_L(true, false); }
            else { if (hl) FB_L(false, true); else FB_L(false, false); }
#undef FB_L
            ctx->launches++;
            CK(cudaGetLastError());
            continue;
        }
#define FB(MDv, HLv, Wv, SLv)                                                                 \
        do {                                                                                  \
            if ((rc = set_smem_attr(ctx, lean_fused_boot_kernel<MDv, HLv, Wv, SLv>, smem)) != 0) return rc;   \
            lean_fused_boot_kernel<MDv, HLv, Wv, SLv><<<grid, kLeanThreads, smem, ctx->stream>>>(A, g);       \
        } while (0)
#define FB2(MDv, HLv, Wv) do { if (slow) FB(MDv, HLv, Wv, true); else FB(MDv, HLv, Wv, false); } while (0)
#define FB3(MDv, HLv) do { if (wide) FB2(MDv, HLv, true); else FB2(MDv, HLv, false); } while (0)
        if (multiday) { if (hl) FB3(true, true); else FB3(true, false); }
        else { if (hl) FB3(false, true); else FB3(false, false); }
#undef FB3
#undef FB2
#undef FB
        ctx->launches++;
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_run_fused_bootstrap(qmmx_ctx* c, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                             const float* window_ohlc, int32_t window_len, int32_t block_len, qmmx_agg* agg_out) {
    int rc = qmmx_run_fused_bootstrap_async(c, seed, path_offset, paths, steps, window_ohlc, window_len, block_len);
    if (rc != 0) return rc;
    return qmmx_fetch_agg(c, agg_out);
}

int qmmx_run_fused_gbm(qmmx_ctx* c, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                       const qmmx_gbm* gbm, qmmx_agg* agg_out) {
    int rc = qmmx_run_fused_gbm_async(c, seed, path_offset, paths, steps, gbm);
    if (rc != 0) return rc;
    return qmmx_fetch_agg(c, agg_out);
}

int qmmx_agg_device(qmmx_ctx* c, void** d_flat, size_t* bytes) {
    ENTER(c);
    { int rc0 = check_agg(ctx); if (rc0 != 0) return rc0; }
    if (d_flat) *d_flat = ctx->d_flat;
    if (bytes) *bytes = ctx->params.size() * QMMX_AGG_SLOTS * 8;
    return QMMX_OK;
}

int qmmx_merge_flat(qmmx_ctx* c, const void* d_parts, int n_parts) {
    ENTER(c);
    { int rc0 = check_agg(ctx); if (rc0 != 0) return rc0; }
    if (!d_parts || n_parts < 1) return fail(ctx, QMMX_E_BADARG, "merge_flat: null parts");
    const int n = (int)ctx->params.size() * QMMX_AGG_SLOTS;
    merge_flat_kernel<<<std::max(1, (n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_flat, static_cast<const unsigned long long*>(d_parts), n_parts, n);
    ctx->launches++;
    CK(cudaGetLastError());
    return QMMX_OK;
}

int qmmx_nccl_unique_id(qmmx_ctx* c, void* id128_out) {
    ENTER(c);
    if (!id128_out) return fail(ctx, QMMX_E_BADARG, "id128_out is null");
    int rc = load_nccl(ctx);
    if (rc != 0) return rc;
    NcclId id;
    const int nr = g_nccl.GetUniqueId(&id);
    if (nr != 0) return nccl_fail(ctx, "ncclGetUniqueId", nr);
    std::memcpy(id128_out, &id, sizeof(id));
    return QMMX_OK;
}

int qmmx_nccl_init(qmmx_ctx* c, const void* id128, int world_size, int rank) {
    ENTER(c);
    if (!id128 || world_size < 1 || rank < 0 || rank >= world_size) return fail(ctx, QMMX_E_BADARG, "nccl_init: bad id / world_size / rank");
    int rc = load_nccl(ctx);
    if (rc != 0) return rc;
    if (ctx->nccl_comm) { g_nccl.CommDestroy(ctx->nccl_comm); ctx->nccl_comm = nullptr; }
    NcclId id;
    std::memcpy(&id, id128, sizeof(id));
    const int nr = g_nccl.CommInitRank(&ctx->nccl_comm, world_size, id, rank);
    if (nr != 0) { ctx->nccl_comm = nullptr; return nccl_fail(ctx, "ncclCommInitRank", nr); }
    return QMMX_OK;
}

int qmmx_allreduce(qmmx_ctx* c, void* nccl_comm) {
    ENTER(c);
    { int rc0 = check_agg(ctx); if (rc0 != 0) return rc0; }
    int rc = load_nccl(ctx);
    if (rc != 0) return rc;
    void* comm = nccl_comm ? nccl_comm : ctx->nccl_comm;
    if (!comm) return fail(ctx, QMMX_E_STATE, "no communicator: pass an ncclComm_t or call qmmx_nccl_init first");
    int world = 0;
    int nr = g_nccl.CommCount(comm, &world);
    if (nr != 0 || world < 1) return nccl_fail(ctx, "ncclCommCount", nr);
    const size_t n = (size_t)ctx->agg_combos * QMMX_AGG_SLOTS;
    if (ctx->gather_slots < n * (size_t)world) {
        if (ctx->d_gather) cudaFree(ctx->d_gather);
        ctx->d_gather = nullptr; ctx->gather_slots = 0;
        CK(cudaMalloc(&ctx->d_gather, n * (size_t)world * 8));
        ctx->gather_slots = n * (size_t)world;
    }
    // same stream as the scan kernels and the merge: the gather reads finished aggregates, the merge reads a landed gather
    nr = g_nccl.AllGather(ctx->d_flat, ctx->d_gather, n, /*ncclUint64*/ 5, comm, ctx->stream);
    if (nr != 0) return nccl_fail(ctx, "ncclAllGather", nr);
    return qmmx_merge_flat(ctx, ctx->d_gather, world);
}

int qmmx_get_stream(qmmx_ctx* c, void** cuda_stream_out) {
    ENTER(c);
    if (!cuda_stream_out) return fail(ctx, QMMX_E_BADARG, "cuda_stream_out is null");
    *cuda_stream_out = (void*)ctx->stream;
    return QMMX_OK;
}

int qmmx_agg_unpack(qmmx_ctx* ctx, const void* flat_host, int n_combos, qmmx_agg* out) {
    if (!flat_host || !out || n_combos < 0) return QMMX_E_BADARG;
    const unsigned long long* f = static_cast<const unsigned long long*>(flat_host);
    for (int i = 0; i < n_combos; ++i) unpack_one(ctx, f + (size_t)i * QMMX_AGG_SLOTS, i, out + i);
    return QMMX_OK;
}

int qmmx_lr_score(qmmx_ctx* c, const float* x, int64_t n, int32_t dim, const float* w, const float* bias,
                  int32_t heads, float* out) {
    ENTER(c);
    if (!x || !w || !out || n < 0 || dim < 1 || dim > kLrMaxDim || heads < 1 || heads > kLrMaxHeads)
        return fail(ctx, QMMX_E_BADARG, "lr_score: dim 1..32, heads 1..8");
    if (n == 0) return QMMX_OK;
    float *dx = nullptr, *dw = nullptr, *db = nullptr, *dout = nullptr;
    cudaError_t e = cudaMalloc(&dx, (size_t)n * dim * 4);
    if (e == cudaSuccess) e = cudaMalloc(&dw, (size_t)heads * dim * 4);
    if (e == cudaSuccess && bias) e = cudaMalloc(&db, (size_t)heads * 4);
    if (e == cudaSuccess) e = cudaMalloc(&dout, (size_t)n * heads * 4);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dx, x, (size_t)n * dim * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dw, w, (size_t)heads * dim * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && bias) e = cudaMemcpyAsync(db, bias, (size_t)heads * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        const int64_t blocks = std::min<int64_t>((n + kThreads - 1) / kThreads, (int64_t)ctx->sm_count * 8);
        cudaEventRecord(ctx->ev0, ctx->stream);
        lr_score_kernel<<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(dx, n, dim, dw, db, heads, dout);
        ctx->launches++;
        e = cudaGetLastError();
        cudaEventRecord(ctx->ev1, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout, (size_t)n * heads * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(dx); cudaFree(dw); cudaFree(db); cudaFree(dout);
    if (e != cudaSuccess) return fail(ctx, e == cudaErrorMemoryAllocation ? QMMX_E_NOMEM : QMMX_E_CUDA, std::string("lr_score: ") + cudaGetErrorString(e));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_run_jitter(qmmx_ctx* c, const double* high, const double* low, int32_t n_bars, const int32_t* cand_bar,
                    const uint8_t* cand_side, const double* cand_level, const double* cand_entry, int32_t n_cand,
                    double stop_padding, double tp_padding, const double* sigmas, int64_t trials, uint64_t seed,
                    uint64_t trial_offset, const double* replay_normals, const double* replay_uniforms,
                    double* out_total_r, double* out_max_dd, int32_t* out_wins, int32_t* out_losses, int32_t* out_opens) {
    ENTER(c);
    if (!high || !low || n_bars < 1 || n_bars > 12000 || n_cand < 0 || trials < 0 || !sigmas ||
        (n_cand > 0 && (!cand_bar || !cand_side || !cand_level || !cand_entry)) ||
        ((replay_normals == nullptr) != (replay_uniforms == nullptr)) ||
        !out_total_r || !out_max_dd || !out_wins || !out_losses || !out_opens)
        return fail(ctx, QMMX_E_BADARG, "run_jitter: bad arguments (n_bars 1..12000)");
    if (trials == 0) return QMMX_OK;
    const size_t nb = (size_t)n_bars, nk = (size_t)n_cand, nt = (size_t)trials;
    // one device arena: high, low, level, entry | replay normals, uniforms | outputs | int arrays
    const size_t n_f64_in = 2 * nb + 2 * nk + (replay_normals ? nt * nk * 5 : 0);
    const size_t n_f64 = n_f64_in + 2 * nt;
    const size_t n_i32 = nk + 3 * nt;
    uint8_t* arena = nullptr;
    const size_t bytes = n_f64 * 8 + n_i32 * 4 + nk + 64;
    CK(cudaMalloc(&arena, bytes));
    double* d64 = reinterpret_cast<double*>(arena);
    int32_t* i32 = reinterpret_cast<int32_t*>(arena + n_f64 * 8);
    uint8_t* u8 = arena + n_f64 * 8 + n_i32 * 4;
    JitterArgs A;
    std::memset(&A, 0, sizeof(A));
    double* p = d64;
    cudaError_t e = cudaSuccess;
    auto up = [&](const void* src, void* dst, size_t n) {
        if (e == cudaSuccess && n) e = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, ctx->stream);
    };
    A.high = p; up(high, p, nb * 8); p += nb;
    A.low = p; up(low, p, nb * 8); p += nb;
    A.cand_level = p; up(cand_level, p, nk * 8); p += nk;
    A.cand_entry = p; up(cand_entry, p, nk * 8); p += nk;
    if (replay_normals) {
        A.replay_normals = p; up(replay_normals, p, nt * nk * 4 * 8); p += nt * nk * 4;
        A.replay_uniforms = p; up(replay_uniforms, p, nt * nk * 8); p += nt * nk;
    }
    A.out_total = p; p += nt;
    A.out_dd = p; p += nt;
    A.cand_bar = i32; up(cand_bar, i32, nk * 4);
    A.out_wins = i32 + nk; A.out_losses = i32 + nk + nt; A.out_opens = i32 + nk + 2 * nt;
    A.cand_side = u8; up(cand_side, u8, nk);
    A.n_bars = n_bars; A.n_cand = n_cand; A.sp = stop_padding; A.tp = tp_padding;
    for (int q = 0; q < 4; ++q) A.sig[q] = sigmas[q];
    A.trials = trials; A.trial_offset = trial_offset; A.k0 = (uint32_t)seed; A.k1 = (uint32_t)(seed >> 32);
    const size_t smem = 2 * nb * 8;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(jitter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) {
        const int64_t blocks = std::min<int64_t>((trials + kThreads - 1) / kThreads, (int64_t)ctx->sm_count * 8);
        cudaEventRecord(ctx->ev0, ctx->stream);
        jitter_kernel<<<(unsigned)blocks, kThreads, smem, ctx->stream>>>(A);
        ctx->launches++;
        e = cudaGetLastError();
        cudaEventRecord(ctx->ev1, ctx->stream);
    }
    auto down = [&](void* dst, const void* src, size_t n) {
        if (e == cudaSuccess && n) e = cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, ctx->stream);
    };
    down(out_total_r, A.out_total, nt * 8); down(out_max_dd, A.out_dd, nt * 8);
    down(out_wins, A.out_wins, nt * 4); down(out_losses, A.out_losses, nt * 4); down(out_opens, A.out_opens, nt * 4);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(arena);
    if (e != cudaSuccess) return fail(ctx, QMMX_E_CUDA, std::string("run_jitter: ") + cudaGetErrorString(e));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_run_last_bars(qmmx_ctx* c, const double* high, const double* low, int32_t n_bars, const int32_t* cand_bar,
                       const uint8_t* cand_side, const double* cand_stop, const double* cand_target, int32_t n_cand,
                       int32_t* out_outcome, int32_t* out_exit_bar) {
    ENTER(c);
    if (!high || !low || n_bars < 1 || n_bars > 12000 || n_cand < 0 ||
        (n_cand > 0 && (!cand_bar || !cand_side || !cand_stop || !cand_target || !out_outcome || !out_exit_bar)))
        return fail(ctx, QMMX_E_BADARG, "run_last_bars: bad arguments (n_bars 1..12000)");
    if (n_cand == 0) return QMMX_OK;
    const size_t nb = (size_t)n_bars, nk = (size_t)n_cand;
    uint8_t* arena = nullptr;                      // f64: high, low, stop, target | i32: bar, outcome, exit_bar | u8: side
    const size_t bytes = (2 * nb + 2 * nk) * 8 + 3 * nk * 4 + nk + 64;
    CK(cudaMalloc(&arena, bytes));
    double* d64 = reinterpret_cast<double*>(arena);
    int32_t* i32 = reinterpret_cast<int32_t*>(arena + (2 * nb + 2 * nk) * 8);
    uint8_t* u8 = reinterpret_cast<uint8_t*>(i32 + 3 * nk);
    cudaError_t e = cudaSuccess;
    auto up = [&](const void* src, void* dst, size_t n) {
        if (e == cudaSuccess && n) e = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, ctx->stream);
    };
    up(high, d64, nb * 8); up(low, d64 + nb, nb * 8);
    up(cand_stop, d64 + 2 * nb, nk * 8); up(cand_target, d64 + 2 * nb + nk, nk * 8);
    up(cand_bar, i32, nk * 4); up(cand_side, u8, nk);
    if (e == cudaSuccess) {
        const int64_t blocks = std::min<int64_t>((nk * 32 + kThreads - 1) / kThreads, (int64_t)ctx->sm_count * 8);
        cudaEventRecord(ctx->ev0, ctx->stream);
        last_bars_kernel<<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(d64, d64 + nb, n_bars, i32, u8, d64 + 2 * nb,
                                                                         d64 + 2 * nb + nk, n_cand, i32 + nk, i32 + 2 * nk);
        ctx->launches++;
        e = cudaGetLastError();
        cudaEventRecord(ctx->ev1, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_outcome, i32 + nk, nk * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_exit_bar, i32 + 2 * nk, nk * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(arena);
    if (e != cudaSuccess) return fail(ctx, QMMX_E_CUDA, std::string("run_last_bars: ") + cudaGetErrorString(e));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_run_last_bars_static(qmmx_ctx* c, const double* high, const double* low, int32_t n_bars, const int32_t* cand_bar,
                              const uint8_t* cand_side, const double* cand_stop, const double* cand_target,
                              int32_t n_cand, int32_t* out_outcome, int32_t* out_exit_bar) {
    ENTER(c);
    if (!high || !low || n_bars < 1 || n_cand < 0 ||
        (n_cand > 0 && (!cand_bar || !cand_side || !cand_stop || !cand_target || !out_outcome || !out_exit_bar)))
        return fail(ctx, QMMX_E_BADARG, "run_last_bars_static: bad arguments");
    if (n_bars > kStaticMaxBars || n_cand > kStaticMaxCand)
        return fail(ctx, QMMX_E_UNSUPPORTED, "run_last_bars_static: n_bars " + std::to_string(n_bars) + " / n_cand " + std::to_string(n_cand) +
                    " beyond the static scratch (QMMX_STATIC_MAX_BARS " + std::to_string(kStaticMaxBars) + ", QMMX_STATIC_MAX_CAND " +
                    std::to_string(kStaticMaxCand) + "): use qmmx_run_last_bars");
    if (ctx->static_scratch_err != cudaSuccess || !ctx->static_scratch)
        return fail(ctx, QMMX_E_CUDA, std::string("run_last_bars_static: cudaGetSymbolAddress(g_static_scratch): ") +
                    cudaGetErrorString(ctx->static_scratch_err));
    if (n_cand == 0) return QMMX_OK;
    const size_t nb = (size_t)n_bars, nk = (size_t)n_cand;
    // same layout as qmmx_run_last_bars' arena, inside the shared array: f64 high, low, stop, target | i32 bar, outcome,
    // exit_bar | u8 side.  Nothing has to start at zero: the inputs are overwritten here and the kernel writes both
    // outputs of every candidate, so what the previous call left behind is never read.
    double* d64 = reinterpret_cast<double*>(ctx->static_scratch);
    int32_t* i32 = reinterpret_cast<int32_t*>(ctx->static_scratch + (2 * nb + 2 * nk) * 8);
    uint8_t* u8 = reinterpret_cast<uint8_t*>(i32 + 3 * nk);
    cudaError_t e = cudaSuccess;
    auto up = [&](const void* src, void* dst, size_t n) {
        if (e == cudaSuccess && n) e = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, ctx->stream);
    };
    up(high, d64, nb * 8); up(low, d64 + nb, nb * 8);
    up(cand_stop, d64 + 2 * nb, nk * 8); up(cand_target, d64 + 2 * nb + nk, nk * 8);
    up(cand_bar, i32, nk * 4); up(cand_side, u8, nk);
    if (e == cudaSuccess) {
        const int64_t blocks = std::min<int64_t>((nk * 32 + kThreads - 1) / kThreads, (int64_t)ctx->sm_count * 8);
        cudaEventRecord(ctx->ev0, ctx->stream);
        last_bars_kernel<<<(unsigned)blocks, kThreads, 0, ctx->stream>>>(d64, d64 + nb, n_bars, i32, u8, d64 + 2 * nb,
                                                                         d64 + 2 * nb + nk, n_cand, i32 + nk, i32 + 2 * nk);
        ctx->launches++;
        e = cudaGetLastError();
        cudaEventRecord(ctx->ev1, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_outcome, i32 + nk, nk * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_exit_bar, i32 + 2 * nk, nk * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return fail(ctx, QMMX_E_CUDA, std::string("run_last_bars_static: ") + cudaGetErrorString(e));
    ctx->ran = true;
    return QMMX_OK;
}

int qmmx_debug_classify(const double* price, const uint8_t* solid, int n_levels, double prox, double qmin,
                        const float* p, int64_t n, int32_t* level_table, int32_t* level_brute,
                        uint8_t* conf_mask, uint8_t* conf_brute, uint64_t* near_table, uint64_t* near_brute) {
    if (n_levels < 0 || n_levels > QMMX_MAX_LEVELS || n < 0 || (n > 0 && !p)) return QMMX_E_BADARG;
    std::vector<DevTables> T(1);
    std::string e;
    if (build_tables(n_levels, price, solid, prox, qmin, &T[0], &e) != 0) return QMMX_E_UNSUPPORTED;
    for (int64_t i = 0; i < n; ++i) {
        const float pf = p[i];
        const int kt = classify_tables(&T[0], pf), kb = classify_bruteforce(n_levels, price, prox, pf);
        if (level_table) level_table[i] = kt;
        if (level_brute) level_brute[i] = kb;
        uint8_t mt = 0, mb = 0;
        if (kb >= 0) {
            for (int b = 0; b < 3; ++b) {
                if (kt >= 0 && pf >= T[0].lev[kt].cpass[b][0] && pf <= T[0].lev[kt].cpass[b][1]) mt |= 1u << b;
                if (!(confidence_f64(std::fabs(price[kb] - (double)pf), solid[kb] != 0, b + 1, prox) < qmin)) mb |= 1u << b;
            }
        }
        if (conf_mask) conf_mask[i] = mt;
        if (conf_brute) conf_brute[i] = mb;
        uint64_t nt = 0, nbm = 0;
        for (int j = 0; j < n_levels; ++j) {
            if (pf >= T[0].lev[j].wlo && pf <= T[0].lev[j].whi) nt |= 1ull << j;
            if (!(std::fabs(price[j] - (double)pf) > prox)) nbm |= 1ull << j;
        }
        if (T[0].ncls >= 0) {
            // the class partition of the lean / tile kernels: its level must be the level table's, and its window mask is what
            // near_table reports (the mask of a far price is empty, as is the set of windows that hold it)
            uint64_t cm = 0;
            const int kf = classify_tables_fine(&T[0], pf, &cm);
            if (kf != kt && level_table) level_table[i] = -100;
            nt = cm;
        }
        if (near_table) near_table[i] = nt;
        if (near_brute) near_brute[i] = nbm;
    }
    return QMMX_OK;
}

int qmmx_debug_class_partition(const double* price, const uint8_t* solid, int n_levels, double prox, double qmin,
                               int32_t* n_classes, int32_t* n_runs, int32_t* n_records) {
    if (n_levels < 0 || n_levels > QMMX_MAX_LEVELS) return QMMX_E_BADARG;
    std::vector<DevTables> T(1);
    std::string e;
    if (build_tables(n_levels, price, solid, prox, qmin, &T[0], &e) != 0) return QMMX_E_UNSUPPORTED;
    if (n_classes) *n_classes = T[0].ncls;
    if (n_runs) *n_runs = T[0].FZ;
    if (n_records) *n_records = T[0].fnmixed;
    return QMMX_OK;
}

int qmmx_last_timing(qmmx_ctx* c, float* ms) {
    ENTER(c);
    if (!ctx->ran || !ms) return fail(ctx, QMMX_E_STATE, "nothing has run");
    CK(cudaEventSynchronize(ctx->ev1));
    CK(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return QMMX_OK;
}

int qmmx_download_policy_weights(qmmx_ctx* c, double* out, int64_t paths) {
    ENTER(c);
    if (!out || paths < 0) return fail(ctx, QMMX_E_BADARG, "download_policy_weights: bad arguments");
    if (!ctx->d_polw || paths != ctx->polw_paths) return fail(ctx, QMMX_E_STATE, "no ONLINE_RETRAIN scan of that many paths has run");
    std::vector<double> tmp((size_t)14 * (size_t)ctx->polw_pitch);
    CK(cudaMemcpyAsync(tmp.data(), ctx->d_polw, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int64_t p = 0; p < paths; ++p)
        for (int q = 0; q < 14; ++q) out[p * 14 + q] = tmp[(size_t)q * (size_t)ctx->polw_pitch + (size_t)p];
    return QMMX_OK;
}

int qmmx_measure_fp32_peak(qmmx_ctx* c, double* tflops) {
    ENTER(c);
    if (!tflops) return fail(ctx, QMMX_E_BADARG, "tflops is null");
    const int blocks = ctx->sm_count * 2, threads = 1024, iters = 4096;
    float* d = nullptr;
    CK(cudaMalloc(&d, (size_t)blocks * threads * sizeof(float)));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 0.f;
    for (int rep = 0; rep < 4; ++rep) {      // first repetition is the warm-up
        cudaEventRecord(a, ctx->stream);
        fp32_probe_kernel<<<blocks, threads, 0, ctx->stream>>>(d, iters, 0.999f, 1e-3f);
        cudaEventRecord(b, ctx->stream);
        cudaEventSynchronize(b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        ctx->launches++;
        if (rep > 0 && (best == 0.f || ms < best)) best = ms;
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    cudaFree(d);
    CK(cudaGetLastError());
    *tflops = 2.0 * 64.0 * (double)iters * (double)blocks * threads / ((double)best * 1e-3) / 1e12;
    return QMMX_OK;
}

int qmmx_measure_issue_peak(qmmx_ctx* c, double* alu_gwips, double* mixed_gwips) {
    ENTER(c);
    if (!alu_gwips || !mixed_gwips) return fail(ctx, QMMX_E_BADARG, "null output");
    const int blocks = ctx->sm_count, threads = 1024, iters = 2048;
    uint32_t* d = nullptr;
    CK(cudaMalloc(&d, (size_t)blocks * threads * sizeof(uint32_t)));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double out[2] = {0.0, 0.0};
    for (int mode = 0; mode < 2; ++mode) {
        float best = 0.f;
        for (int rep = 0; rep < 4; ++rep) {      // first repetition is the warm-up
            cudaEventRecord(a, ctx->stream);
            if (mode == 0) issue_probe_kernel<0><<<blocks, threads, 0, ctx->stream>>>(d, iters, 0x9e3779b9u, 0x7f4a7c15u, 0.999f, 1e-3f);
            else issue_probe_kernel<1><<<blocks, threads, 0, ctx->stream>>>(d, iters, 0x9e3779b9u, 0x7f4a7c15u, 0.999f, 1e-3f);
            cudaEventRecord(b, ctx->stream);
            cudaEventSynchronize(b);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, a, b);
            ctx->launches++;
            if (rep > 0 && (best == 0.f || ms < best)) best = ms;
        }
        // warp instructions of the probed kind(s) per second, whole device, in units of 1e9
        const double per_thread = 64.0 * (double)iters * (mode == 0 ? 1.0 : 2.0);
        out[mode] = per_thread * (double)blocks * (threads / 32) / ((double)best * 1e-3) / 1e9;
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    cudaFree(d);
    CK(cudaGetLastError());
    *alu_gwips = out[0]; *mixed_gwips = out[1];
    return QMMX_OK;
}

int qmmx_launch_count(qmmx_ctx* c, uint64_t* n) {
    if (!c || !n) return QMMX_E_BADARG;
    *n = c->launches;
    return QMMX_OK;
}

}  // extern "C"
