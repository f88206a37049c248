// Device-side QMMX engine: one thread walks one price path through the entry gates
// (qmmx_monolithic.py:1492-1771), the live loop's exit block (M:2966-3014) and the OnlinePolicy gate
// (M:3046-3109).  Semantics frozen in SURVEY.md Appendix A; checked bit-for-bit against oracle/.
//
// Arithmetic contract: prices are fp32 in memory and the reference evaluates them in float64.  Every
// decision that depends only on the price (nearest level, proximity, latch release, confidence threshold,
// stop/target touch) is taken with fp32 compares against thresholds the host derived EXACTLY from the
// float64 predicates (tables.h); the logistic gates use host-evaluated float64 probabilities (same libm as
// CPython) and a float64 Neumaier dot product on the device.  The file is compiled with --fmad=false so no
// multiply-add is ever contracted.
//
// Performance discipline (measured with ncu, profiles/): a warp takes nearly every branch of the state
// machine on every step (some lane enters, some lane exits, some lane touches a level), so the cost of a step
// is the SUM of all branches.  Therefore: per-path state is a bag of scalars whose address is never taken (one
// address-taken member sends the whole struct to local memory); run constants are kernel parameters (constant
// bank operands, no loads); block tables are STATIC shared memory in structure-of-arrays form (absolute LDS
// addresses, bank = level index); rare work sits in __noinline__ functions.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "qmmx_mc.h"
#include "tables.h"

namespace qmmx {

constexpr int U_NPATHS = 0, U_NSTEPS = 1, U_NTRADES = 2, U_NWINS = 3, U_NOPEN = 4, U_NOVERFLOW = 5;
constexpr int U_ENTRY = 6, U_EXIT = U_ENTRY + QMMX_N_ENTRY_CODES, U_HIST = U_EXIT + QMMX_N_EXIT_CODES;
constexpr int D_SUM_PNL = 0, D_SUM_PNL2 = 1, D_SUM_PATH2 = 2, D_SUM_MTM = 3;

constexpr uint32_t kCbHasMl = 1u << 16, kCbHasProb = 1u << 17;   // internal DevCombo.flags bits

// Kernel parameters of one scan launch = one parameter combo over one set of paths (constant bank).
struct ScanArgs {
    const float* close; const float* high; const float* low;   // [steps][pitch] fp32
    const uint8_t* regime;        // AccumulationBreakoutGuard regime per tick [steps][pitch] (regime_kernel), or null
    const float* vol;             // per-bar volume [steps][pitch], or null: the soft veto is inert (M:1796-1811)
    double* polw;                 // QMMX_F_ONLINE_RETRAIN: learnt go_long / go_short heads [14][pitch], or null
    int64_t pitch, paths;
    int32_t steps, combo_index;
    const DevTables* tables;                                    // the table of this combo
    unsigned long long* flat;                                   // reduction buffer of this combo
    const uint32_t* mlmask;       // [n_days][kMaxLevels]: bit side*16+tc = ML gate FAILS; null = gate off / no model
    const double* mlprob;         // [n_days][kMaxLevels][2][16] P(win) for blend mode; null when unused
    const double* pol;            // [n_pol][3][7], then [n_pol][3]: sum of |w| per head
    double z_go, z_skip;          // sigmoid(z) >= 0.60 <=> z >= z_go ; sigmoid(z) < 0.55 <=> z < z_skip
    double hist_lo, hist_inv_w;
    // parity-mode outputs (may be null)
    qmmx_trade* trades; int32_t* tcounts; uint8_t* codes;
    int32_t max_trades, n_days, n_pol, n_levels;
    int32_t cell_obase, cell_shift, cell_ncell;        // grid of the table (tables.h cell_of_bits), as constants
    uint32_t cell_lim;                                 // (ncell << shift) - 1: cell = min(bits - obase, cell_lim) >> shift (one add-min + one shift)
    float guard[4];                                    // far zones adjoining the grid (tables.h DevTables::guard)
    uint32_t mix_base;                                 // class partition: cell bytes >= mix_base are not class ids (tables.h DevTables::ncls)
    int32_t n_out;                                     // class runs outside the grid (tables.h DevTables::n_out), < 0: use the zone search
    float grid_lo, grid_hi, out_lo[4], out_hi[4];
    uint32_t out_cls[4];
    int32_t cell_bytes;                                // dynamic shared memory reserved for the cell bytes (scan_lean.cuh carves the rest)
    unsigned int* tile_counter;                        // scan_tile.cuh: next unclaimed 32-path tile of this launch (zeroed by init_flat_kernel)
    DevCombo combo;
};

// ---- shared-memory block state (static; the grid-cell bytes are the only dynamic part) -------------------
struct ColdLevel { double price; float wlo, whi; uint32_t flags, pad; };

struct BlockShared {
    // contact path, structure of arrays indexed by level: bank = level (mod 32)
    uint32_t cand_lo[kMaxLevels], cand_hi[kMaxLevels];   // levels whose contact window overlaps this level's
    float2 cpass[3][kMaxLevels];        // touch bucket: conf >= Q_MIN_PROB  <=>  x <= p <= y
    float2 thr[2][kMaxLevels];          // [side] {lo_thr, hi_thr}: long {stop v, target ^}, short {target v, stop ^}
    ColdLevel cold[kMaxLevels];
    float zhi[kMaxZones];
    uint8_t zlev[kMaxZones];
    int32_t L, Z, ncell, shift, obase, pad0;
    // block aggregate, flushed to the global reduction buffer with atomics
    unsigned long long u[U_HIST];
    uint32_t hist[QMMX_PNL_HIST_BINS];  // per-path PnL histogram (32-bit: native shared-memory atomics)
    double d[QMMX_AGG_ND];
    double xmax, xnegmin;
};

__shared__ BlockShared g_S;
// per-thread touch counters of the lockstep-v1 engine: 4-bit saturating, 64 levels -> 8 words per thread,
// [word][thread].  A separate symbol so that kernels which do not use it (scan_lean.cuh) do not pay for it.
__shared__ uint32_t g_touch[8 * kThreads];
extern __shared__ __align__(16) uint8_t g_cell[];   // level index / kFar / kMixed per grid cell

__device__ __forceinline__ void smem_inc(unsigned long long* p, unsigned long long v = 1ull) { atomicAdd(p, v); }

// OnlinePolicy._dot under CPython >= 3.12 (Neumaier-compensated builtin sum), M:334-335
__device__ __noinline__ double dot_py312(const double* __restrict__ w, double x1, double x3, double x4, double x5, double x6) {
    // x = [1, x1, 0, x3, x4, x5, x6]
    double f = 0.0 + w[0] * 1.0;
    double c = 0.0;
#pragma unroll
    for (int i = 1; i < QMMX_POLICY_DIM; ++i) {
        const double xi = i == 1 ? x1 : i == 2 ? 0.0 : i == 3 ? x3 : i == 4 ? x4 : i == 5 ? x5 : x6;
        const double v = w[i] * xi;
        const double t = f + v;
        if (fabs(f) >= fabs(v)) c += (f - t) + v; else c += (v - t) + f;
        f = t;
    }
    if (c != 0.0 && isfinite(c)) f += c;
    return f;
}

__device__ __forceinline__ double compute_confidence(double dist, bool solid, int tc, double pmax) {   // M:1415-1427
    double base = 1.0 - (dist / pmax);
    if (!(base > 0.0)) base = 0.0;
    base += solid ? 0.08 : 0.02;
    if (tc <= 1) base += 0.10;
    else if (tc == 2) base -= 0.08;
    else base -= (0.08 * 2);
    base += 0.03;
    const double m = base < 1.0 ? base : 1.0;
    return m > 0.0 ? m : 0.0;
}

// rare: neighbouring levels are latched while level k is touched -> each stays latched only while p is still
// inside its own contact window (M:1572-1576).  Returns the latch bits to keep.
__device__ __noinline__ unsigned long long keep_neighbours(float pf, uint32_t klo, uint32_t khi) {
    uint32_t llo = 0u, lhi = 0u;
    while (klo) {
        const int j = __ffs((int)klo) - 1; klo &= klo - 1;
        if (pf >= g_S.cold[j].wlo && pf <= g_S.cold[j].whi) llo |= 1u << j;
    }
    while (khi) {
        const int j = __ffs((int)khi) - 1; khi &= khi - 1;
        if (pf >= g_S.cold[32 + j].wlo && pf <= g_S.cold[32 + j].whi) lhi |= 1u << j;
    }
    return ((unsigned long long)lhi << 32) | llo;
}

// rare: the grid cell straddles a zone boundary -> exact binary search over the zone table
__device__ __noinline__ uint32_t classify_zones(float pf) {
    int a = 0, b = g_S.Z - 1;
    while (a < b) {
        const int m = (a + b) >> 1;
        if (pf <= g_S.zhi[m]) b = m; else a = m + 1;
    }
    return g_S.zlev[a];
}

// rare: blend gate (M:1711-1742) in float64; true when the blended score is below Q_MIN_PROB
__device__ __noinline__ bool blend_fails(const ScanArgs* A, int k, float pf, int tc, int side, int day_off) {
    const double lp = g_S.cold[k].price;
    const double conf = compute_confidence(fabs(lp - (double)pf), (g_S.cold[k].flags & kLvSolid) != 0, tc, A->combo.pmax);
    double mlp = conf;
    if (A->mlprob) mlp = A->mlprob[(((size_t)day_off + k) * 2 + side) * 16 + tc];
    const double blended = A->combo.wr * conf + A->combo.wm * mlp;
    const bool fails = blended < A->combo.qmin;
    if (fails) smem_inc(&g_S.u[U_ENTRY + QMMX_COMBINED_LOW]);
    return fails;
}

// rare: accumulation / breakout regime gate (M:1652-1666, int code 904): counter-trend entries are blocked after a
// volume-confirmed break of the 60-minute box.  Returns 0 or the reason code.
__device__ __noinline__ int regime_gate(const ScanArgs* A, int side, int i, int64_t path) {
    const int r = __ldg(A->regime + (size_t)i * (size_t)A->pitch + path);
    if ((r == 2 && side == 1) || (r == 3 && side == 0)) {
        smem_inc(&g_S.u[U_ENTRY + QMMX_ACC_BREAKOUT_GATE]);
        return QMMX_ACC_BREAKOUT_GATE;
    }
    return 0;
}

// builtin sum() of n <= 3 floats under CPython >= 3.12 (Neumaier-compensated), as _calc_volume_slope uses it
__device__ __forceinline__ double sum3_py312(double a, double b, double c3, int n) {
    double f = 0.0 + a, c = 0.0;
    {
        const double t = f + b;
        if (fabs(f) >= fabs(b)) c += (f - t) + b; else c += (b - t) + f;
        f = t;
    }
    if (n > 2) {
        const double t = f + c3;
        if (fabs(f) >= fabs(c3)) c += (f - t) + c3; else c += (c3 - t) + f;
        f = t;
    }
    if (c != 0.0 && isfinite(c)) f += c;
    return f;
}

// rare: behavioural soft veto (M:1678-1705; _soft_veto M:1773-1794; _calc_volume_slope M:1796-1811 over
// engine.recent_bars = the last min(i+1, 240) bars, window 6).  Returns 0 (allow) or the veto reason code.
// The side follows the direction (M:1645), so of the four branches of _soft_veto only "approaching with falling
// volume" can fire: from_below & long & slope < -strong, from_above & short & slope < -strong.
__device__ __noinline__ int veto_check(const ScanArgs* A, int k, float pf, int side, int i, int64_t path) {
    const int len = i + 1 < 240 ? i + 1 : 240;
    double vs = 0.0;
    if (len >= 3) {
        const int n = len < 6 ? len : 6;
        double v0 = 0, v1 = 0, v2 = 0, v3 = 0, v4 = 0, v5 = 0;      // newest six, oldest first from index 6 - n
        const float* col = A->vol + path;
        const size_t pitch = (size_t)A->pitch;
        v5 = (double)__ldg(col + (size_t)i * pitch); v4 = (double)__ldg(col + (size_t)(i - 1) * pitch);
        v3 = (double)__ldg(col + (size_t)(i - 2) * pitch);
        if (n > 3) v2 = (double)__ldg(col + (size_t)(i - 3) * pitch);
        if (n > 4) v1 = (double)__ldg(col + (size_t)(i - 4) * pitch);
        if (n > 5) v0 = (double)__ldg(col + (size_t)(i - 5) * pitch);
        // `b.get("v") or ... or 0.0`: a zero volume of either sign becomes +0.0; adding 0.0 does exactly that
        v0 += 0.0; v1 += 0.0; v2 += 0.0; v3 += 0.0; v4 += 0.0; v5 += 0.0;
        // vols = [oldest .. newest] of length n; half = max(2, n // 2); first half vs last half
        double a1, a2;
        if (n == 6) { a1 = sum3_py312(v0, v1, v2, 3) / 3.0; a2 = sum3_py312(v3, v4, v5, 3) / 3.0; }
        else if (n == 5) { a1 = sum3_py312(v1, v2, 0.0, 2) / 2.0; a2 = sum3_py312(v4, v5, 0.0, 2) / 2.0; }
        else if (n == 4) { a1 = sum3_py312(v2, v3, 0.0, 2) / 2.0; a2 = sum3_py312(v4, v5, 0.0, 2) / 2.0; }
        else { a1 = sum3_py312(v3, v4, 0.0, 2) / 2.0; a2 = sum3_py312(v4, v5, 0.0, 2) / 2.0; }
        vs = (a1 == 0.0 && a2 == 0.0) ? 0.0 : (a2 - a1) / (fabs(a1) + 1e-9);
    }
    const bool confl = (g_S.cold[k].flags & kLvConfl015) != 0;        // _has_confluence_near(level, 0.15), M:1885
    if (fabs(vs) < 0.05 && !confl) return 0;                          // INCONCLUSIVE
    const double dist = fabs(g_S.cold[k].price - (double)pf);
    const double pw = A->combo.prox * 0.12;
    const bool near_ = dist <= (A->combo.vprox > pw ? A->combo.vprox : pw);
    if (near_ && vs < -A->combo.vstrong) {
        const int code = side == 0 ? QMMX_VETO_CONTRA_VOL_LONG : QMMX_VETO_CONTRA_VOL_SHORT;
        smem_inc(&g_S.u[U_ENTRY + code]);
        return code;
    }
    return 0;
}

// rare: OnlinePolicy gate of the live loop (M:3046-3109); true = entry allowed.  LEARNT: the head of the entry's side is
// the one this path has learnt so far (QMMX_F_ONLINE_RETRAIN, A->polw) instead of the table's.
template <bool LEARNT>
__device__ __noinline__ bool policy_passes(const ScanArgs* A, int k, float pf, int side, int i, int64_t path) {
    if (A->n_pol <= 0) { smem_inc(&g_S.u[U_ENTRY + QMMX_ONLINE_POLICY]); return false; }   // zero weights: sigmoid(0) = 0.5 < 0.60
    const int spd = A->combo.spd;
    const double lp = g_S.cold[k].price;
    const double pa = fabs((double)pf - lp);
    const int day = i < spd ? 0 : i / spd;                                   // one division, and none on the first day
    const double tod = (double)(i - day * spd) / 390.0;
    const double x1 = pa < 1.0 ? pa : 1.0;
    const double x3 = side == 1 ? 1.0 : 0.0, x4 = side == 0 ? 1.0 : 0.0;
    const bool confl = (g_S.cold[k].flags & kLvConfl06) != 0;
    const double x5 = confl ? 1.0 : 0.0, x6 = tod < 1.0 ? tod : 1.0;
    const int dsel = day < A->n_pol ? day : A->n_pol - 1;
    const double* w = A->pol + QMMX_POLICY_HEADS * QMMX_POLICY_DIM * dsel;
    // Fast path: the plain float64 sum of the same products differs from CPython's compensated sum by at most a few
    // roundings of the partial sums, each below 2^-53 * sum|w| (every |x| <= 1), so outside a band of 2^-48 * sum|w|
    // around the threshold both sums are on the same side of it; inside the band (practically never) the exact sum decides.
    const double* sabs = A->pol + (size_t)QMMX_POLICY_HEADS * QMMX_POLICY_DIM * A->n_pol + QMMX_POLICY_HEADS * dsel;
    auto plain = [&](const double* q) { return q[0] + q[1] * x1 + (side == 1 ? q[3] : q[4]) + (confl ? q[5] : 0.0) + q[6] * x6; };
    const double band = 3.552713678800501e-15;                              // 2^-48
    bool go;
    if (LEARNT) {
        double wl[QMMX_POLICY_DIM];
#pragma unroll
        for (int q = 0; q < QMMX_POLICY_DIM; ++q) wl[q] = A->polw[(size_t)(side * QMMX_POLICY_DIM + q) * (size_t)A->pitch + (size_t)path];
        go = dot_py312(wl, x1, x3, x4, x5, x6) >= A->z_go;
    } else {
        const double* q = w + QMMX_POLICY_DIM * side;
        const double z = plain(q), b = sabs[side] * band;
        if (z >= A->z_go + b) go = true;
        else if (z < A->z_go - b) go = false;
        else go = dot_py312(q, x1, x3, x4, x5, x6) >= A->z_go;
    }
    bool pass = false;
    if (go) {                                                                // the skip head matters only when the go head passes
        const double* q = w + QMMX_POLICY_DIM * 2;
        const double z = plain(q), b = sabs[2] * band;
        if (z < A->z_skip - b) pass = true;
        else if (z >= A->z_skip + b) pass = false;
        else pass = dot_py312(q, x1, x3, x4, x5, x6) < A->z_skip;
    }
    if (!pass) smem_inc(&g_S.u[U_ENTRY + QMMX_ONLINE_POLICY]);
    return pass;
}

// M:20-26
__device__ __forceinline__ double sigmoid_policy(double x) {
    if (x < -50.0) return 0.0;
    if (x > 50.0) return 1.0;
    return 1.0 / (1.0 + exp(-x));
}

// QMMX_F_ONLINE_RETRAIN, at the close of a trade: the trade's two labelled policy_events rows train the head of its
// action (update_entry M:357-363 / _sgd_update M:337-341, lr 0.03, l2 1e-6 of M:290), in row order: the PRE_OPEN row
// of evaluate_entry (M:1757-1768 -> features [1, min(1, |p - L|), 0, from_above, from_below, 0, 0]), then the live
// loop's {"x": x} row (M:3139-3150), whose keys _retrain_from_labeled_events cannot read: [1, 0, 0, 1, 0, 0, 0].
// Restated from oracle/headless.py run_path; exp() is CUDA's (<= 1 ulp from libm's), so the learnt weights are
// compared with a tolerance and the decisions they drive exactly.
__device__ __noinline__ void policy_learn(const ScanArgs* A, int64_t path, int side, double dist, bool win) {
    double w[QMMX_POLICY_DIM];
    double* g = A->polw + (size_t)(side * QMMX_POLICY_DIM) * (size_t)A->pitch + (size_t)path;
#pragma unroll
    for (int q = 0; q < QMMX_POLICY_DIM; ++q) w[q] = g[(size_t)q * (size_t)A->pitch];
    const double y = win ? 1.0 : 0.0;
#pragma unroll 1
    for (int row = 0; row < 2; ++row) {
        const double x1 = row == 0 ? (dist < 1.0 ? dist : 1.0) : 0.0;
        const double x3 = row == 0 ? (side == 1 ? 1.0 : 0.0) : 1.0, x4 = row == 0 ? (side == 0 ? 1.0 : 0.0) : 0.0;
        const double pred = sigmoid_policy(dot_py312(w, x1, x3, x4, 0.0, 0.0));
#pragma unroll
        for (int q = 0; q < QMMX_POLICY_DIM; ++q) {
            const double xq = q == 0 ? 1.0 : q == 1 ? x1 : q == 3 ? x3 : q == 4 ? x4 : 0.0;
            const double grad = (pred - y) * xq + 1e-6 * w[q];
            w[q] -= 0.03 * grad;
        }
    }
#pragma unroll
    for (int q = 0; q < QMMX_POLICY_DIM; ++q) g[(size_t)q * (size_t)A->pitch] = w[q];
}

// parity mode: one trade record
__device__ __noinline__ void record_trade(const ScanArgs* A, int64_t path, int ntr, int entry_bar, int exit_bar, int side,
                                          int reason, int k, float entry, float exit, float pnl) {
    if (!A->trades) return;
    if (ntr < A->max_trades) {
        qmmx_trade t;
        t.entry_bar = entry_bar; t.exit_bar = exit_bar; t.side = (uint8_t)side; t.reason = (uint8_t)reason;
        t.level_id = (uint16_t)k; t.entry = entry; t.exit = exit; t.pnl = pnl;
        A->trades[path * A->max_trades + ntr] = t;
    } else smem_inc(&g_S.u[U_NOVERFLOW]);
}

template <bool WIDE> struct LatchType { typedef uint32_t T; };
template <> struct LatchType<true> { typedef unsigned long long T; };

// ---- per-path state machine ---------------------------------------------------------------------------
// WIDE = more than 32 levels (64-bit contact latch).  PARITY adds per-step codes and trade records.
template <bool HL, bool PARITY, bool TINY, bool WIDE>
struct PathEngine {
    typedef typename LatchType<WIDE>::T latch_t;
    uint32_t* tptr;        // &g_touch[tid]: this thread's column of touch counters
    // EngineState (M:259-270) + contact latch (M:1376)
    float prev;            // NaN until the first tick
    int d;                 // last_direction: 0 unknown, 1 up, 2 down
    int cool_until;        // first step index at which the cooldown is over
    int pos;               // -1 flat, else level | side << 8
    float lo_thr, hi_thr;  // exit when low <= lo_thr or high >= hi_thr (directed-rounded stop/target)
    float entry;
    latch_t latch;
    // per-path tallies, two 16-bit counters per register (steps per path <= 65535); reduced per warp at path end
    uint32_t c_gate;       // LEVEL_OVERTOUCHED | CONF_LOW << 16
    uint32_t c_ml;         // ML_CONF_LOW | winning trades << 16
    uint32_t c_exit;       // STOP exits | TARGET exits << 16
    uint32_t c_span;       // IN_POSITION steps | COOLDOWN steps << 16
    uint32_t c_dir;        // DIR_UNKNOWN
    // per-path accumulators
    double pnl, pnl2, mtm;
    // parity mode only
    int64_t path; int entry_bar; int ntr;

    __device__ __forceinline__ void reset(int tid, int64_t path_) {
        tptr = &g_touch[tid]; path = path_;
        prev = __int_as_float(0x7fc00000);
        d = 0; cool_until = 0; pos = -1; lo_thr = 0.f; hi_thr = 0.f; entry = 0.f; entry_bar = 0;
        latch = 0; c_gate = 0u; c_ml = 0u; c_exit = 0u; c_span = 0u; c_dir = 0u;
        pnl = 0.0; pnl2 = 0.0; mtm = 0.0; ntr = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) tptr[w * kThreads] = 0u;
    }

    // contact path: price is nearest to, and within CONTACT_PROX of, level k
    __device__ __forceinline__ int contact(const ScanArgs& A, int k, float pf, int dir, int i, int day_off) {
        // latch + touch count (M:1562-1576)
        latch_t bit, cand;
        if (WIDE) {
            bit = (latch_t)(1ull << k);
            cand = (latch_t)(((unsigned long long)g_S.cand_hi[k] << 32) | g_S.cand_lo[k]);
        } else {
            bit = (latch_t)(1u << k);
            cand = (latch_t)g_S.cand_lo[k];
        }
        const bool isnew = (latch & bit) == 0;
        const latch_t keep = latch & cand;
        latch = bit;
        if (keep) latch |= (latch_t)keep_neighbours(pf, (uint32_t)keep, WIDE ? (uint32_t)((unsigned long long)keep >> 32) : 0u);
        uint32_t* tw = tptr + (k >> 3) * kThreads;
        const int sh = (k & 7) * 4;
        const uint32_t word = *tw;
        int tc = (word >> sh) & 15;
        if (isnew && tc < 15) { ++tc; *tw = word + (1u << sh); }
        if (tc >= A.combo.overtouch) { c_gate += 1u; return QMMX_LEVEL_OVERTOUCHED; }     // M:1579
        const float2 cp = g_S.cpass[min(tc, 3) - 1][k];
        if (!(pf >= cp.x && pf <= cp.y)) { c_gate += 1u << 16; return QMMX_CONF_LOW; }   // M:1628
        const int side = dir - 1;                                            // M:1645: up -> long(0), down -> short(1)
        if (A.regime) { const int rc = regime_gate(&A, side, i, path); if (rc) return rc; }       // M:1652-1666
        if (A.vol) { const int vc = veto_check(&A, k, pf, side, i, path); if (vc) return vc; }   // M:1678-1705
        if (A.combo.flags & QMMX_F_USE_BLEND) {                              // M:1730-1742
            if (blend_fails(&A, k, pf, tc, side, day_off)) return QMMX_COMBINED_LOW;
        } else if (A.mlmask) {                                               // M:1454-1466, M:1745-1755
            const uint32_t m = __ldg(&A.mlmask[day_off + k]);
            if ((m >> (side * 16 + tc)) & 1u) { c_ml += 1u; return QMMX_ML_CONF_LOW; }
        }
        // evaluate_entry returned OK; OnlinePolicy gate of the live loop (M:3046-3109)
        if (!(A.combo.flags & QMMX_F_DISABLE_POLICY_GATE)) {
            if (!policy_passes<false>(&A, k, pf, side, i, path)) return QMMX_ONLINE_POLICY;
        }
        // open_trade (M:3112-3115, M:1888-1897): entry at the tick price; first exit test on the next tick.
        // thresholds (M:1669-1674) were rounded per level at CTA setup:
        // p <= stop (float64, p fp32)  <=>  p <= round_down_fp32(stop);  p >= target  <=>  p >= round_up_fp32(target)
        const float2 th = g_S.thr[side][k];
        lo_thr = th.x; hi_thr = th.y;
        pos = k | (side << 8);
        entry = pf;
        if (PARITY) entry_bar = i;
        return QMMX_OK_ENTRY;
    }

    __device__ __forceinline__ void do_exit(const ScanArgs& A, float pf, bool hit_lo, bool hit_hi, int i) {
        const int side = pos >> 8;
        // long: stop below (tested first), target above; short: stop above (tested first), target below
        const bool is_stop = (side == 0) ? hit_lo : hit_hi;
        double tp;
        float exit_f = pf;
        if (!HL) {
            // close exit (M:2979-3012): exit at the tick price; fp32 values, float64 difference (M:1926)
            const double diff = (double)pf - (double)entry;
            tp = side == 0 ? diff : -diff;       // e - a == -(a - e) exactly
        } else {    // exit at the stop/target level itself (M:3636)
            const double lp = g_S.cold[pos & 0xff].price;
            double exit_px;
            if (side == 0) exit_px = is_stop ? lp - A.combo.sp : lp + A.combo.tp;
            else exit_px = is_stop ? lp + A.combo.sp : lp - A.combo.tp;
            const double e = (double)entry;
            tp = (side == 0) ? (exit_px - e) : (e - exit_px);
            exit_f = (float)exit_px;
        }
        pnl += tp; pnl2 += tp * tp;
        cool_until = i + A.combo.cool_steps;                                 // M:2982, M:269-270
        if (tp > 0) c_ml += 1u << 16;                                        // label, M:1936
        c_exit += is_stop ? 1u : (1u << 16);
        if (PARITY) record_trade(&A, path, ntr, entry_bar, i, side, is_stop ? QMMX_EXIT_STOP : QMMX_EXIT_TARGET, pos & 0xff, entry, exit_f, (float)tp);
        ++ntr;
        pos = -1;
    }

    // price -> level index touched, or kFar (exact: see tables.h)
    __device__ __forceinline__ uint32_t classify(const ScanArgs& A, float pf) const {
        uint32_t v = g_cell[cell_of_bits(__float_as_int(pf), A.cell_obase, A.cell_shift, A.cell_ncell)];
        if (v >= kMixBase && v != kFar) v = classify_zones(pf);   // any straddling cell: exact zone search
        return v;
    }

    // one tick: price pf = bar i's close (hh/ll = its high/low when HL); day_off = min(day, n_days-1) * kMaxLevels
    __device__ __forceinline__ void step(const ScanArgs& A, float pf, float hh, float ll, int i, int day_off) {
        int code = QMMX_TOO_FAR;
        // direction (M:1529-1537) and last_direction update (M:2954-2955)
        int dir = d;
        if (!TINY) {
            // |p|,|prev| >= 2^-5 (QMMX_F_TINY_PRICES unset): fp32 spacing > 1.8e-9, so the +-1e-9 guard band of
            // the reference cannot change any comparison and dir doubles as the updated last_direction
            if (pf < prev) dir = 2;
            if (pf > prev) dir = 1;
        } else {
            dir = 0;
            if (prev == prev) {
                const double p = (double)pf, q = (double)prev;
                if (p > q + 1e-9) dir = 1; else if (p < q - 1e-9) dir = 2; else dir = d;
            }
        }
        if (pos >= 0) {
            code = QMMX_IN_POSITION;                                         // M:1508
            c_span += 1u;
            const bool hit_lo = (HL ? ll : pf) <= lo_thr;
            const bool hit_hi = (HL ? hh : pf) >= hi_thr;
            if (hit_lo | hit_hi) do_exit(A, pf, hit_lo, hit_hi, i);
        } else if (i < cool_until) {
            code = QMMX_COOLDOWN;                                            // M:1516
            c_span += 1u << 16;
        } else if (dir == 0) {
            code = QMMX_DIR_UNKNOWN;                                         // M:1538 (NOLEVELS when L == 0, fixed at flush)
            ++c_dir;
        } else {
            const uint32_t v = classify(A, pf);
            if (v != kFar) code = contact(A, (int)v, pf, dir, i, day_off);   // else TOO_FAR (M:1545), or NOLEVELS (M:1524)
        }
        if (!TINY) d = dir;
        else if (prev == prev && pf != prev) d = (pf > prev) ? 1 : 2;
        prev = pf;
        if (PARITY && A.codes) {
            if (A.n_levels == 0 && (code == QMMX_TOO_FAR || code == QMMX_DIR_UNKNOWN)) code = QMMX_NOLEVELS;
            A.codes[(int64_t)i * A.paths + path] = (uint8_t)code;
        }
    }

    // end of path: open position => OPEN, realised 0, mark-to-market reported separately
    __device__ __forceinline__ void finish(const ScanArgs& A) {
        if (pos >= 0) {
            const int side = pos >> 8;
            const double e = (double)entry, last = (double)prev;
            mtm = (side == 0) ? (last - e) : (e - last);
            if (PARITY) record_trade(&A, path, ntr, entry_bar, -1, side, QMMX_EXIT_OPEN, pos & 0xff, entry, prev, 0.0f);
            ++ntr;
        }
        if (PARITY && A.tcounts) A.tcounts[path] = ntr;
        double fb = floor((pnl - A.hist_lo) * A.hist_inv_w);
        const int bin = fb < 0 ? 0 : (fb > QMMX_PNL_HIST_BINS - 1 ? QMMX_PNL_HIST_BINS - 1 : (int)fb);
        atomicAdd(&g_S.hist[bin], 1u);
    }
};

// ---- block-level helpers ---------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const double t = __shfl_down_sync(0xffffffffu, v, o); v = t > v ? t : v; }
    return v;
}

__device__ __forceinline__ void atomic_max_f64(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (__longlong_as_double((long long)old) < v) {
        const unsigned long long assumed = old;
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
        if (old == assumed) break;
    }
}

// add one finished path per lane (inactive lanes contribute neutral values) to the block aggregate
__device__ __noinline__ void warp_accumulate(bool active, double pnl, double pnl2, double mtm, uint32_t cg, uint32_t cm,
                                             uint32_t ce, uint32_t cs, uint32_t cd, bool open) {
    BlockShared& S = g_S;
    const int lane = threadIdx.x & 31;
    const double neg_inf = __longlong_as_double((long long)0xfff0000000000000ull);
    if (!active) { pnl = 0.0; pnl2 = 0.0; mtm = 0.0; cg = cm = ce = cs = cd = 0u; open = false; }
    const double s1 = warp_sum(pnl);
    const double s2 = warp_sum(pnl2);
    const double s3 = warp_sum(pnl * pnl);
    const double s4 = warp_sum(mtm);
    const double mx = warp_max(active ? pnl : neg_inf);
    const double mn = warp_max(active ? -pnl : neg_inf);
    // 16-bit fields: a warp sum of 32 values < 65536 each fits 21 bits, so reduce the halves separately
    const unsigned n_over = __reduce_add_sync(0xffffffffu, cg & 0xffffu), n_conf = __reduce_add_sync(0xffffffffu, cg >> 16);
    const unsigned n_ml = __reduce_add_sync(0xffffffffu, cm & 0xffffu), n_wins = __reduce_add_sync(0xffffffffu, cm >> 16);
    const unsigned n_stop = __reduce_add_sync(0xffffffffu, ce & 0xffffu), n_tgt = __reduce_add_sync(0xffffffffu, ce >> 16);
    const unsigned n_inpos = __reduce_add_sync(0xffffffffu, cs & 0xffffu), n_cool = __reduce_add_sync(0xffffffffu, cs >> 16);
    const unsigned n_dir = __reduce_add_sync(0xffffffffu, cd);
    const unsigned n_open = __popc(__ballot_sync(0xffffffffu, open));
    const unsigned n = __popc(__ballot_sync(0xffffffffu, active));
    if (lane == 0 && n) {
        atomicAdd(&S.d[D_SUM_PNL], s1); atomicAdd(&S.d[D_SUM_PNL2], s2);
        atomicAdd(&S.d[D_SUM_PATH2], s3); atomicAdd(&S.d[D_SUM_MTM], s4);
        atomic_max_f64(&S.xmax, mx); atomic_max_f64(&S.xnegmin, mn);
        smem_inc(&S.u[U_NPATHS], n);
        if (n_over) smem_inc(&S.u[U_ENTRY + QMMX_LEVEL_OVERTOUCHED], n_over);
        if (n_conf) smem_inc(&S.u[U_ENTRY + QMMX_CONF_LOW], n_conf);
        if (n_ml) smem_inc(&S.u[U_ENTRY + QMMX_ML_CONF_LOW], n_ml);
        if (n_dir) smem_inc(&S.u[U_ENTRY + QMMX_DIR_UNKNOWN], n_dir);
        if (n_wins) smem_inc(&S.u[U_NWINS], n_wins);
        if (n_stop) smem_inc(&S.u[U_EXIT + QMMX_EXIT_STOP], n_stop);
        if (n_tgt) smem_inc(&S.u[U_EXIT + QMMX_EXIT_TARGET], n_tgt);
        if (n_stop + n_tgt) smem_inc(&S.u[U_NTRADES], n_stop + n_tgt);
        if (n_inpos) smem_inc(&S.u[U_ENTRY + QMMX_IN_POSITION], n_inpos);
        if (n_cool) smem_inc(&S.u[U_ENTRY + QMMX_COOLDOWN], n_cool);
        if (n_open) { smem_inc(&S.u[U_NOPEN], n_open); smem_inc(&S.u[U_EXIT + QMMX_EXIT_OPEN], n_open); }
    }
}

// per-CTA setup: zero the aggregate and build the shared tables from DevTables + this launch's combo
// fine = the class partition (lean / tile kernels): cell bytes, zone list and zone values are class ids
__device__ inline void block_setup(const ScanArgs& A, bool fine = false) {
    BlockShared& S = g_S;
    const DevTables* __restrict__ T = A.tables;
    const double neg_inf = __longlong_as_double((long long)0xfff0000000000000ull);
    for (int t = threadIdx.x; t < U_HIST; t += blockDim.x) S.u[t] = 0ull;
    for (int t = threadIdx.x; t < QMMX_PNL_HIST_BINS; t += blockDim.x) S.hist[t] = 0u;
    if (threadIdx.x < QMMX_AGG_ND) S.d[threadIdx.x] = 0.0;
    if (threadIdx.x == 0) {
        S.xmax = neg_inf; S.xnegmin = neg_inf;
        S.L = T->L; S.Z = fine ? T->FZ : T->Z; S.ncell = T->ncell; S.shift = T->shift; S.obase = T->obase;
    }
    for (int k = threadIdx.x; k < kMaxLevels; k += blockDim.x) {
        const DevLevel D = T->lev[k];
        // M:1669-1674: long stop = L - SP, target = L + TP; short stop = L + SP, target = L - TP  (float64, then
        // rounded toward the inside of the no-exit band so the fp32 compare equals the float64 compare)
        S.thr[0][k] = make_float2(__double2float_rd(D.price - A.combo.sp), __double2float_ru(D.price + A.combo.tp));
        S.thr[1][k] = make_float2(__double2float_rd(D.price - A.combo.tp), __double2float_ru(D.price + A.combo.sp));
#pragma unroll
        for (int b = 0; b < 3; ++b) S.cpass[b][k] = make_float2(D.cpass[b][0], D.cpass[b][1]);
        S.cand_lo[k] = D.cand_lo; S.cand_hi[k] = D.cand_hi;
        ColdLevel Cl; Cl.price = D.price; Cl.wlo = D.wlo; Cl.whi = D.whi; Cl.flags = D.flags; Cl.pad = 0u;
        S.cold[k] = Cl;
    }
    for (int t = threadIdx.x; t < kMaxZones; t += blockDim.x) {
        S.zhi[t] = fine ? T->fzhi[t] : T->zhi[t]; S.zlev[t] = fine ? T->fzcls[t] : T->zlev[t];
    }
    const int n16 = (T->ncell + 15) / 16;
    const uint4* src = reinterpret_cast<const uint4*>(fine ? T->fcell : T->cell);
    uint4* dst = reinterpret_cast<uint4*>(g_cell);
    for (int t = threadIdx.x; t < n16; t += blockDim.x) dst[t] = src[t];
    __syncthreads();
}

// flush the block aggregate into the flat global buffer of this launch's combo; caller syncs before
__device__ inline void block_flush_agg(unsigned long long* gu, int steps) {
    BlockShared& S = g_S;
    double* gd = reinterpret_cast<double*>(gu + QMMX_AGG_NU);
    if (threadIdx.x == 0) {
        // derived counters: n_steps, OK entries (every opened position ends as a trade or as OPEN), and the
        // remainder code (TOO_FAR, or NOLEVELS without levels)
        const unsigned long long nsteps = S.u[U_NPATHS] * (unsigned long long)steps;
        S.u[U_ENTRY + QMMX_OK_ENTRY] = S.u[U_NTRADES] + S.u[U_NOPEN];
        if (S.L == 0) S.u[U_ENTRY + QMMX_DIR_UNKNOWN] = 0ull;      // NOLEVELS precedes DIR_UNKNOWN (M:1524)
        unsigned long long others = 0;
        for (int c = 0; c < QMMX_N_ENTRY_CODES; ++c) others += S.u[U_ENTRY + c];
        S.u[U_NSTEPS] = nsteps;
        S.u[U_ENTRY + (S.L == 0 ? QMMX_NOLEVELS : QMMX_TOO_FAR)] += nsteps - others;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < QMMX_AGG_NU; t += blockDim.x) {
        const unsigned long long v = t < U_HIST ? S.u[t] : (unsigned long long)S.hist[t - U_HIST];
        if (v) atomicAdd(&gu[t], v);
    }
    if (threadIdx.x < QMMX_AGG_ND) { const double v = S.d[threadIdx.x]; if (v != 0.0) atomicAdd(&gd[threadIdx.x], v); }
    if (threadIdx.x == 0 && S.u[U_NPATHS]) {
        atomic_max_f64(&gd[QMMX_AGG_ND + 0], S.xmax);
        atomic_max_f64(&gd[QMMX_AGG_ND + 1], S.xnegmin);
    }
}

}  // namespace qmmx
