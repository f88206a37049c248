// Device-side QMMX engine: one thread walks one price path through the entry gates
// (qmmx_monolithic.py:1492-1771), the live loop's exit block (M:2966-3014) and the OnlinePolicy gate
// (M:3046-3109).  Semantics frozen in SURVEY.md Appendix A; checked bit-for-bit against oracle/.
//
// Arithmetic contract: prices are fp32 in memory and are evaluated in float64 exactly like CPython
// wherever a decision is taken on the contact path.  The file is compiled with --fmad=false so no
// multiply-add is ever contracted; the per-step fast path replaces the O(L) float64 nearest-level search
// by the exact host-built classification tables (tables.h) and float64 threshold compares by
// equivalent fp32 compares against directed-rounded thresholds.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "qmmx_mc.h"
#include "tables.h"

namespace qmmx {

// ---- shared-memory block state -------------------------------------------------------------------------
struct BlockShared {
    // tables (subset of DevTables the kernels need)
    int32_t L, Z, ncell, pad0;
    float glo, inv_w;
    double price[kMaxLevels];
    uint8_t solid[kMaxLevels];
    uint8_t confl06[kMaxLevels];
    float zhi[kMaxZones];
    uint8_t zlev[kMaxZones];
    DevCombo combo;
    // block aggregate, flushed to the global reduction buffer with atomics
    unsigned long long u[QMMX_AGG_NU];
    double d[QMMX_AGG_ND];
    double xmax, xnegmin;
    // per-thread touch counters: 4-bit saturating, 64 levels -> 8 words per thread, [word][thread]
    uint32_t touch[8 * kThreads];
    // grid cells follow (dynamic)
    uint8_t cell[1];
};

constexpr int U_NPATHS = 0, U_NSTEPS = 1, U_NTRADES = 2, U_NWINS = 3, U_NOPEN = 4, U_NOVERFLOW = 5;
constexpr int U_ENTRY = 6, U_EXIT = U_ENTRY + QMMX_N_ENTRY_CODES, U_HIST = U_EXIT + QMMX_N_EXIT_CODES;
constexpr int D_SUM_PNL = 0, D_SUM_PNL2 = 1, D_SUM_PATH2 = 2, D_SUM_MTM = 3;

struct EngineGlobals {            // read-only, global memory
    const double* lr;  int n_lr;      // [n_lr][5]
    const double* pol; int n_pol;     // [n_pol][3][7]
    double hist_lo, hist_inv_w;
    int32_t steps;
    // parity-mode outputs (may be null)
    qmmx_trade* trades; int32_t* tcounts; int32_t max_trades; uint8_t* codes;
    int64_t paths;                    // row stride of codes
};

__device__ __forceinline__ void smem_inc(unsigned long long* p, unsigned long long v = 1ull) { atomicAdd(p, v); }

// OnlinePolicy._dot under CPython >= 3.12 (Neumaier-compensated builtin sum), M:334-335
__device__ inline double dot_py312(const double* __restrict__ w, const double* x) {
    double f = 0.0 + w[0] * x[0];
    double c = 0.0;
#pragma unroll
    for (int i = 1; i < QMMX_POLICY_DIM; ++i) {
        const double v = w[i] * x[i];
        const double t = f + v;
        if (fabs(f) >= fabs(v)) c += (f - t) + v; else c += (v - t) + f;
        f = t;
    }
    if (c != 0.0 && isfinite(c)) f += c;
    return f;
}

__device__ inline double sigmoid_policy(double x) {   // M:20-26
    if (x < -50) return 0.0;
    if (x > 50) return 1.0;
    return 1.0 / (1.0 + exp(-x));
}

__device__ inline double compute_confidence(double dist, bool solid, int tc, double pmax) {   // M:1415-1427
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

// ---- per-path state machine ---------------------------------------------------------------------------
template <bool HL, bool PARITY, bool TINY>
struct PathEngine {
    BlockShared& S;
    const EngineGlobals& G;
    const int tid;
    int64_t path;          // global index into the resident bars (parity outputs)
    // EngineState (M:259-270) + contact latch (M:1376)
    float prev;            // NaN until the first tick
    int last_dir;          // 0 unknown, 1 up, 2 down
    int cool_until;        // first step index at which the cooldown is over
    int pos;               // -1 flat, else level | side << 8
    float lo_thr, hi_thr;  // exit when low <= lo_thr or high >= hi_thr (directed-rounded stop/target)
    float entry;
    int entry_bar;
    uint64_t latch;
    // per-path accumulators
    double pnl, pnl2, mtm;
    int ntr;

    __device__ PathEngine(BlockShared& s, const EngineGlobals& g, int t) : S(s), G(g), tid(t) {}

    __device__ __forceinline__ void reset(int64_t path_) {
        path = path_;
        prev = __int_as_float(0x7fc00000);
        last_dir = 0; cool_until = 0; pos = -1; lo_thr = 0.f; hi_thr = 0.f; entry = 0.f; entry_bar = 0;
        latch = 0ull; pnl = 0.0; pnl2 = 0.0; mtm = 0.0; ntr = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) S.touch[w * kThreads + tid] = 0u;
    }

    __device__ __forceinline__ int touch_get(int k) const {
        return (S.touch[(k >> 3) * kThreads + tid] >> ((k & 7) * 4)) & 15;
    }
    __device__ __forceinline__ int touch_inc(int k) {
        uint32_t& w = S.touch[(k >> 3) * kThreads + tid];
        const int sh = (k & 7) * 4;
        int v = (w >> sh) & 15;
        if (v < 15) { ++v; w = (w & ~(15u << sh)) | ((uint32_t)v << sh); }
        return v;
    }

    __device__ __forceinline__ void record(int entry_bar_, int exit_bar, int side, int reason, int k,
                                           float entry_, float exit_, float pnl_) {
        if (PARITY && G.trades) {
            if (ntr < G.max_trades) {
                qmmx_trade t;
                t.entry_bar = entry_bar_; t.exit_bar = exit_bar; t.side = (uint8_t)side; t.reason = (uint8_t)reason;
                t.level_id = (uint16_t)k; t.entry = entry_; t.exit = exit_; t.pnl = pnl_;
                G.trades[path * G.max_trades + ntr] = t;
            } else smem_inc(&S.u[U_NOVERFLOW]);
        }
        ++ntr;
    }

    // contact path: price is within CONTACT_PROX of level k (rare: a few % of steps)
    __device__ __forceinline__ int contact(int k, float pf, int d, int i) {
        const DevCombo& C = S.combo;
        const double p = (double)pf;
        const double lp = S.price[k];
        const double dist = fabs(lp - p);
        const uint64_t bit = 1ull << k;
        int tc;
        if (!(latch & bit)) { latch |= bit; tc = touch_inc(k); }           // M:1562-1566
        else tc = touch_get(k);
        uint64_t others = latch & ~bit;                                     // M:1572-1576
        while (others) {
            const int j = __ffsll((long long)others) - 1;
            others &= others - 1;
            if (fabs(S.price[j] - p) > C.prox) latch &= ~(1ull << j);
        }
        if (tc >= C.overtouch) return QMMX_LEVEL_OVERTOUCHED;                // M:1579
        const double conf = compute_confidence(dist, S.solid[k] != 0, tc, C.pmax);
        if (conf < C.qmin) return QMMX_CONF_LOW;                             // M:1628
        const int side = (d == 1) ? 0 : 1;                                   // M:1645
        double stop, target;                                                 // M:1669-1674
        if (side == 0) { stop = lp - C.sp; target = lp + C.tp; }
        else { stop = lp + C.sp; target = lp - C.tp; }
        // M:1677-1705 soft veto: volume slope is identically 0 without per-bar volume => never vetoes
        const int day = i / C.spd;
        double mlp = conf;
        bool ok_ml = true;
        if (!(C.flags & QMMX_F_DISABLE_ML_GATE) && G.n_lr > 0) {             // M:1454-1466
            const double* m = G.lr + QMMX_LR_DIM * (day < G.n_lr ? day : G.n_lr - 1);
            const double distf = fabs(lp - stop);
            double z = m[4];
            z = z + m[0] * 1.0;
            z = z + m[1] * distf;
            z = z + m[2] * (double)tc;
            z = z + m[3] * (d == 1 ? 1.0 : 0.0);
            const double prob = 1.0 / (1.0 + exp(-z));
            ok_ml = prob >= C.qmin;
            mlp = prob;
        }
        if (C.flags & QMMX_F_USE_BLEND) {                                    // M:1730-1742
            const double blended = C.wr * conf + C.wm * mlp;
            if (blended < C.qmin) return QMMX_COMBINED_LOW;
        } else if (!(C.flags & QMMX_F_DISABLE_ML_GATE) && !ok_ml) return QMMX_ML_CONF_LOW;   // M:1745-1755
        // evaluate_entry returned OK; OnlinePolicy gate of the live loop (M:3046-3109)
        if (!(C.flags & QMMX_F_DISABLE_POLICY_GATE)) {
            double x[QMMX_POLICY_DIM];
            const double pa = fabs(p - lp);
            const double tod = (double)(i % C.spd) / 390.0;
            x[0] = 1.0; x[1] = pa < 1.0 ? pa : 1.0; x[2] = 0.0;
            x[3] = side == 1 ? 1.0 : 0.0; x[4] = side == 0 ? 1.0 : 0.0;
            x[5] = S.confl06[k] ? 1.0 : 0.0; x[6] = tod < 1.0 ? tod : 1.0;
            double s_go = 0.5, s_skip = 0.5;                                 // zero weights => sigmoid(0)
            if (G.n_pol > 0) {
                const double* w = G.pol + QMMX_POLICY_HEADS * QMMX_POLICY_DIM * (day < G.n_pol ? day : G.n_pol - 1);
                s_go = sigmoid_policy(dot_py312(w + QMMX_POLICY_DIM * side, x));
                s_skip = sigmoid_policy(dot_py312(w + QMMX_POLICY_DIM * 2, x));
            }
            if (!((s_go >= 0.60) && (s_skip < 0.55))) return QMMX_ONLINE_POLICY;
        }
        // open_trade (M:3112-3115, M:1888-1897): entry at the tick price; first exit test on the next tick
        pos = k | (side << 8);
        entry = pf; entry_bar = i;
        if (side == 0) { lo_thr = __double2float_rd(stop); hi_thr = __double2float_ru(target); }
        else { hi_thr = __double2float_ru(stop); lo_thr = __double2float_rd(target); }
        return QMMX_OK_ENTRY;
    }

    __device__ __forceinline__ void do_exit(float pf, bool hit_lo, bool hit_hi, int i) {
        const DevCombo& C = S.combo;
        const int k = pos & 0xff, side = pos >> 8;
        const int reason = (side == 0) ? (hit_lo ? QMMX_EXIT_STOP : QMMX_EXIT_TARGET)
                                       : (hit_hi ? QMMX_EXIT_STOP : QMMX_EXIT_TARGET);
        double exit_px = (double)pf;
        if (HL) {   // exit at the stop/target level itself (M:3636)
            const double lp = S.price[k];
            if (side == 0) exit_px = (reason == QMMX_EXIT_STOP) ? lp - C.sp : lp + C.tp;
            else exit_px = (reason == QMMX_EXIT_STOP) ? lp + C.sp : lp - C.tp;
        }
        const double e = (double)entry;
        const double tp = (side == 0) ? (exit_px - e) : (e - exit_px);       // M:1926
        pnl += tp; pnl2 += tp * tp;
        cool_until = i + C.cool_steps;                                       // M:2982, M:269-270
        smem_inc(&S.u[U_NTRADES]);
        if (tp > 0) smem_inc(&S.u[U_NWINS]);                                 // label, M:1936
        smem_inc(&S.u[U_EXIT + reason]);
        smem_inc(&S.u[U_ENTRY + QMMX_IN_POSITION], (unsigned long long)(i - entry_bar));
        int cd = C.cool_steps < G.steps - i ? C.cool_steps : G.steps - i;
        cd -= 1;
        if (cd > 0) smem_inc(&S.u[U_ENTRY + QMMX_COOLDOWN], (unsigned long long)cd);
        record(entry_bar, i, side, reason, k, entry, (float)exit_px, (float)tp);
        pos = -1;
    }

    // one tick: price pf = bar i's close (hh/ll = its high/low when HL)
    __device__ __forceinline__ void step(float pf, float hh, float ll, int i) {
        int code = QMMX_TOO_FAR;
        // direction (M:1529-1537) and last_direction update (M:2954-2955)
        int d;
        if (!TINY) {
            // |p|,|prev| >= 2^-5 (guaranteed by the host): fp32 spacing > 2e-9, so the +-1e-9 guard band of
            // the reference cannot change any comparison and d doubles as the updated last_direction
            d = (pf > prev) ? 1 : ((pf < prev) ? 2 : last_dir);
        } else {
            d = 0;
            if (prev == prev) {
                const double p = (double)pf, q = (double)prev;
                if (p > q + 1e-9) d = 1; else if (p < q - 1e-9) d = 2; else d = last_dir;
            }
        }
        if (pos >= 0) {
            code = QMMX_IN_POSITION;                                         // M:1508
            const bool hit_lo = (HL ? ll : pf) <= lo_thr;
            const bool hit_hi = (HL ? hh : pf) >= hi_thr;
            if (hit_lo | hit_hi) do_exit(pf, hit_lo, hit_hi, i);
        } else if (i < cool_until) {
            code = QMMX_COOLDOWN;                                            // M:1516
        } else if (S.L == 0) {
            code = QMMX_NOLEVELS;                                            // M:1524 (tallied as remainder)
        } else if (d == 0) {
            code = QMMX_DIR_UNKNOWN;                                         // M:1538
            smem_inc(&S.u[U_ENTRY + QMMX_DIR_UNKNOWN]);
        } else {
            uint32_t v = S.cell[cell_of(pf, S.glo, S.inv_w, S.ncell)];
            if (v == kMixed) {
                int a = 0, b = S.Z - 1;
                while (a < b) {
                    const int m = (a + b) >> 1;
                    if (pf <= S.zhi[m]) b = m; else a = m + 1;
                }
                v = S.zlev[a];
            }
            if (v != kFar) {                                                 // else TOO_FAR, M:1545
                code = contact((int)v, pf, d, i);
                smem_inc(&S.u[U_ENTRY + code]);
            }
        }
        if (!TINY) last_dir = d;
        else if (prev == prev && pf != prev) last_dir = (pf > prev) ? 1 : 2;
        prev = pf;
        if (PARITY && G.codes) G.codes[(int64_t)i * G.paths + path] = (uint8_t)code;
    }

    // end of path: open position => OPEN, realised 0, mark-to-market reported separately
    __device__ __forceinline__ void finish() {
        if (pos >= 0) {
            const int k = pos & 0xff, side = pos >> 8;
            const double e = (double)entry, last = (double)prev;
            mtm = (side == 0) ? (last - e) : (e - last);
            smem_inc(&S.u[U_NOPEN]);
            smem_inc(&S.u[U_EXIT + QMMX_EXIT_OPEN]);
            smem_inc(&S.u[U_ENTRY + QMMX_IN_POSITION], (unsigned long long)(G.steps - 1 - entry_bar));
            record(entry_bar, -1, side, QMMX_EXIT_OPEN, k, entry, prev, 0.0f);
        }
        if (PARITY && G.tcounts) G.tcounts[path] = ntr;
        double fb = floor((pnl - G.hist_lo) * G.hist_inv_w);
        const int bin = fb < 0 ? 0 : (fb > QMMX_PNL_HIST_BINS - 1 ? QMMX_PNL_HIST_BINS - 1 : (int)fb);
        smem_inc(&S.u[U_HIST + bin]);
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

__device__ __forceinline__ void smem_max(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (__longlong_as_double((long long)old) < v) {
        const unsigned long long assumed = old;
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
        if (old == assumed) break;
    }
}

__device__ __forceinline__ void global_max(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (__longlong_as_double((long long)old) < v) {
        const unsigned long long assumed = old;
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
        if (old == assumed) break;
    }
}

// add one finished path (active lanes only contribute non-neutral values) to the block aggregate
__device__ __forceinline__ void block_accumulate(BlockShared& S, bool active, double pnl, double pnl2, double mtm) {
    const int lane = threadIdx.x & 31;
    const double neg_inf = __longlong_as_double((long long)0xfff0000000000000ull);
    const double s1 = warp_sum(active ? pnl : 0.0);
    const double s2 = warp_sum(active ? pnl2 : 0.0);
    const double s3 = warp_sum(active ? pnl * pnl : 0.0);
    const double s4 = warp_sum(active ? mtm : 0.0);
    const double mx = warp_max(active ? pnl : neg_inf);
    const double mn = warp_max(active ? -pnl : neg_inf);
    const unsigned n = __popc(__ballot_sync(0xffffffffu, active));
    if (lane == 0 && n) {
        atomicAdd(&S.d[D_SUM_PNL], s1); atomicAdd(&S.d[D_SUM_PNL2], s2);
        atomicAdd(&S.d[D_SUM_PATH2], s3); atomicAdd(&S.d[D_SUM_MTM], s4);
        smem_max(&S.xmax, mx); smem_max(&S.xnegmin, mn);
        smem_inc(&S.u[U_NPATHS], n);
    }
}

__device__ inline void block_zero_agg(BlockShared& S) {
    const double neg_inf = __longlong_as_double((long long)0xfff0000000000000ull);
    for (int t = threadIdx.x; t < QMMX_AGG_NU; t += blockDim.x) S.u[t] = 0ull;
    if (threadIdx.x < QMMX_AGG_ND) S.d[threadIdx.x] = 0.0;
    if (threadIdx.x == 0) { S.xmax = neg_inf; S.xnegmin = neg_inf; }
}

// flush the block aggregate into the flat global buffer of `combo`; caller syncs before and after
__device__ inline void block_flush_agg(BlockShared& S, unsigned long long* flat, int combo, int steps) {
    unsigned long long* gu = flat + (size_t)combo * QMMX_AGG_SLOTS;
    double* gd = reinterpret_cast<double*>(gu + QMMX_AGG_NU);
    if (threadIdx.x == 0) {
        // derived counters: n_steps, and the remainder code (TOO_FAR, or NOLEVELS without levels)
        const unsigned long long nsteps = S.u[U_NPATHS] * (unsigned long long)steps;
        unsigned long long others = 0;
        for (int c = 0; c < QMMX_N_ENTRY_CODES; ++c) others += S.u[U_ENTRY + c];
        S.u[U_NSTEPS] = nsteps;
        S.u[U_ENTRY + (S.L == 0 ? QMMX_NOLEVELS : QMMX_TOO_FAR)] += nsteps - others;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < QMMX_AGG_NU; t += blockDim.x) {
        const unsigned long long v = S.u[t];
        if (v) atomicAdd(&gu[t], v);
    }
    if (threadIdx.x < QMMX_AGG_ND) { const double v = S.d[threadIdx.x]; if (v != 0.0) atomicAdd(&gd[threadIdx.x], v); }
    if (threadIdx.x == 0 && S.u[U_NPATHS]) {
        global_max(&gd[QMMX_AGG_ND + 0], S.xmax);
        global_max(&gd[QMMX_AGG_ND + 1], S.xnegmin);
    }
}

__device__ inline void block_load_tables(BlockShared& S, const DevTables* T) {
    if (threadIdx.x == 0) {
        S.L = T->L; S.Z = T->Z; S.ncell = T->ncell; S.glo = T->glo; S.inv_w = T->inv_w;
    }
    for (int t = threadIdx.x; t < kMaxLevels; t += blockDim.x) {
        S.price[t] = T->price[t]; S.solid[t] = T->solid[t]; S.confl06[t] = T->confl06[t];
    }
    for (int t = threadIdx.x; t < kMaxZones; t += blockDim.x) { S.zhi[t] = T->zhi[t]; S.zlev[t] = T->zlev[t]; }
    const int n4 = (T->ncell + 3) / 4;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(T->cell);
    uint32_t* dst = reinterpret_cast<uint32_t*>(S.cell);
    for (int t = threadIdx.x; t < n4; t += blockDim.x) dst[t] = src[t];
}

}  // namespace qmmx
