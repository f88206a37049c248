// Lean lockstep engine scan (the production scan kernel; engine.cuh's PathEngine stays as the TINY_PRICES path).
//
// Same semantics as PathEngine (qmmx_monolithic.py:1492-1771 gates, M:2966-3014 exits),
// restructured around what ncu showed for the first kernel (profiles/r1b_scan_close_summary.txt): a warp takes
// every branch of the state machine on almost every step, so the price of a step is the SUM of the branches and
// the only lever is to make each branch short.
//
//  * flat / in-position / cooldown are not branches any more: a flat path carries exit thresholds (-inf, +inf), a
//    path in a position carries cool = INT_MAX, so "exit test" and "may enter" are two compares on every step.
//  * span codes are not counted per step: IN_POSITION steps = exit bar - entry bar and COOLDOWN steps =
//    min(cool, steps) - exit bar - 1 are added at the exit; the DIR_UNKNOWN prefix (steps before the first price
//    change) is found by a short prologue and becomes the initial `cool`.
//  * direction is the sign of the last non-zero fp32 difference (exact: fp32 subtraction of distinct values
//    is never 0 without flush-to-zero), one FADD + FSETP + FSEL.
//  * the three gates LEVEL_OVERTOUCHED (M:1579), CONF_LOW (M:1628) and ML_CONF_LOW (M:1745-1755) are ONE
//    shared-memory float2 per (level, side, touch count): pass <=> x <= p <= |y|; the sign of y and the
//    in-range bit select the reason code (tables built per CTA from the exact host intervals).
//  * touch counters are one byte per (level, thread) in shared memory, rows rotated over the banks.
#pragma once
#include "kernels.cuh"

namespace qmmx {

constexpr int kLeanThreads = 512;
constexpr int kLeanBlocksPerSM = 2;
constexpr int kLeanWarps = kLeanThreads / 32;
constexpr int kTouchStride = kLeanThreads + 4;      // bytes per level row; stride/4 is odd, so level rows rotate over the banks
constexpr int kNoCool = 0x7fffffff;

__host__ __device__ inline int lean_lpad(int L) { return L > 32 ? 64 : 32; }
__host__ __device__ inline size_t lean_smem_bytes(int cell_bytes, int L) {
    return (size_t)cell_bytes + (size_t)lean_lpad(L) * 2 * 16 * sizeof(float4) + (size_t)lean_lpad(L) * kTouchStride;
}

struct LeanTables {
    const float4* gate;      // [level][side][tc]: {lo, +-hi, exit-low threshold, exit-high threshold}
    uint8_t* touch;          // [level][kTouchStride]
};

// per-CTA: gate[(k*2+side)*16+tc] from the exact confidence intervals, the overtouch limit and (single day, AND
// mode) the host-evaluated logistic mask.  Encoding: in range & y>0 -> OK; in range & y<0 -> ML_CONF_LOW;
// out of range & y = -inf -> LEVEL_OVERTOUCHED (x = +inf: never in range); out of range otherwise -> CONF_LOW.  z, w = the exit thresholds an entry at
// this (level, side) arms (M:1669-1674; float64 stop/target rounded toward the inside of the band, as block_setup).
template <bool MULTIDAY>
__device__ inline LeanTables lean_setup(const ScanArgs& A) {
    float4* gate = reinterpret_cast<float4*>(g_cell + A.cell_bytes);
    const int lpad = lean_lpad(A.n_levels);
    uint8_t* touch = reinterpret_cast<uint8_t*>(gate + lpad * 32);
    const DevTables* __restrict__ T = A.tables;
    const float inf = __int_as_float(0x7f800000);
    const bool fold_ml = !MULTIDAY && A.mlmask && !(A.combo.flags & QMMX_F_USE_BLEND);
    for (int t = threadIdx.x; t < lpad * 32; t += blockDim.x) {
        const int k = t >> 5, side = (t >> 4) & 1, tc = t & 15;
        float4 g = make_float4(inf, inf, -inf, inf);            // never in range, CONF_LOW
        if (k < A.n_levels) {
            const int b = (tc < 1 ? 1 : (tc > 3 ? 3 : tc)) - 1;
            const float lo = T->lev[k].cpass[b][0], hi = T->lev[k].cpass[b][1];
            if (lo <= hi) { g.x = lo; g.y = hi; }
            if (tc >= A.combo.overtouch) { g.x = inf; g.y = -inf; }
            else if (fold_ml && lo <= hi && ((A.mlmask[k] >> (side * 16 + tc)) & 1u)) g.y = -hi;
            const double lp = T->lev[k].price;
            g.z = __double2float_rd(side == 0 ? lp - A.combo.sp : lp - A.combo.tp);
            g.w = __double2float_ru(side == 0 ? lp + A.combo.tp : lp + A.combo.sp);
        }
        gate[t] = g;
    }
    LeanTables S; S.gate = gate; S.touch = touch;
    return S;
}

// the sign encoding needs every finite upper bound of a confidence interval to be positive (prices are)
// -- checked on the host before this kernel is chosen (qmmx_mc.cu lean_ok).

// SLOW = the blend gate or the OnlinePolicy gate is enabled (float64 slow paths compiled in)
template <bool MULTIDAY, bool HL, bool PARITY, bool WIDE, bool SLOW>
struct LeanEngine {
    typedef typename LatchType<WIDE>::T latch_t;
    float prev, d;         // last price; last non-zero price change (0 = direction unknown, M:1529-1538)
    int cool;              // first step at which an entry may be evaluated; kNoCool while in a position
    float lo_thr, hi_thr;  // exit when low <= lo_thr or high >= hi_thr; (-inf, +inf) when flat
    float entry;
    int pos, entry_bar;    // level | side << 8
    latch_t latch;
    uint32_t c_gate, c_ml, c_exit, c_span, c_dir;    // packed 16-bit tallies, as PathEngine
    double pnl, pnl2, mtm;
    int ntr; int64_t path;
    uint8_t* tcol;         // this thread's column of touch bytes

    __device__ __forceinline__ void reset(const LeanTables& S, int lpad, int64_t path_) {
        path = path_;
        tcol = S.touch + threadIdx.x;
        prev = 0.f; d = 0.f; cool = 0; pos = -1; entry = 0.f; entry_bar = 0;
        lo_thr = __int_as_float(0xff800000); hi_thr = __int_as_float(0x7f800000);
        latch = 0; c_gate = c_ml = c_exit = c_span = c_dir = 0u;
        pnl = 0.0; pnl2 = 0.0; mtm = 0.0; ntr = 0;
        for (int k = 0; k < lpad; ++k) tcol[k * kTouchStride] = 0;
    }

    // steps 0 .. j0-1 carry no direction (DIR_UNKNOWN, M:1538): j0 = index of the first close that differs from close[0]
    __device__ __forceinline__ void prologue(const ScanArgs& A, float p0, int j0) {
        prev = p0; cool = j0; c_dir = (uint32_t)j0;
    }

    __device__ __forceinline__ void do_exit(const ScanArgs& A, float pf, bool hit_lo, bool hit_hi, int i) {
        const int side = pos >> 8;
        const bool is_stop = (side == 0) ? hit_lo : hit_hi;      // stop is tested first (M:2975-3010, M:3621-3636)
        double tp;
        float exit_f = pf;
        if (!HL) {
            const double diff = (double)pf - (double)entry;      // fp32 values, float64 difference (M:1926)
            tp = side == 0 ? diff : -diff;
        } else {
            const double lp = g_S.cold[pos & 0xff].price;
            double exit_px;
            if (side == 0) exit_px = is_stop ? lp - A.combo.sp : lp + A.combo.tp;
            else exit_px = is_stop ? lp + A.combo.sp : lp - A.combo.tp;
            const double e = (double)entry;
            tp = (side == 0) ? (exit_px - e) : (e - exit_px);
            exit_f = (float)exit_px;
        }
        pnl += tp; pnl2 += tp * tp;
        cool = i + A.combo.cool_steps;                            // M:2982, M:269-270
        const int ncool = max(min(cool, A.steps) - i - 1, 0);     // steps coded COOLDOWN (M:1516)
        c_span += (uint32_t)(i - entry_bar) + ((uint32_t)ncool << 16);
        if (tp > 0) c_ml += 1u << 16;                             // label, M:1936
        c_exit += is_stop ? 1u : (1u << 16);
        if (PARITY) record_trade(&A, path, ntr, entry_bar, i, side, is_stop ? QMMX_EXIT_STOP : QMMX_EXIT_TARGET, pos & 0xff, entry, exit_f, (float)tp);
        ++ntr;
        pos = -1;
        lo_thr = __int_as_float(0xff800000); hi_thr = __int_as_float(0x7f800000);
    }

    __device__ __forceinline__ int contact(const ScanArgs& A, const LeanTables& S, int k, float pf, int i, int day_off) {
        const int side = d < 0.f ? 1 : 0;                         // M:1645: up -> long(0), down -> short(1)
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
        uint8_t* tp = tcol + k * kTouchStride;
        int tc = *tp;
        if (isnew) { tc = min(tc + 1, 15); *tp = (uint8_t)tc; }
        const float4 g = S.gate[(k * 2 + side) * 16 + tc];
        const bool neg = __float_as_int(g.y) < 0;
        if (!(pf >= g.x && pf <= fabsf(g.y))) {
            const bool over = __float_as_int(g.y) == (int)0xff800000;   // y = -inf marks tc >= overtouch limit
            c_gate += over ? 1u : (1u << 16);                     // M:1579 / M:1628
            return over ? QMMX_LEVEL_OVERTOUCHED : QMMX_CONF_LOW;
        }
        if (neg) { c_ml += 1u; return QMMX_ML_CONF_LOW; }         // M:1745-1755 (single day, AND mode)
        if (SLOW && (A.combo.flags & QMMX_F_USE_BLEND)) {         // M:1730-1742
            if (blend_fails(&A, k, pf, tc, side, day_off)) return QMMX_COMBINED_LOW;
        } else if (MULTIDAY && A.mlmask) {
            const uint32_t m = __ldg(&A.mlmask[day_off + k]);
            if ((m >> (side * 16 + tc)) & 1u) { c_ml += 1u; return QMMX_ML_CONF_LOW; }
        }
        if (SLOW && !(A.combo.flags & QMMX_F_DISABLE_POLICY_GATE)) {   // OnlinePolicy gate of the live loop (M:3046-3109)
            if (!policy_passes(&A, k, pf, side, i)) return QMMX_ONLINE_POLICY;
        }
        lo_thr = g.z; hi_thr = g.w;                               // M:1669-1674
        pos = k | (side << 8);
        entry = pf; entry_bar = i;
        cool = kNoCool;
        return QMMX_OK_ENTRY;
    }

    __device__ __forceinline__ void step(const ScanArgs& A, const LeanTables& S, float pf, float hh, float ll, int i, int day_off) {
        const bool elig = i >= cool;
        int code = QMMX_TOO_FAR;
        if (PARITY) code = (cool == kNoCool) ? QMMX_IN_POSITION : (elig ? QMMX_TOO_FAR : QMMX_COOLDOWN);
        const float xl = HL ? ll : pf, xh = HL ? hh : pf;
        const bool hit_lo = xl <= lo_thr, hit_hi = xh >= hi_thr;
        if (hit_lo | hit_hi) do_exit(A, pf, hit_lo, hit_hi, i);
        const float diff = pf - prev;
        if (diff != 0.f) d = diff;
        prev = pf;
        if (PARITY && d == 0.f) code = QMMX_DIR_UNKNOWN;
        // nearest level within CONTACT_PROX (M:1543-1545): the cell byte is read by every lane, the branch is taken by few
        const uint32_t c = min((uint32_t)(__float_as_int(pf) - A.cell_obase) >> A.cell_shift, (uint32_t)(A.cell_ncell - 1));
        uint32_t v = g_cell[c];
        if (elig && v != kFar) {
            if (v == kMixed) v = classify_zones(pf);
            if (v != kFar) code = contact(A, S, (int)v, pf, i, day_off);
        }
        if (PARITY && A.codes) {
            if (A.n_levels == 0 && (code == QMMX_TOO_FAR || code == QMMX_DIR_UNKNOWN)) code = QMMX_NOLEVELS;
            A.codes[(int64_t)i * A.paths + path] = (uint8_t)code;
        }
    }

    __device__ __forceinline__ void finish(const ScanArgs& A) {
        if (pos >= 0) {
            const int side = pos >> 8;
            const double e = (double)entry, last = (double)prev;
            mtm = (side == 0) ? (last - e) : (e - last);
            c_span += (uint32_t)(A.steps - 1 - entry_bar);
            if (PARITY) record_trade(&A, path, ntr, entry_bar, -1, side, QMMX_EXIT_OPEN, pos & 0xff, entry, prev, 0.0f);
            ++ntr;
        }
        if (PARITY && A.tcounts) A.tcounts[path] = ntr;
        double fb = floor((pnl - A.hist_lo) * A.hist_inv_w);
        const int bin = fb < 0 ? 0 : (fb > QMMX_PNL_HIST_BINS - 1 ? QMMX_PNL_HIST_BINS - 1 : (int)fb);
        atomicAdd(&g_S.hist[bin], 1u);
    }
};

// Bars are [step][pitch]: one coalesced 128-byte line per warp, step and field; 2U rows are kept in flight per
// thread in registers (register double buffer) -- at 1024 threads per SM a warp returns to the loop every ~0.3 us,
// HBM latency is ~0.8 us.
template <bool MULTIDAY, bool HL, bool PARITY, bool WIDE, bool SLOW, int U>
__device__ __forceinline__ void lean_scan_path(LeanEngine<MULTIDAY, HL, PARITY, WIDE, SLOW>& E, const ScanArgs& A,
                                               const LeanTables& S, int64_t path) {
    const size_t pitch = (size_t)A.pitch;
    const float* pc = A.close + path;
    const float* ph = HL ? A.high + path : nullptr;
    const float* pl = HL ? A.low + path : nullptr;
    const int steps = A.steps;
    {   // DIR_UNKNOWN prefix
        const float p0 = __ldg(pc);
        int j0 = 1;
        while (j0 < steps && __ldg(pc + (size_t)j0 * pitch) == p0) ++j0;
        E.prologue(A, p0, j0);
    }
    float c0[U], c1[U], h0[U], h1[U], l0[U], l1[U];
    const int nfull = steps / (2 * U);
    DayTracker<MULTIDAY> day;
    day.init(A);
    int i = 0;
    if (nfull > 0) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            c0[u] = __ldg(pc); pc += pitch;
            if (HL) { h0[u] = __ldg(ph); l0[u] = __ldg(pl); ph += pitch; pl += pitch; }
        }
#pragma unroll 1
        for (int n = 0; n < nfull; ++n, i += 2 * U) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                c1[u] = __ldg(pc); pc += pitch;
                if (HL) { h1[u] = __ldg(ph); l1[u] = __ldg(pl); ph += pitch; pl += pitch; }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) E.step(A, S, c0[u], HL ? h0[u] : 0.f, HL ? l0[u] : 0.f, i + u, day.at(A, i + u));
            if (n + 1 < nfull) {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    c0[u] = __ldg(pc); pc += pitch;
                    if (HL) { h0[u] = __ldg(ph); l0[u] = __ldg(pl); ph += pitch; pl += pitch; }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) E.step(A, S, c1[u], HL ? h1[u] : 0.f, HL ? l1[u] : 0.f, i + U + u, day.at(A, i + U + u));
        }
    }
#pragma unroll 1
    for (; i < steps; ++i) {
        const float c = __ldg(pc); pc += pitch;
        float h = 0.f, l = 0.f;
        if (HL) { h = __ldg(ph); l = __ldg(pl); ph += pitch; pl += pitch; }
        E.step(A, S, c, h, l, i, day.at(A, i));
    }
}

template <bool MULTIDAY, bool HL, bool PARITY, bool WIDE, bool SLOW>
__global__ void __launch_bounds__(kLeanThreads, kLeanBlocksPerSM) lean_scan_kernel(const __grid_constant__ ScanArgs A) {
    block_setup(A);
    const LeanTables S = lean_setup<MULTIDAY>(A);
    __syncthreads();
    LeanEngine<MULTIDAY, HL, PARITY, WIDE, SLOW> E;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lpad = lean_lpad(A.n_levels);
    const int64_t ntiles = (A.paths + 31) / 32;
    for (int64_t tile = (int64_t)blockIdx.x * kLeanWarps + warp; tile < ntiles; tile += (int64_t)gridDim.x * kLeanWarps) {
        const int64_t path = tile * 32 + lane;
        const bool active = path < A.paths;
        E.reset(S, lpad, path);
        if (active) { lean_scan_path<MULTIDAY, HL, PARITY, WIDE, SLOW, 2>(E, A, S, path); E.finish(A); }
        __syncwarp();
        accumulate_path(active, E);
    }
    __syncthreads();
    block_flush_agg(A.flat, A.steps);
}

}  // namespace qmmx
