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
//
// Kernels: lean_scan_kernel (resident bars; LEARN variants train the per-path OnlinePolicy heads), lean_fused_gbm_kernel
// and lean_fused_boot_kernel (bars generated and consumed in registers), lean_sweep_kernel (a parameter sweep: the
// 32 lanes of a warp are 32 combos of ONE path, steps are classified 32 at a time and far steps are skipped).
#pragma once
#include "kernels.cuh"

namespace qmmx {

#ifndef QMMX_LEAN_THREADS
#define QMMX_LEAN_THREADS 1024
#define QMMX_LEAN_BLOCKS 1
#endif
#ifndef QMMX_LEAN_U
#define QMMX_LEAN_U 2                 // rows per half of the register double buffer of the lockstep scan (2 * U rows in flight)
#endif
constexpr int kLeanThreads = QMMX_LEAN_THREADS;
constexpr int kLeanBlocksPerSM = QMMX_LEAN_BLOCKS;
constexpr int kLeanWarps = kLeanThreads / 32;
constexpr int kTouchStride = kLeanThreads + 4;      // bytes per level row; stride/4 is odd, so level rows rotate over the banks
constexpr int kNoCool = 0x7fffffff;
constexpr int kSmemDays = 16;         // per-day logistic masks of up to this many days are staged in shared memory
static_assert(kMaxLevels == 64, "DayTracker offsets are day * 64 (day_off >> 6 below)");

// Dynamic shared memory of the lean kernels: ONE base address (a register) + compile-time offsets, read with
// ld.shared [reg + imm].  (Static __shared__ symbols make ptxas rebuild the CTA's shared-window base from
// SR_CgaCtaId at every use inside the divergent blocks: ~10 issue slots per step in the first version.)
template <int LPAD>
struct LeanLayout {
    static constexpr int kCell = 0;                                  // uint8 [kMaxCells]
    static constexpr int kGate = kCell + kMaxCells;                  // float4 [LPAD][2][16]: {lo, +-hi, exit-low, exit-high}
    static constexpr int kTouch = kGate + LPAD * 32 * 16;            // uint8 [LPAD][kTouchStride]
    static constexpr int kPrice = kTouch + LPAD * kTouchStride;      // double [LPAD]
    // one record per cell-byte value v (tables.h, class partition), kRecShift = log2 of its size:
    //   v < mix_base: class  {level touched, levels whose contact window holds p (32 bits; 64 bits in .z/.w when LPAD = 64)}
    //   else:         single-boundary cell {float bound, uint32 lo | hi << 8}: p <= bound ? lo : hi (class ids or kFar)
    static constexpr int kRecShift = LPAD > 32 ? 4 : 3;
    static constexpr int kRec = kPrice + LPAD * 8;
    static constexpr int kMask = kRec + (256 << kRecShift);          // uint32 [kSmemDays][LPAD]: per-day logistic masks (MULTIDAY)
    static constexpr int kEnd = kMask + kSmemDays * LPAD * 4;
    // exits on high/low only (lockstep kernels; the tile kernels end at kEnd): {double +-exit price, uint32 c_exit increment}
    // per [level][side][low hit][high hit], 16 bytes each -- the stop/target level an exit pays (M:3636), negated for a short
    static constexpr int kExit = (kEnd + 15) & ~15;
    static constexpr int kEndExit = kExit + LPAD * 8 * 16;
    static constexpr int kBoot = kEndExit;                           // float [3][kMaxWindow]: bootstrap return ratios (fused bootstrap only)
    static constexpr int kEndBoot = kBoot + 3 * kMaxWindow * 4;
    static_assert(kGate % 16 == 0 && kTouch % 16 == 0 && kPrice % 8 == 0 && kRec % 16 == 0, "alignment");
};
inline size_t lean_smem_bytes(int L) { return L > 32 ? (size_t)LeanLayout<64>::kEndExit : (size_t)LeanLayout<32>::kEndExit; }

// -DQMMX_DEBUG_BOUNDS (build.py build_library(defines=...)): every shared-memory access of the lean and tile kernels
// is checked against the kernel's dynamic window and its natural alignment, a violation traps (compute-sanitizer is
// closed on this pool: profiles/r2/sanitizer_closed.log).  The product build compiles the check away.
#ifdef QMMX_DEBUG_BOUNDS
__device__ __forceinline__ void smem_chk(uint32_t a, uint32_t bytes) {
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(g_cell);
    uint32_t size;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(size));
    if (a < base || a + bytes > base + size || (a & (bytes - 1u)) != 0u) __trap();
}
#else
__device__ __forceinline__ void smem_chk(uint32_t, uint32_t) {}
#endif

#define QMMX_LDS(type, ptxty, cons, nbytes)                                                                 \
    template <int OFF> __device__ __forceinline__ type lds_##ptxty(uint32_t a) {                            \
        smem_chk(a + OFF, nbytes);                                                                          \
        type v; asm("ld.shared." #ptxty " %0, [%1+%2];" : "=" cons(v) : "r"(a), "n"(OFF)); return v; }
QMMX_LDS(uint32_t, u8, "r", 1)
QMMX_LDS(uint32_t, u32, "r", 4)
QMMX_LDS(unsigned long long, u64, "l", 8)
QMMX_LDS(double, f64, "d", 8)
QMMX_LDS(float, f32, "f", 4)
#undef QMMX_LDS
template <int OFF> __device__ __forceinline__ float2 lds_f2(uint32_t a) {
    smem_chk(a + OFF, 8);
    float2 v; asm("ld.shared.v2.f32 {%0, %1}, [%2+%3];" : "=f"(v.x), "=f"(v.y) : "r"(a), "n"(OFF)); return v; }
template <int OFF> __device__ __forceinline__ uint4 lds_u4(uint32_t a) {
    smem_chk(a + OFF, 16);
    uint4 v; asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4+%5];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a), "n"(OFF)); return v; }
template <int OFF> __device__ __forceinline__ float4 lds_f4(uint32_t a) {
    smem_chk(a + OFF, 16);
    float4 v; asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a), "n"(OFF)); return v; }
// touch bytes are read-modify-written by their own thread only; volatile keeps the asm statements in program order
__device__ __forceinline__ uint32_t touch_ld(uint32_t a) { smem_chk(a, 1); uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void touch_st(uint32_t a, uint32_t v) { smem_chk(a, 1); asm volatile("st.shared.u8 [%0], %1;" :: "r"(a), "r"(v)); }

// per-CTA tables.  gate[(k*2+side)*16+tc] from the exact confidence intervals, the overtouch limit and (single day,
// AND mode) the host-evaluated logistic mask.  Encoding: in range & y>0 -> OK; in range & y<0 -> ML_CONF_LOW;
// out of range & y = -inf -> LEVEL_OVERTOUCHED (x = +inf: never in range); out of range otherwise -> CONF_LOW.
// z, w = the exit thresholds an entry at this (level, side) arms (M:1669-1674; float64 stop/target rounded toward
// the inside of the band, as block_setup does for the first-generation kernel).
// The sign encoding needs every non-empty confidence interval to have a positive upper bound (prices are
// positive) -- checked on the host before a lean kernel is chosen (qmmx_mc.cu lean_ok).
template <bool MULTIDAY, int LPAD>
__device__ inline void lean_setup(const ScanArgs& A, bool exit_table = true) {
    typedef LeanLayout<LPAD> Y;
    const DevTables* __restrict__ T = A.tables;
    const float inf = __int_as_float(0x7f800000);
    const bool fold_ml = !MULTIDAY && A.mlmask && !(A.combo.flags & QMMX_F_USE_BLEND);
    float4* gate = reinterpret_cast<float4*>(g_cell + Y::kGate);
    for (int t = threadIdx.x; t < LPAD * 32; t += blockDim.x) {
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
    double* price = reinterpret_cast<double*>(g_cell + Y::kPrice);
    for (int k = threadIdx.x; k < LPAD; k += blockDim.x) price[k] = k < A.n_levels ? T->lev[k].price : 0.0;
    for (int v = threadIdx.x; v < 256; v += blockDim.x) {
        uint4 r = make_uint4(0u, 0u, 0u, 0u);
        if (v < T->ncls) r = make_uint4(T->cls_lev[v], T->cls_mask_lo[v], T->cls_mask_lo[v], T->cls_mask_hi[v]);
        else if (v < T->ncls + T->fnmixed) r.x = (uint32_t)__float_as_int(T->fmixb[v]), r.y = (uint32_t)T->fmixlo[v] | ((uint32_t)T->fmixhi[v] << 8);
        else r.y = (uint32_t)kFar | ((uint32_t)kFar << 8);
        if (LPAD > 32) reinterpret_cast<uint4*>(g_cell + Y::kRec)[v] = r;
        else reinterpret_cast<uint2*>(g_cell + Y::kRec)[v] = make_uint2(r.x, r.y);
    }
    if (exit_table && (A.combo.flags & QMMX_F_EXIT_ON_HL)) {      // not for the tile and sweep kernels: other data lives past kEnd there
        // stop is tested first (M:3621-3636): a long stops on the low, a short on the high
        for (int t = threadIdx.x; t < LPAD * 8; t += blockDim.x) {
            const int k = t >> 3, side = (t >> 2) & 1, lo = (t >> 1) & 1, hi = t & 1;
            const bool is_stop = side ? hi != 0 : lo != 0;
            const double lp = k < A.n_levels ? T->lev[k].price : 0.0;
            double px;
            if (side == 0) px = is_stop ? lp - A.combo.sp : lp + A.combo.tp;
            else px = is_stop ? lp + A.combo.sp : lp - A.combo.tp;
            const unsigned long long bits = (unsigned long long)__double_as_longlong(side ? -px : px);
            reinterpret_cast<uint4*>(g_cell + Y::kExit)[t] = make_uint4((uint32_t)bits, (uint32_t)(bits >> 32), is_stop ? 1u : 0x10000u, 0u);
        }
    }
    if (MULTIDAY && A.mlmask && A.n_days <= kSmemDays) {
        uint32_t* mk = reinterpret_cast<uint32_t*>(g_cell + Y::kMask);
        for (int t = threadIdx.x; t < A.n_days * LPAD; t += blockDim.x) mk[t] = A.mlmask[(t / LPAD) * kMaxLevels + (t % LPAD)];
    }
}

// SLOW = the blend gate, the OnlinePolicy gate or the volume veto is enabled (float64 slow paths compiled in)
// SWEEP = the lanes of a warp are parameter combos of ONE path (lean_sweep_kernel): stop/target paddings, cooldown and
// the logistic mask are per-lane members instead of launch constants.
// LEARN = QMMX_F_ONLINE_RETRAIN: an exit also trains this path's OnlinePolicy head (its own kernels: the call costs the
// other SLOW kernels 13-18 % in spills when it is merely compiled in)
template <bool MULTIDAY, bool HL, bool PARITY, bool WIDE, bool SLOW, bool SWEEP = false, bool LEARN = false>
struct LeanEngine {
    typedef typename LatchType<WIDE>::T latch_t;
    static constexpr int LPAD = WIDE ? 64 : 32;
    typedef LeanLayout<LPAD> Y;
    double l_sp, l_tp;     // SWEEP only: this lane's STOP_PADDING / TP_PADDING
    int l_cs;              // SWEEP only: this lane's cooldown in steps
    uint32_t l_mask;       // SWEEP only: shared address of this lane's logistic-mask row (0 = gate off)
    __device__ __forceinline__ double sp_(const ScanArgs& A) const { return SWEEP ? l_sp : A.combo.sp; }
    __device__ __forceinline__ double tp_(const ScanArgs& A) const { return SWEEP ? l_tp : A.combo.tp; }
    __device__ __forceinline__ int cs_(const ScanArgs& A) const { return SWEEP ? l_cs : A.combo.cool_steps; }
    uint32_t sb, st;       // shared addresses: table base; this thread's column of touch bytes
    float prev, d;         // last price; last non-zero price change (0 = direction unknown, M:1529-1538)
    int cool;              // first step at which an entry may be evaluated; kNoCool while in a position
    float lo_thr, hi_thr;  // exit when low <= lo_thr or high >= hi_thr; (-inf, +inf) when flat
    float entry_s;         // entry price, sign flipped for a short: pnl = (+-exit) - entry_s
    uint32_t sgn;          // 0 long, 0x80000000 short
    int pos, entry_bar;    // level | side << 8 (stale when flat); entry_bar is only read in PARITY mode
    uint32_t xoff;         // HL: byte offset of this position's (level, side) rows of the exit table
    uint32_t inc_lo;       // c_exit increment of an exit through the LOW threshold: 1 (stop, long) or 0x10000 (target, short)
    int a_inpos;           // sum of (exit bar - entry bar) = steps coded IN_POSITION (M:1508), kept as running -entry +exit
    latch_t latch;
    uint32_t c_gate, c_ml, c_exit, c_span, c_dir;    // packed 16-bit tallies, as PathEngine (c_span is built in finish)
    double pnl, pnl2, mtm;
    int ntr; int64_t path;

    __device__ __forceinline__ bool is_open() const { return cool == kNoCool; }

    __device__ __forceinline__ void reset(uint32_t sb_, int64_t path_) {
        path = path_;
        sb = sb_; st = sb_ + Y::kTouch + threadIdx.x;
        prev = 0.f; d = 0.f; cool = 0; pos = -1; entry_s = 0.f; sgn = 0u; entry_bar = 0; inc_lo = 1u; a_inpos = 0; xoff = 0u;
        lo_thr = __int_as_float(0xff800000); hi_thr = __int_as_float(0x7f800000);
        latch = 0; c_gate = c_ml = c_exit = c_span = c_dir = 0u;
        pnl = 0.0; pnl2 = 0.0; mtm = 0.0; ntr = 0;
#pragma unroll 8
        for (int k = 0; k < LPAD; ++k) touch_st(st + k * kTouchStride, 0u);
    }

    // steps 0 .. j0-1 carry no direction (DIR_UNKNOWN, M:1538): j0 = index of the first close that differs from close[0]
    __device__ __forceinline__ void prologue(float p0, int j0) { prev = p0; cool = j0; c_dir = (uint32_t)j0; }

    __device__ __forceinline__ float entry_price() const { return __int_as_float(__float_as_int(entry_s) ^ sgn); }

    __device__ __forceinline__ void do_exit(const ScanArgs& A, float pf, bool hit_lo, bool hit_hi, int i) {
        const bool is_short = sgn != 0u;
        const bool is_stop = is_short ? hit_hi : hit_lo;         // stop is tested first (M:2975-3010, M:3621-3636)
        double tp;
        float exit_f = pf;
        bool win;
        if (!HL) {
            // close exit (M:2979-3012): fp32 values, float64 difference (M:1926); a short negates both operands (exact)
            const float xs = __int_as_float(__float_as_int(pf) ^ sgn);
            tp = (double)xs - (double)entry_s;
            win = xs > entry_s;                                   // pnl > 0 (label, M:1936), decided on the fp32 operands
            c_exit += hit_lo ? inc_lo : (inc_lo ^ 0x10001u);      // exactly one threshold is hit by a single price
        } else if (!SWEEP) {                                      // exit at the stop/target level itself (M:3636), from the exit table
            // +-exit price (negated for a short, as entry_s is): pnl = (+-exit) - entry_s, exactly exit - entry or entry - exit
            uint32_t xa = sb + xoff;
            if (hit_lo) xa += 32u;
            if (hit_hi) xa += 16u;
            const uint4 r = lds_u4<Y::kExit>(xa);
            const double sx = __longlong_as_double((long long)(((unsigned long long)r.y << 32) | r.x));
            tp = sx - (double)entry_s;
            win = tp > 0;
            c_exit += r.z;
            if (PARITY) exit_f = (float)(is_short ? -sx : sx);
        } else {                                                  // the sweep kernel: this lane's own paddings
            const double lp = lds_f64<Y::kPrice>(sb + (pos & 0xff) * 8);
            double exit_px;
            if (!is_short) exit_px = is_stop ? lp - sp_(A) : lp + tp_(A);
            else exit_px = is_stop ? lp + sp_(A) : lp - tp_(A);
            const double e = (double)entry_price();
            tp = is_short ? (e - exit_px) : (exit_px - e);
            exit_f = (float)exit_px;
            win = tp > 0;
            c_exit += is_stop ? 1u : (1u << 16);
        }
        if (LEARN && A.polw) {                                     // QMMX_F_ONLINE_RETRAIN: learn from this trade (engine.cuh policy_learn)
            const double lp = lds_f64<Y::kPrice>(sb + (pos & 0xff) * 8);
            policy_learn(&A, path, is_short ? 1 : 0, fabs(lp - (double)entry_price()), win);
        }
        pnl += tp; pnl2 += tp * tp;
        a_inpos += i;                                             // IN_POSITION steps of this trade: i - entry bar
        cool = i + cs_(A);                                        // M:2982, M:269-270; COOLDOWN steps are counted in finish
        if (win) c_ml += 1u << 16;
        if (PARITY) record_trade(&A, path, ntr, entry_bar, i, is_short ? 1 : 0, is_stop ? QMMX_EXIT_STOP : QMMX_EXIT_TARGET,
                                 pos & 0xff, entry_price(), exit_f, (float)tp);
        ++ntr;
        lo_thr = __int_as_float(0xff800000); hi_thr = __int_as_float(0x7f800000);
    }

    // cls: class id of the price (tables.h): the level touched and the levels whose contact window holds it.
    // side: M:1645, up -> long (0), down -> short (1)
    __device__ __forceinline__ int contact(const ScanArgs& A, int cls, float pf, int i, int day_off, int side) {
        // latch + touch count (M:1562-1576): a latched level stays latched only while p is inside its own window, which is
        // what the class mask says for every level at once (the touched level's own bit is in it)
        int k;
        latch_t win;
        if (WIDE) {
            const uint4 r = lds_u4<Y::kRec>(sb + (cls << 4));
            k = (int)r.x; win = (latch_t)(((unsigned long long)r.w << 32) | r.z);
        } else {
            const unsigned long long r = lds_u64<Y::kRec>(sb + (cls << 3));
            k = (int)(uint32_t)r; win = (latch_t)(r >> 32);
        }
        const latch_t bit = (latch_t)1 << k;
        const bool isnew = (latch & bit) == 0;
        latch = (latch & win) | bit;
        const uint32_t ta = st + k * kTouchStride;
        uint32_t tc = touch_ld(ta);
        if (isnew) { tc = min(tc + 1u, 15u); touch_st(ta, tc); }
        const float4 g = lds_f4<Y::kGate>(sb + (((k * 2 + side) * 16 + (int)tc) << 4));
        const bool neg = __float_as_int(g.y) < 0;
        if (!(pf >= g.x && pf <= fabsf(g.y))) {
            const bool over = __float_as_int(g.y) == (int)0xff800000;   // y = -inf marks tc >= overtouch limit
            c_gate += over ? 1u : (1u << 16);                     // M:1579 / M:1628
            return over ? QMMX_LEVEL_OVERTOUCHED : QMMX_CONF_LOW;
        }
        if (SLOW && A.regime) { const int rc = regime_gate(&A, side, i, path); if (rc) return rc; }      // M:1652-1666
        if (SLOW && A.vol) { const int vc = veto_check(&A, k, pf, side, i, path); if (vc) return vc; }   // M:1678-1705
        if (neg) { c_ml += 1u; return QMMX_ML_CONF_LOW; }         // M:1745-1755 (single day, AND mode)
        if (SWEEP && l_mask) {                                    // the same gate with this lane's mask (depends on STOP_PADDING)
            if ((lds_u32<0>(l_mask + k * 4) >> (side * 16 + (int)tc)) & 1u) { c_ml += 1u; return QMMX_ML_CONF_LOW; }
        }
        if (SLOW && (A.combo.flags & QMMX_F_USE_BLEND)) {         // M:1730-1742
            if (blend_fails(&A, k, pf, (int)tc, side, day_off)) return QMMX_COMBINED_LOW;
        } else if (MULTIDAY && A.mlmask) {
            // day_off = day * kMaxLevels (DayTracker); masks of short horizons sit in shared memory
            const uint32_t m = A.n_days <= kSmemDays ? lds_u32<Y::kMask>(sb + ((day_off >> 6) * LPAD + k) * 4)
                                                     : __ldg(&A.mlmask[day_off + k]);
            if ((m >> (side * 16 + (int)tc)) & 1u) { c_ml += 1u; return QMMX_ML_CONF_LOW; }
        }
        if (SLOW && !(A.combo.flags & QMMX_F_DISABLE_POLICY_GATE)) {   // OnlinePolicy gate of the live loop (M:3046-3109)
            if (!policy_passes<LEARN>(&A, k, pf, side, i, path)) return QMMX_ONLINE_POLICY;
        }
        if (SWEEP) {                                              // M:1669-1674 with this lane's paddings (as lean_setup rounds them)
            const double lp = lds_f64<Y::kPrice>(sb + k * 8);
            lo_thr = __double2float_rd(side == 0 ? lp - l_sp : lp - l_tp);
            hi_thr = __double2float_ru(side == 0 ? lp + l_tp : lp + l_sp);
        } else { lo_thr = g.z; hi_thr = g.w; }                    // M:1669-1674
        pos = k | (side << 8);
        if (HL && !SWEEP) xoff = (uint32_t)(k * 2 + side) << 6;
        sgn = (uint32_t)side << 31;
        entry_s = __int_as_float(__float_as_int(pf) ^ sgn);
        inc_lo = side ? 0x10000u : 1u;
        a_inpos -= i;
        if (PARITY) entry_bar = i;
        cool = kNoCool;
        return QMMX_OK_ENTRY;
    }

    __device__ __forceinline__ void step(const ScanArgs& A, float pf, float hh, float ll, int i, int day_off) {
        int code = QMMX_TOO_FAR;
        if (PARITY) code = (cool == kNoCool) ? QMMX_IN_POSITION : (i >= cool ? QMMX_TOO_FAR : QMMX_COOLDOWN);
        const float xl = HL ? ll : pf, xh = HL ? hh : pf;
        const bool hit_lo = xl <= lo_thr, hit_hi = xh >= hi_thr;
        if (hit_lo | hit_hi) do_exit(A, pf, hit_lo, hit_hi, i);
        // decided after the exit on purpose (one live copy of `cool`): an exit leaves cool = i + cool_steps > i, so the
        // exit tick itself is never eligible, exactly as when the test is made before the exit (cool_steps >= 1, qmmx_mc.cu)
        const bool elig = i >= cool;
        const float diff = pf - prev;
        if (diff != 0.f) d = diff;
        prev = pf;
        if (PARITY && d == 0.f) code = QMMX_DIR_UNKNOWN;
        // nearest level within CONTACT_PROX (M:1543-1545): the cell byte is read by every lane, the branch is taken by few
        const uint32_t c = min((uint32_t)(__float_as_int(pf) - A.cell_obase), A.cell_lim) >> A.cell_shift;
        uint32_t v = lds_u8<Y::kCell>(sb + c);
        if (elig && v != kFar) {
            if (v >= A.mix_base) {                                // the cell straddles a class boundary
                if (v == kMixed) {                                // several boundaries, or an end cell (price left the grid)
                    v = resolve_mixed(A, pf);
                } else {
                    const float2 r = lds_f2<Y::kRec>(sb + (v << Y::kRecShift));
                    const uint32_t lohi = (uint32_t)__float_as_int(r.y);
                    v = pf <= r.x ? (lohi & 0xffu) : (lohi >> 8);
                }
            }
            if (v != kFar) code = contact(A, (int)v, pf, i, day_off, d < 0.f ? 1 : 0);
        }
        if (PARITY && A.codes) {
            if (A.n_levels == 0 && (code == QMMX_TOO_FAR || code == QMMX_DIR_UNKNOWN)) code = QMMX_NOLEVELS;
            A.codes[(int64_t)i * A.paths + path] = (uint8_t)code;
        }
    }

    // a kMixed cell: the far zones next to the grid, then the few class runs outside the grid (a price that wandered off:
    // bootstrap paths drift for days), then the zone search (interior cells with several boundaries)
    __device__ __forceinline__ uint32_t resolve_mixed(const ScanArgs& A, float pf) const {
        if ((pf > A.guard[0] && pf <= A.guard[1]) || (pf > A.guard[2] && pf <= A.guard[3])) return kFar;
        if (A.n_out >= 0 && (pf < A.grid_lo || pf > A.grid_hi)) {
            uint32_t v = kFar;
#pragma unroll
            for (int j = 0; j < kMaxOut; ++j) if (pf >= A.out_lo[j] && pf <= A.out_hi[j]) v = A.out_cls[j];
            return v;
        }
        return classify_zones(pf);
    }

    // class of p (the argument contact() takes), or kFar: the cell byte of step(), fully resolved
    __device__ __forceinline__ uint32_t classify(const ScanArgs& A, float pf) const {
        const uint32_t c = min((uint32_t)(__float_as_int(pf) - A.cell_obase), A.cell_lim) >> A.cell_shift;
        uint32_t v = lds_u8<Y::kCell>(sb + c);
        if (v >= A.mix_base && v != kFar) {
            if (v == kMixed) {
                // tile-event and sweep kernels: without the list of class runs outside the grid (their prices rarely leave it, and
                // the longer inline form costs the close-only tile kernel 1.3 % through the register allocation of its rounds;
                // an out-of-line call costs 9 %)
                if ((pf > A.guard[0] && pf <= A.guard[1]) || (pf > A.guard[2] && pf <= A.guard[3])) v = kFar;
                else v = classify_zones(pf);
            } else {
                const float2 r = lds_f2<Y::kRec>(sb + (v << Y::kRecShift));
                const uint32_t lohi = (uint32_t)__float_as_int(r.y);
                v = pf <= r.x ? (lohi & 0xffu) : (lohi >> 8);
            }
        }
        return v;
    }

    __device__ __forceinline__ void finish(const ScanArgs& A) {
        // COOLDOWN steps (M:1516): every exit is followed by cool_steps - 1 of them, except that the last one may be cut
        // by the end of the path (only the last: no entry can follow an exit whose cooldown outlasts the path)
        const int n_exits = (int)((c_exit & 0xffffu) + (c_exit >> 16));
        int n_cool = n_exits * max(cs_(A) - 1, 0);
        if (!is_open() && n_exits > 0 && cool > A.steps) n_cool -= cool - A.steps;
        if (is_open()) {
            const bool is_short = sgn != 0u;
            const double e = (double)entry_price(), last = (double)prev;
            mtm = is_short ? (e - last) : (last - e);
            a_inpos += A.steps - 1;
            if (PARITY) record_trade(&A, path, ntr, entry_bar, -1, is_short ? 1 : 0, QMMX_EXIT_OPEN, pos & 0xff, entry_price(), prev, 0.0f);
            ++ntr;
        }
        c_span = (uint32_t)a_inpos | ((uint32_t)n_cool << 16);
        if (PARITY && A.tcounts) A.tcounts[path] = ntr;
        if (!SWEEP) atomicAdd(&g_S.hist[hist_bin(A)], 1u);        // the sweep kernel keeps one histogram per lane
    }

    __device__ __forceinline__ int hist_bin(const ScanArgs& A) const {
        const double fb = floor((pnl - A.hist_lo) * A.hist_inv_w);
        return fb < 0 ? 0 : (fb > QMMX_PNL_HIST_BINS - 1 ? QMMX_PNL_HIST_BINS - 1 : (int)fb);
    }
};

// Tiles are claimed from a per-launch counter instead of being dealt out round-robin: 1M paths are 6.6 tiles per resident
// warp, so a static deal leaves 6 % of the machine idle behind the warps that drew 7, and tiles differ in work.
__device__ __forceinline__ int64_t next_tile(const ScanArgs& A, int lane) {
    unsigned int t = 0u;
    if (lane == 0) t = atomicAdd(A.tile_counter, 1u);
    return (int64_t)__shfl_sync(0xffffffffu, t, 0);
}

// Bars are [step][pitch]: one coalesced 128-byte line per warp, step and field; 2U rows are kept in flight per
// thread in registers (register double buffer) -- at 1024 threads per SM a warp returns to the loop every ~0.3 us,
// HBM latency is ~0.8 us.
template <bool MULTIDAY, bool HL, bool PARITY, bool WIDE, bool SLOW, int U, bool LEARN = false>
__device__ __forceinline__ void lean_scan_path(LeanEngine<MULTIDAY, HL, PARITY, WIDE, SLOW, false, LEARN>& E, const ScanArgs& A,
                                               int64_t path) {
    const size_t pitch = (size_t)A.pitch;
    const float* pc = A.close + path;
    const float* ph = HL ? A.high + path : nullptr;
    const float* pl = HL ? A.low + path : nullptr;
    const int steps = A.steps;
    {   // DIR_UNKNOWN prefix
        const float p0 = __ldg(pc);
        int j0 = 1;
        while (j0 < steps && __ldg(pc + (size_t)j0 * pitch) == p0) ++j0;
        E.prologue(p0, j0);
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
            for (int u = 0; u < U; ++u) E.step(A, c0[u], HL ? h0[u] : 0.f, HL ? l0[u] : 0.f, i + u, day.at(A, i + u));
            if (n + 1 < nfull) {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    c0[u] = __ldg(pc); pc += pitch;
                    if (HL) { h0[u] = __ldg(ph); l0[u] = __ldg(pl); ph += pitch; pl += pitch; }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) E.step(A, c1[u], HL ? h1[u] : 0.f, HL ? l1[u] : 0.f, i + U + u, day.at(A, i + U + u));
        }
    }
#pragma unroll 1
    for (; i < steps; ++i) {
        const float c = __ldg(pc); pc += pitch;
        float h = 0.f, l = 0.f;
        if (HL) { h = __ldg(ph); l = __ldg(pl); ph += pitch; pl += pitch; }
        E.step(A, c, h, l, i, day.at(A, i));
    }
}

template <bool MULTIDAY, bool HL, bool PARITY, bool WIDE, bool SLOW, bool LEARN = false>
__global__ void __launch_bounds__(kLeanThreads, kLeanBlocksPerSM) lean_scan_kernel(const __grid_constant__ ScanArgs A) {
    block_setup(A, true);
    lean_setup<MULTIDAY, WIDE ? 64 : 32>(A);
    __syncthreads();
    LeanEngine<MULTIDAY, HL, PARITY, WIDE, SLOW, false, LEARN> E;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // through a shuffle so that ptxas keeps the base in a register instead of rebuilding it from SR_CgaCtaId at every use
    const uint32_t sb = __shfl_sync(0xffffffffu, (uint32_t)__cvta_generic_to_shared(g_cell), 0);
    const int64_t ntiles = (A.paths + 31) / 32;
    for (;;) {
        const int64_t tile = next_tile(A, lane);
        if (tile >= ntiles) break;
        const int64_t path = tile * 32 + lane;
        const bool active = path < A.paths;
        E.reset(sb, path);
        if (active) { lean_scan_path<MULTIDAY, HL, PARITY, WIDE, SLOW, QMMX_LEAN_U, LEARN>(E, A, path); E.finish(A); }
        __syncwarp();
        warp_accumulate(active, E.pnl, E.pnl2, E.mtm, E.c_gate, E.c_ml, E.c_exit, E.c_span, E.c_dir, active && E.is_open());
    }
    __syncthreads();
    block_flush_agg(A.flat, A.steps);
}

// ---- fused GBM generate + scan on the lean engine: bars live in registers only (Philox4x32-10 + Box-Muller) ----------
// The DIR_UNKNOWN prefix needs close[0] and the index of the first differing close before the walk starts: the
// first Philox group(s) are generated twice (one extra group per path, ~1 % of the generator work).
template <bool MULTIDAY, bool HL, bool WIDE, bool SLOW, bool LEARN = false>
__global__ void __launch_bounds__(kLeanThreads, kLeanBlocksPerSM) lean_fused_gbm_kernel(const __grid_constant__ ScanArgs A, const GbmDev g) {
    block_setup(A, true);
    lean_setup<MULTIDAY, WIDE ? 64 : 32>(A);
    __syncthreads();
    LeanEngine<MULTIDAY, HL, false, WIDE, SLOW, false, LEARN> E;      // LEARN: QMMX_F_ONLINE_RETRAIN, heads in A.polw [14][A.pitch]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t sb = __shfl_sync(0xffffffffu, (uint32_t)__cvta_generic_to_shared(g_cell), 0);
    const int64_t ntiles = (A.paths + 31) / 32;
    const int steps = A.steps;
    for (;;) {
        const int64_t tile = next_tile(A, lane);
        if (tile >= ntiles) break;
        const int64_t path = tile * 32 + lane;
        const bool active = path < A.paths;
        E.reset(sb, path);
        if (active) {
            const uint64_t gid = g.path_offset + (uint64_t)path;
            float z[4];
            {   // prologue: close[0] and the first step whose close differs from it
                float P = g.s0, p0 = 0.f;
                int j0 = steps;
                for (int i = 0; i < steps && j0 == steps; i += HL ? 1 : 4) {
                    gbm_normals(g, gid, (uint32_t)(HL ? i : (i >> 2)), z);
                    if (HL) {
#pragma unroll
                        for (int u = 0; u < 4; ++u) P = gbm_advance(g, P, z[u]);
                        const float c = g.cents ? round_cents(P) : P;
                        if (i == 0) p0 = c; else if (c != p0) j0 = i;
                    } else {
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (i + u < steps && j0 == steps) {
                                P = gbm_advance(g, P, z[u]);
                                const float c = g.cents ? round_cents(P) : P;
                                if (i + u == 0) p0 = c; else if (c != p0) j0 = i + u;
                            }
                        }
                    }
                }
                E.prologue(p0, j0);
            }
            float P = g.s0;
            DayTracker<MULTIDAY> day;
            day.init(A);
            if (!HL) {
                for (int i = 0; i < steps; i += 4) {
                    gbm_normals(g, gid, (uint32_t)(i >> 2), z);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (i + u < steps) {
                            P = gbm_advance(g, P, z[u]);
                            E.step(A, g.cents ? round_cents(P) : P, 0.f, 0.f, i + u, day.at(A, i + u));
                        }
                    }
                }
            } else {
#pragma unroll 1
                for (int i = 0; i < steps; ++i) {
                    gbm_normals(g, gid, (uint32_t)i, z);
                    float hi = P, lo = P;
#pragma unroll
                    for (int u = 0; u < 4; ++u) { P = gbm_advance(g, P, z[u]); hi = fmaxf(hi, P); lo = fminf(lo, P); }
                    E.step(A, g.cents ? round_cents(P) : P, g.cents ? round_cents(hi) : hi, g.cents ? round_cents(lo) : lo, i, day.at(A, i));
                }
            }
            E.finish(A);
        }
        __syncwarp();
        warp_accumulate(active, E.pnl, E.pnl2, E.mtm, E.c_gate, E.c_ml, E.c_exit, E.c_span, E.c_dir, active && E.is_open());
    }
    __syncthreads();
    block_flush_agg(A.flat, A.steps);
}

// ---- fused block bootstrap + scan (BASELINE config 5 without the 12 B/step round trip through HBM) ----------------
// Same draws and the same IEEE products as gen_bootstrap_kernel (Philox counter (path id, block group, stream 1),
// start = umulhi(r, nret), circular blocks), consumed in registers.  Bars carry no volume here: the soft veto is inert.
template <bool MULTIDAY, bool HL, bool WIDE, bool SLOW, bool LEARN = false>
__global__ void __launch_bounds__(kLeanThreads, kLeanBlocksPerSM) lean_fused_boot_kernel(const __grid_constant__ ScanArgs A, const BootDev g) {
    typedef LeanLayout<WIDE ? 64 : 32> Y;
    block_setup(A, true);
    lean_setup<MULTIDAY, WIDE ? 64 : 32>(A);
    {
        float* rb = reinterpret_cast<float*>(g_cell + Y::kBoot);
        for (int t = threadIdx.x; t < g.nret; t += blockDim.x) {
            rb[t] = g.ratios[t]; rb[kMaxWindow + t] = g.ratios[g.nret + t]; rb[2 * kMaxWindow + t] = g.ratios[2 * g.nret + t];
        }
    }
    __syncthreads();
    LeanEngine<MULTIDAY, HL, false, WIDE, SLOW, false, LEARN> E;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t sb = __shfl_sync(0xffffffffu, (uint32_t)__cvta_generic_to_shared(g_cell), 0);
    const int64_t ntiles = (A.paths + 31) / 32;
    const int steps = A.steps, nret = g.nret, block_len = g.block_len;
    for (;;) {
        const int64_t tile = next_tile(A, lane);
        if (tile >= ntiles) break;
        const int64_t path = tile * 32 + lane;
        const bool active = path < A.paths;
        E.reset(sb, path);
        if (active) {
            const uint64_t gid = g.path_offset + (uint64_t)path;
            uint4 x = make_uint4(0, 0, 0, 0);
            int idx = 0, off = 0, blk = 0;                   // idx = (start + off) mod nret as a running circular index
            float S = g.s0;
            auto draw = [&](float& hi, float& lo) {          // one bar: advances (S, idx, off, blk)
                if (off == 0) {
                    if ((blk & 3) == 0) x = philox4x32_10(make_uint4((uint32_t)gid, (uint32_t)(gid >> 32), (uint32_t)(blk >> 2), 1u), g.k0, g.k1);
                    const uint32_t r = (blk & 3) == 0 ? x.x : (blk & 3) == 1 ? x.y : (blk & 3) == 2 ? x.z : x.w;
                    idx = (int)__umulhi(r, (uint32_t)nret);
                } else if (++idx == nret) idx = 0;
                const uint32_t a = sb + (uint32_t)idx * 4u;
                if (HL) { hi = __fmul_rn(S, lds_f32<Y::kBoot>(a)); lo = __fmul_rn(S, lds_f32<Y::kBoot + kMaxWindow * 4>(a)); }
                S = __fmul_rn(S, lds_f32<Y::kBoot + 2 * kMaxWindow * 4>(a));
                if (++off == block_len) { off = 0; ++blk; }
            };
            {   // prologue: close[0] and the first step whose close differs from it (DIR_UNKNOWN prefix)
                float hi, lo, p0 = 0.f;
                int j0 = steps;
                for (int i = 0; i < steps && j0 == steps; ++i) {
                    draw(hi, lo);
                    if (i == 0) p0 = S; else if (S != p0) j0 = i;
                }
                E.prologue(p0, j0);
                S = g.s0; idx = 0; off = 0; blk = 0;
            }
            DayTracker<MULTIDAY> day;
            day.init(A);
#pragma unroll 2
            for (int i = 0; i < steps; ++i) {
                float hi = 0.f, lo = 0.f;
                draw(hi, lo);
                E.step(A, S, hi, lo, i, day.at(A, i));
            }
            E.finish(A);
        }
        __syncwarp();
        warp_accumulate(active, E.pnl, E.pnl2, E.mtm, E.c_gate, E.c_ml, E.c_exit, E.c_span, E.c_dir, active && E.is_open());
    }
    __syncthreads();
    block_flush_agg(A.flat, A.steps);
}

// ---- parameter sweep: the 32 lanes of a warp are 32 parameter combos of ONE path ------------------------------------
// A sweep over stop/target paddings and cooldowns (BASELINE C4: 256 combos over the same resident bars) evaluates the
// same prices under every combo.  With lane = combo, all lanes see the same price on the same step: the cell byte,
// the "far from every level" case and the contact windows are warp-uniform, so the contact block runs on the ~20 % of
// steps where the path touches a level instead of on every step (32 independent paths always have SOME lane near a
// level).  Bars are read once per 32 combos: lane l loads step i0 + l of the path and each step is a warp shuffle.
// Combos of one launch share the tables (levels, CONTACT_PROX, Q_MIN_PROB) and all flags; stop/target paddings, the
// cooldown and the logistic mask (its feature |level - stop| is STOP_PADDING, M:1459) are per lane.
struct SweepLane { double sp, tp; int32_t cool_steps, combo; };
struct SweepArgs {
    SweepLane lane[32];
    int32_t n_lanes, pad;
    const uint32_t* mlmask;            // [combo][mask_stride] or null (no model / gate off)
    int64_t mask_stride;
    unsigned long long* flat;          // [combo][QMMX_AGG_SLOTS]
};

// a warp-step of the sweep kernel costs about half of a warp-step of the per-combo kernel, whatever the lane count
constexpr int kSweepMinLanes = 20;
constexpr int kSwU = 11, kSwD = 6;     // per-lane tallies (uint32) and float64 sums of one CTA
enum { SW_NPATHS, SW_OVER, SW_CONF, SW_ML, SW_DIR, SW_WINS, SW_STOP, SW_TGT, SW_INPOS, SW_COOL, SW_OPEN };
enum { SWD_PNL, SWD_PNL2, SWD_PATH2, SWD_MTM, SWD_MAX, SWD_NEGMIN };

template <int LPAD>
struct SweepLayout {
    typedef LeanLayout<LPAD> Y;
    static constexpr int kRow = LPAD + 1;                                     // mask row stride in words: lanes land on distinct banks
    static constexpr int kSMask = (Y::kEnd + 15) & ~15;                       // uint32 [32][kRow]
    static constexpr int kSD = (kSMask + 32 * kRow * 4 + 7) & ~7;             // double [kSwD][32]
    static constexpr int kSU = kSD + kSwD * 32 * 8;                           // uint32 [kSwU][32]
    static constexpr int kSHist = kSU + kSwU * 32 * 4;                        // uint32 [QMMX_PNL_HIST_BINS][32]
    static constexpr int kSEnd = kSHist + QMMX_PNL_HIST_BINS * 32 * 4;
};
inline size_t sweep_smem_bytes(int L) { return L > 32 ? (size_t)SweepLayout<64>::kSEnd : (size_t)SweepLayout<32>::kSEnd; }

template <bool HL, bool WIDE>
__global__ void __launch_bounds__(kLeanThreads, 1) lean_sweep_kernel(const __grid_constant__ ScanArgs A,
                                                                     const __grid_constant__ SweepArgs W) {
    constexpr int LPAD = WIDE ? 64 : 32;
    typedef SweepLayout<LPAD> Z;
    block_setup(A, true);
    lean_setup<false, LPAD>(A, false);     // A.mlmask is null here: the gate table carries no logistic sign; exits use the lane's own paddings
    uint32_t* smask = reinterpret_cast<uint32_t*>(g_cell + Z::kSMask);
    double* sd = reinterpret_cast<double*>(g_cell + Z::kSD);
    uint32_t* su = reinterpret_cast<uint32_t*>(g_cell + Z::kSU);
    uint32_t* shist = reinterpret_cast<uint32_t*>(g_cell + Z::kSHist);
    for (int t = threadIdx.x; t < 32 * Z::kRow; t += blockDim.x) {
        const int ln = t / Z::kRow, k = t % Z::kRow;
        smask[t] = (W.mlmask && ln < W.n_lanes && k < LPAD) ? W.mlmask[(size_t)W.lane[ln].combo * (size_t)W.mask_stride + k] : 0u;
    }
    const double neg_inf = __longlong_as_double((long long)0xfff0000000000000ull);
    for (int t = threadIdx.x; t < kSwD * 32; t += blockDim.x) sd[t] = t >= SWD_MAX * 32 ? neg_inf : 0.0;
    for (int t = threadIdx.x; t < kSwU * 32 + QMMX_PNL_HIST_BINS * 32; t += blockDim.x) su[t] = 0u;   // tallies + histogram (adjacent)
    __syncthreads();

    LeanEngine<false, HL, false, WIDE, false, true> E;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t sb = __shfl_sync(0xffffffffu, (uint32_t)__cvta_generic_to_shared(g_cell), 0);
    const bool live = lane < W.n_lanes;
    const int ln = live ? lane : 0;        // spare lanes shadow lane 0 and are not counted
    E.l_sp = W.lane[ln].sp; E.l_tp = W.lane[ln].tp; E.l_cs = W.lane[ln].cool_steps;
    E.l_mask = W.mlmask ? sb + Z::kSMask + ln * Z::kRow * 4 : 0u;
    const size_t pitch = (size_t)A.pitch;
    const int steps = A.steps;
    // the 32 warps of a CTA walk 32 neighbouring paths: the 128-byte line of a step is shared through L1/L2
    for (int64_t path = (int64_t)blockIdx.x * kLeanWarps + warp; path < A.paths; path += (int64_t)gridDim.x * kLeanWarps) {
        E.reset(sb, path);
        const float* pc = A.close + path;
        const float* ph = HL ? A.high + path : nullptr;
        const float* pl = HL ? A.low + path : nullptr;
        float cv = lane < steps ? __ldg(pc + (size_t)lane * pitch) : 0.f, hv = 0.f, lv = 0.f;
        if (HL && lane < steps) { hv = __ldg(ph + (size_t)lane * pitch); lv = __ldg(pl + (size_t)lane * pitch); }
        {   // DIR_UNKNOWN prefix: index of the first close that differs from close[0]
            const float p0 = __shfl_sync(0xffffffffu, cv, 0);
            int j0 = steps;
            unsigned m = __ballot_sync(0xffffffffu, lane < steps && cv != p0);
            if (m) j0 = __ffs((int)m) - 1;
            else {
                for (int i0 = 32; i0 < steps; i0 += 32) {
                    const int idx = i0 + lane;
                    const float v = idx < steps ? __ldg(pc + (size_t)idx * pitch) : p0;
                    m = __ballot_sync(0xffffffffu, v != p0);
                    if (m) { j0 = i0 + __ffs((int)m) - 1; break; }
                }
            }
            E.prologue(p0, j0);
        }
        // Each lane classifies ONE step of the 32-step chunk (price change, contact zone); three ballots then tell the
        // whole warp which steps can matter.  While no lane holds a position the walk jumps from contact step to contact
        // step; while some lane does, every step is an exit test (two compares) and nothing else.
        float carry_p = __shfl_sync(0xffffffffu, cv, 0);          // price before the chunk's first step (close[0] at the start)
        uint32_t carry_side = 0u;                                 // side of the last non-zero price change so far
        for (int i0 = 0; i0 < steps; i0 += 32) {
            const int nidx = i0 + 32 + lane;                      // next chunk in flight while this one is walked
            float cn = 0.f, hn = 0.f, lnx = 0.f;
            if (nidx < steps) {
                cn = __ldg(pc + (size_t)nidx * pitch);
                if (HL) { hn = __ldg(ph + (size_t)nidx * pitch); lnx = __ldg(pl + (size_t)nidx * pitch); }
            }
            const int n = min(32, steps - i0);
            const float up = __shfl_up_sync(0xffffffffu, cv, 1);
            const float df = lane < n ? cv - (lane == 0 ? carry_p : up) : 0.f;
            const unsigned nz = __ballot_sync(0xffffffffu, df != 0.f), ng = __ballot_sync(0xffffffffu, df < 0.f);
            const uint32_t vv = lane < n ? E.classify(A, cv) : (uint32_t)kFar;
            const unsigned cm = __ballot_sync(0xffffffffu, vv != kFar);
            // side of the last non-zero price change up to this lane's step (M:1529-1538, M:1645), packed beside the level
            const unsigned mine = nz & (0xffffffffu >> (31 - lane));
            const uint32_t vs = vv | ((mine ? (ng >> (31 - __clz((int)mine))) & 1u : carry_side) << 8);
            int u = 0;
            while (u < n) {
                __syncwarp();
                if (!__any_sync(0xffffffffu, E.cool == kNoCool)) {
                    const unsigned rem = cm >> u;
                    if (!rem) break;
                    u += __ffs((int)rem) - 1;
                }
                const int i = i0 + u;
                const float pf = __shfl_sync(0xffffffffu, cv, u);
                const float xl = HL ? __shfl_sync(0xffffffffu, lv, u) : pf, xh = HL ? __shfl_sync(0xffffffffu, hv, u) : pf;
                const bool elig = i >= E.cool;                    // as step(): decided before this step's exit
                const bool hit_lo = xl <= E.lo_thr, hit_hi = xh >= E.hi_thr;
                if (hit_lo | hit_hi) E.do_exit(A, pf, hit_lo, hit_hi, i);
                if ((cm >> u) & 1u) {
                    const uint32_t w = __shfl_sync(0xffffffffu, vs, u);
                    if (elig) E.contact(A, (int)(w & 0xffu), pf, i, 0, (int)(w >> 8));
                }
                ++u;
            }
            if (nz) carry_side = (ng >> (31 - __clz((int)nz))) & 1u;
            carry_p = __shfl_sync(0xffffffffu, cv, n - 1);
            cv = cn; hv = hn; lv = lnx;
        }
        E.prev = carry_p;                                         // last close (mark-to-market of an open position)
        E.finish(A);
        if (live) {
            uint32_t* u = su + lane;
            double* dd = sd + lane;
            atomicAdd(u + SW_NPATHS * 32, 1u);
            if (E.c_gate & 0xffffu) atomicAdd(u + SW_OVER * 32, E.c_gate & 0xffffu);
            if (E.c_gate >> 16) atomicAdd(u + SW_CONF * 32, E.c_gate >> 16);
            if (E.c_ml & 0xffffu) atomicAdd(u + SW_ML * 32, E.c_ml & 0xffffu);
            if (E.c_ml >> 16) atomicAdd(u + SW_WINS * 32, E.c_ml >> 16);
            if (E.c_dir) atomicAdd(u + SW_DIR * 32, E.c_dir);
            if (E.c_exit & 0xffffu) atomicAdd(u + SW_STOP * 32, E.c_exit & 0xffffu);
            if (E.c_exit >> 16) atomicAdd(u + SW_TGT * 32, E.c_exit >> 16);
            if (E.c_span & 0xffffu) atomicAdd(u + SW_INPOS * 32, E.c_span & 0xffffu);
            if (E.c_span >> 16) atomicAdd(u + SW_COOL * 32, E.c_span >> 16);
            if (E.is_open()) atomicAdd(u + SW_OPEN * 32, 1u);
            if (E.pnl != 0.0) { atomicAdd(dd + SWD_PNL * 32, E.pnl); atomicAdd(dd + SWD_PATH2 * 32, E.pnl * E.pnl); }
            if (E.pnl2 != 0.0) atomicAdd(dd + SWD_PNL2 * 32, E.pnl2);
            if (E.mtm != 0.0) atomicAdd(dd + SWD_MTM * 32, E.mtm);
            atomic_max_f64(dd + SWD_MAX * 32, E.pnl); atomic_max_f64(dd + SWD_NEGMIN * 32, -E.pnl);
            atomicAdd(shist + E.hist_bin(A) * 32 + lane, 1u);
        }
        __syncwarp();
    }
    __syncthreads();
    if (threadIdx.x < W.n_lanes) {        // one thread per combo folds the CTA's tallies into that combo's reduction buffer
        const uint32_t* u = su + threadIdx.x;
        const double* dd = sd + threadIdx.x;
        unsigned long long* gu = W.flat + (size_t)W.lane[threadIdx.x].combo * QMMX_AGG_SLOTS;
        double* gd = reinterpret_cast<double*>(gu + QMMX_AGG_NU);
        const unsigned long long np = u[SW_NPATHS * 32];
        if (np) {
            const unsigned long long nsteps = np * (unsigned long long)steps;
            const unsigned long long ntr = (unsigned long long)u[SW_STOP * 32] + u[SW_TGT * 32], nopen = u[SW_OPEN * 32];
            const unsigned long long coded = (unsigned long long)u[SW_OVER * 32] + u[SW_CONF * 32] + u[SW_ML * 32] + u[SW_DIR * 32] +
                                             u[SW_INPOS * 32] + u[SW_COOL * 32] + ntr + nopen;
            auto add = [&](int slot, unsigned long long v) { if (v) atomicAdd(&gu[slot], v); };
            add(U_NPATHS, np); add(U_NSTEPS, nsteps); add(U_NTRADES, ntr); add(U_NWINS, u[SW_WINS * 32]); add(U_NOPEN, nopen);
            add(U_ENTRY + QMMX_OK_ENTRY, ntr + nopen);
            add(U_ENTRY + QMMX_LEVEL_OVERTOUCHED, u[SW_OVER * 32]); add(U_ENTRY + QMMX_CONF_LOW, u[SW_CONF * 32]);
            add(U_ENTRY + QMMX_ML_CONF_LOW, u[SW_ML * 32]); add(U_ENTRY + QMMX_DIR_UNKNOWN, u[SW_DIR * 32]);
            add(U_ENTRY + QMMX_IN_POSITION, u[SW_INPOS * 32]); add(U_ENTRY + QMMX_COOLDOWN, u[SW_COOL * 32]);
            add(U_ENTRY + QMMX_TOO_FAR, nsteps - coded);
            add(U_EXIT + QMMX_EXIT_STOP, u[SW_STOP * 32]); add(U_EXIT + QMMX_EXIT_TARGET, u[SW_TGT * 32]);
            add(U_EXIT + QMMX_EXIT_OPEN, nopen);
            if (dd[SWD_PNL * 32] != 0.0) atomicAdd(&gd[D_SUM_PNL], dd[SWD_PNL * 32]);
            if (dd[SWD_PNL2 * 32] != 0.0) atomicAdd(&gd[D_SUM_PNL2], dd[SWD_PNL2 * 32]);
            if (dd[SWD_PATH2 * 32] != 0.0) atomicAdd(&gd[D_SUM_PATH2], dd[SWD_PATH2 * 32]);
            if (dd[SWD_MTM * 32] != 0.0) atomicAdd(&gd[D_SUM_MTM], dd[SWD_MTM * 32]);
            atomic_max_f64(&gd[QMMX_AGG_ND + 0], dd[SWD_MAX * 32]);
            atomic_max_f64(&gd[QMMX_AGG_ND + 1], dd[SWD_NEGMIN * 32]);
        }
    }
    for (int t = threadIdx.x; t < QMMX_PNL_HIST_BINS * 32; t += blockDim.x) {
        const int ln2 = t & 31;
        const uint32_t v = shist[t];
        if (v && ln2 < W.n_lanes)
            atomicAdd(&W.flat[(size_t)W.lane[ln2].combo * QMMX_AGG_SLOTS + U_HIST + (t >> 5)], (unsigned long long)v);
    }
}

}  // namespace qmmx
