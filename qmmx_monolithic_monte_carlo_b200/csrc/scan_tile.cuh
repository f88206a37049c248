// Tile-event engine scan: the production scan for close-only exits (resident bars and fused Philox GBM).
//
// Why: ncu on the lockstep lean kernel (profiles/r1_ncu_lean_scan_close_summary.txt) shows 91.5 warp-instructions per
// warp-step at 15.4 of 32 active lanes: with 32 independent paths per warp SOME lane touches a level on ~100 % of the
// steps and some lane exits on ~70 %, so every warp-step pays the contact block and the exit block at half-empty
// lanes.  The oracle's traces say a path has 89 evaluated contacts, 43 exit-tested steps and 16 exits per 390 steps
// (profiles/sim_event_schedule.py), i.e. 62 % of the steps need nothing but "is this price near a level".
//
// Structure (same reference semantics as scan_lean.cuh: M:1492-1771 gates, M:2966-3014 exits):
//  * a warp owns 32 paths; the last kRing = 32 rows of their closes live in a per-warp shared-memory ring
//    [row][lane] (bank = lane: a lane may read ANY row of its own column without a bank conflict);
//  * phase 1, lockstep, all lanes active, per chunk of kChunk rows: stage the rows (cp.async from the resident bars,
//    prefetched one chunk ahead; or Philox + Box-Muller + exp for the fused kernel), look up the contact cell of every
//    price and shift the answer into a per-lane 32-bit far mask, track the direction (last non-zero close-to-close
//    difference, M:1529-1538) and shift its sign into a second mask: ~14 instructions per step at full occupancy;
//  * phase 2, rounds: every lane jumps straight to ITS next event -- a flat lane to its next eligible contact
//    (one clz on the far mask; steps in cooldown are masked off), a lane in a position tests its next two rows
//    for its stop/target -- and the warp runs the contact block with all lanes that stand at a contact, the exit block
//    with all lanes that hit.  Lanes may lag kLag rows behind the newest row, so a lane with a cluster of contacts
//    does not hold the others back at every chunk boundary (simulated: 0.37 rounds per step against 0.48 for fixed
//    32-row tiles).
// Decisions are the lean engine's own contact() / do_exit() on the same tables, so every aggregate is bit-identical
// (tests/test_gpu_parity.py::test_gpu_tile_kernel_*).
#pragma once
#include <cuda.h>      // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)

#include "scan_lean.cuh"

namespace qmmx {

constexpr int kRing = 32;                 // rows in the ring = bits in the masks
#ifndef QMMX_TILE_CHUNK
#define QMMX_TILE_CHUNK 8
#endif
constexpr int kChunk = QMMX_TILE_CHUNK;   // rows staged per phase-1 pass
constexpr int kNoDir = 0x7ffffffe;        // `cool` of a flat path whose direction is still unknown (never eligible)
constexpr int kRingBytes = kRing * 128;   // per warp

constexpr int kTileRingOff = (LeanLayout<32>::kEnd + 127) & ~127;
constexpr int kTileBarOff = kTileRingOff + kLeanWarps * kRingBytes;      // one 8-byte mbarrier per warp (TMA staging)
inline size_t tile_smem_bytes() { return (size_t)kTileBarOff + (size_t)kLeanWarps * 8 + 128; }   // + slack: the rings are aligned to 128 bytes at run time
// the dynamic window starts after the kernel's static shared memory, at an address that is only 16-byte aligned
__device__ __forceinline__ uint32_t tile_ring_base(uint32_t sb) { return (sb + kTileRingOff + 127u) & ~127u; }

// ---- TMA (cp.async.bulk.tensor) + mbarrier: one elected lane stages a whole chunk [kChunk rows][32 paths] of the
// resident closes into the warp's ring with ONE instruction; the lanes wait on the warp's mbarrier.
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0u;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    int spins = 0;
    while (!mbar_try_wait(bar, parity)) if (++spins > (1 << 24)) __trap();     // a lost copy must not hang the device
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int x, int y, uint32_t bar) {
    smem_chk(dst, 128); smem_chk(dst + (kChunk - 1) * 128, 128); smem_chk(bar, 8);
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :: "r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar) : "memory");
}

__device__ __forceinline__ void cp_async4(uint32_t dst, const char* src) {
    smem_chk(dst, 4);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ float ring_ld(uint32_t a) { smem_chk(a, 4); float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
__device__ __forceinline__ void ring_st(uint32_t a, float v) { smem_chk(a, 4); asm volatile("st.shared.f32 [%0], %1;" :: "r"(a), "f"(v)); }

// SRC: 0 = resident bars [step][pitch], 1 = Philox GBM generated in registers (close-only, sub_steps = 1)
template <bool MULTIDAY, bool SLOW, int SRC>
struct TileScan {
    typedef LeanEngine<MULTIDAY, false, false, false, SLOW> Eng;
    typedef LeanLayout<32> Y;
    static constexpr int kLag = SRC == 0 ? kRing - 2 * kChunk : kRing - kChunk;   // a prefetched chunk occupies its slots early

    Eng E;
    uint32_t rb;          // shared address of this lane's column of the ring
    uint32_t fm, dm;      // far mask / direction mask: bit j = row F-1-j (far from every level / last price change was down)
    int u;                // first step this lane has not dealt with yet
    float dcur;           // last non-zero close-to-close difference up to the newest row (0 = none yet)
    bool found;           // the first price change has been seen (direction known, M:1529-1538)

    __device__ __forceinline__ uint32_t row_addr(int step) const { return rb + ((uint32_t)(step & (kRing - 1)) << 7); }

    // phase 1 for one row: price p of step i (already in the ring)
    template <bool SEARCH>
    __device__ __forceinline__ void row(const ScanArgs& A, float p, int i) {
        if (SEARCH && i == 0) E.prev = p;
        const float diff = p - E.prev;
        if (diff != 0.f) {
            dcur = diff;
            if (SEARCH && !found) { found = true; E.cool = i; E.c_dir = (uint32_t)i; }   // DIR_UNKNOWN prefix ends here
        }
        E.prev = p;
        dm = __funnelshift_l(__float_as_uint(dcur), dm, 1);
        const uint32_t c = min((uint32_t)(__float_as_int(p) - A.cell_obase), A.cell_lim) >> A.cell_shift;
        const uint32_t v = lds_u8<Y::kCell>(E.sb + c);
        fm = __funnelshift_l((v + 1u) << 23, fm, 1);              // v == kFar (0xff) -> bit 31
    }

    // phase 2: rounds until every lane has dealt with all its events before `must`; events up to F (exclusive) may run.
    // Invariant: u <= F; a flat lane's u is its next eligible contact (F when there is none among the loaded rows).
    __device__ __forceinline__ void rounds(const ScanArgs& A, int F, int must) {
#pragma unroll 1
        for (;;) {
            // next eligible contact of a flat lane: rows [max(u, cool), F) whose far bit is clear.  A lane in a position
            // has cool = INT_MAX: no candidate rows, m = 0.
            const int a = max(u, E.cool);
            const int n = max(F - a, 0);
            const uint32_t m = ~fm & __funnelshift_lc(0xffffffffu, 0u, (uint32_t)n);   // low n bits (n >= 32: all)
            const int j = 31 - __clz((int)m);                      // oldest set bit = earliest row; -1 when m == 0
            const int ev = F - 1 - j;                              // F when m == 0
            if (E.cool != kNoCool) u = ev;
            if (!__any_sync(0xffffffffu, u < must)) break;
            if (m) {
                const float p = ring_ld(row_addr(ev));
                const uint32_t v = E.classify(A, p);
                if (v != kFar)
                    E.contact(A, (int)v, p, ev, MULTIDAY ? min(ev / A.combo.spd, A.n_days - 1) * kMaxLevels : 0, (int)((dm >> j) & 1u));
                u = ev + 1;
            }
            // lanes in a position (the ones that just entered included) test their next two rows for the stop / target
            const bool scan = E.cool == kNoCool && u < F;
            {   // no warp vote around this block: a plain divergent branch costs 4 issue slots less per round (measured: C2 1.024 -> 0.998 ms)
                if (scan) {
                    const uint32_t t0 = ((uint32_t)u << 7) & (kRingBytes - 128), t1 = (((uint32_t)u << 7) + 128u) & (kRingBytes - 128);
                    const float p0 = ring_ld(rb + t0), p1 = ring_ld(rb + t1);
                    const bool h0 = (p0 <= E.lo_thr) | (p0 >= E.hi_thr);
                    const bool h1 = (u + 1 < F) & ((p1 <= E.lo_thr) | (p1 >= E.hi_thr));
                    if (h0 | h1) {
                        const float xp = h0 ? p0 : p1;
                        u += h0 ? 0 : 1;
                        const bool xlo = xp <= E.lo_thr;
                        E.do_exit(A, xp, xlo, !xlo, u);
                        ++u;
                    } else u = min(u + 2, F);
                }
            }
        }
    }

    __device__ __forceinline__ void begin(uint32_t sb, uint32_t rb_, int64_t path, bool active) {
        E.reset(sb, path);
        E.cool = kNoDir; E.c_dir = 0u;
        rb = rb_; fm = 0xffffffffu; dm = 0u; u = 0; dcur = 0.f; found = !active;
    }

    __device__ __forceinline__ void end(const ScanArgs& A, bool active) {
        if (active) {
            if (!found) E.c_dir = (uint32_t)A.steps;               // the close never changed: every step is DIR_UNKNOWN
            E.finish(A);
        }
        __syncwarp();
        warp_accumulate(active, E.pnl, E.pnl2, E.mtm, E.c_gate, E.c_ml, E.c_exit, E.c_span, E.c_dir, active && E.is_open());
    }
};

template <bool MULTIDAY, bool SLOW>
__global__ void __launch_bounds__(kLeanThreads, 1) tile_scan_kernel(const __grid_constant__ ScanArgs A) {
    block_setup(A, true);
    lean_setup<MULTIDAY, 32>(A, false);
    __syncthreads();
    TileScan<MULTIDAY, SLOW, 0> T;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t sb = __shfl_sync(0xffffffffu, (uint32_t)__cvta_generic_to_shared(g_cell), 0);
    const uint32_t rb = tile_ring_base(sb) + warp * kRingBytes + lane * 4;
    const int64_t ntiles = (A.paths + 31) / 32;
    const int steps = A.steps;
    const uint32_t pitch4 = (uint32_t)A.pitch * 4u;                 // bytes per row; kChunk rows fit 32 bits (checked on the host)
    for (;;) {
        const int64_t tile = next_tile(A, lane);                    // dynamic: a warp takes the next unclaimed tile when it is done
        if (tile >= ntiles) break;
        const int64_t path = tile * 32 + lane;
        const bool active = path < A.paths;
        T.begin(sb, rb, path, active);
        const char* src = reinterpret_cast<const char*>(A.close + path);   // row of the next chunk to stage
        auto stage = [&](int f0) {                                  // rows [f0, f0 + kChunk) -> their ring slots
            const int nr = min(kChunk, steps - f0);
            if (active) {
                const uint32_t a0 = T.row_addr(f0);
                if (nr == kChunk) {
#pragma unroll
                    for (int r = 0; r < kChunk; ++r) cp_async4(a0 + r * 128, src + (uint32_t)r * pitch4);
                } else {
                    for (int r = 0; r < nr; ++r) cp_async4(a0 + r * 128, src + (uint32_t)r * pitch4);
                }
            }
            src += (size_t)kChunk * pitch4;
        };
        stage(0);
        bool searching = true;
#pragma unroll 1
        for (int f = 0; f < steps; f += kChunk) {
            const int nr = min(kChunk, steps - f);
            cp_async_wait_all();
            const uint32_t a0 = T.row_addr(f);
            if (nr == kChunk && !searching) {
#pragma unroll
                for (int r = 0; r < kChunk; ++r) T.template row<false>(A, ring_ld(a0 + r * 128), f + r);
            } else {
#pragma unroll 1
                for (int r = 0; r < nr; ++r) T.template row<true>(A, ring_ld(a0 + r * 128), f + r);
                searching = __any_sync(0xffffffffu, !T.found);
            }
            const int F = f + nr;
            if (F < steps) stage(F);
            T.rounds(A, F, F < steps ? F - T.kLag : steps);
        }
        T.end(A, active);
    }
    __syncthreads();
    block_flush_agg(A.flat, A.steps);
}

// The same scan with TMA staging (the production form; the cp.async kernel above stays for row pitches that are not a
// multiple of 16 bytes, which a tensor map cannot describe).  tmap: fp32 [steps][paths] with row pitch A.pitch, box
// {32 paths, kChunk rows}; rows and paths past the end arrive as zeros and are never evaluated.
template <bool MULTIDAY, bool SLOW>
__global__ void __launch_bounds__(kLeanThreads, 1) tile_scan_tma_kernel(const __grid_constant__ ScanArgs A,
                                                                         const __grid_constant__ CUtensorMap tmap) {
    block_setup(A, true);
    lean_setup<MULTIDAY, 32>(A, false);
    TileScan<MULTIDAY, SLOW, 0> T;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t sb = __shfl_sync(0xffffffffu, (uint32_t)__cvta_generic_to_shared(g_cell), 0);
    const uint32_t ring = tile_ring_base(sb) + warp * kRingBytes, rb = ring + lane * 4;
    const uint32_t bar = tile_ring_base(sb) + kLeanWarps * kRingBytes + warp * 8;
    if (lane == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t phase = 0;
    const int64_t ntiles = (A.paths + 31) / 32;
    const int steps = A.steps;
    for (;;) {
        const int64_t tile = next_tile(A, lane);                    // dynamic: a warp takes the next unclaimed tile when it is done
        if (tile >= ntiles) break;
        const int64_t path = tile * 32 + lane;
        const bool active = path < A.paths;
        T.begin(sb, rb, path, active);
        const int x0 = (int)(tile * 32);
        auto stage = [&](int f0) {                                  // rows [f0, f0 + kChunk) -> their ring slots
            __syncwarp();                                           // every lane is done reading the slots being replaced
            if (lane == 0) {
                mbar_expect_tx(bar, kChunk * 128);
                tma_load_2d(ring + ((uint32_t)(f0 & (kRing - 1)) << 7), &tmap, x0, f0, bar);
            }
        };
        stage(0);
        bool searching = true;
#pragma unroll 1
        for (int f = 0; f < steps; f += kChunk) {
            const int nr = min(kChunk, steps - f);
            mbar_wait(bar, phase);
            phase ^= 1u;
            const uint32_t a0 = T.row_addr(f);
            if (nr == kChunk && !searching) {
#pragma unroll
                for (int r = 0; r < kChunk; ++r) T.template row<false>(A, ring_ld(a0 + r * 128), f + r);
            } else {
#pragma unroll 1
                for (int r = 0; r < nr; ++r) T.template row<true>(A, ring_ld(a0 + r * 128), f + r);
                searching = __any_sync(0xffffffffu, !T.found);
            }
            const int F = f + nr;
            if (F < steps) stage(F);
            T.rounds(A, F, F < steps ? F - T.kLag : steps);
        }
        T.end(A, active);
    }
    __syncthreads();
    block_flush_agg(A.flat, A.steps);
}

// fused Philox GBM + tile scan: the generator runs in phase 1 with all 32 lanes active (the lockstep fused kernel pays
// Philox + Box-Muller + exp inside a loop whose other half runs at 15 active lanes), bars never leave the SM
template <bool MULTIDAY, bool SLOW>
__global__ void __launch_bounds__(kLeanThreads, 1) tile_fused_gbm_kernel(const __grid_constant__ ScanArgs A, const GbmDev g) {
    block_setup(A, true);
    lean_setup<MULTIDAY, 32>(A, false);
    __syncthreads();
    TileScan<MULTIDAY, SLOW, 1> T;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t sb = __shfl_sync(0xffffffffu, (uint32_t)__cvta_generic_to_shared(g_cell), 0);
    const uint32_t rb = tile_ring_base(sb) + warp * kRingBytes + lane * 4;
    const int64_t ntiles = (A.paths + 31) / 32;
    const int steps = A.steps;
    for (;;) {
        const int64_t tile = next_tile(A, lane);                    // dynamic: a warp takes the next unclaimed tile when it is done
        if (tile >= ntiles) break;
        const int64_t path = tile * 32 + lane;
        const bool active = path < A.paths;
        T.begin(sb, rb, path, active);
        const uint64_t gid = g.path_offset + (uint64_t)path;
        float P = g.s0;
        bool searching = true;
#pragma unroll 1
        for (int f = 0; f < steps; f += kChunk) {
            const int nr = min(kChunk, steps - f);
            const uint32_t a0 = T.row_addr(f);
            float pr[kChunk];
#pragma unroll
            for (int q = 0; q < kChunk / 4; ++q) {                 // same counters as gen_gbm_kernel: group = step / 4
                float z[4];
                gbm_normals(g, gid, (uint32_t)((f >> 2) + q), z);
#pragma unroll
                for (int w = 0; w < 4; ++w) { P = gbm_advance(g, P, z[w]); pr[q * 4 + w] = g.cents ? round_cents(P) : P; }
            }
#pragma unroll
            for (int r = 0; r < kChunk; ++r) ring_st(a0 + r * 128, pr[r]);
            if (nr == kChunk && !searching) {
#pragma unroll
                for (int r = 0; r < kChunk; ++r) T.template row<false>(A, pr[r], f + r);
            } else {
#pragma unroll
                for (int r = 0; r < kChunk; ++r) if (r < nr) T.template row<true>(A, pr[r], f + r);
                searching = __any_sync(0xffffffffu, !T.found);
            }
            const int F = f + nr;
            T.rounds(A, F, F < steps ? F - T.kLag : steps);
        }
        T.end(A, active);
    }
    __syncthreads();
    block_flush_agg(A.flat, A.steps);
}

}  // namespace qmmx
