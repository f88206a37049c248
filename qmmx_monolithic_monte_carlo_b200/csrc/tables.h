// Host/device table layouts shared by the table builder (host) and the kernels (device).
//
// The nearest-level search of the reference (`min(levels, key=|L.price - p|)`, first minimal element,
// qmmx_monolithic.py:1543) and its proximity test (`dist > CONTACT_PROX`, M:1545) depend only on the
// price p.  For fp32-representable p evaluated in float64 the answer is a piecewise-constant function of
// p: the host builds that function EXACTLY (by bisection over the ordered set of fp32 values against the
// brute-force float64 predicate) as a list of zones, plus a uniform grid over the price axis whose cells
// say "every price in this cell is far from all levels" / "touches level k" / "mixed, consult the zones".
// The kernels then classify a price with one shared-memory byte load instead of an O(L) float64 search.
#pragma once
#include <cstdint>

namespace qmmx {

constexpr int kThreads = 256;           // threads per CTA = paths per tile
constexpr int kMaxLevels = 64;
constexpr int kMaxCells = 16384;
constexpr int kMaxZones = 256;
constexpr int kMaxTables = 8;           // distinct CONTACT_PROX values across the combos of one run
constexpr uint8_t kFar = 0xFF;
constexpr uint8_t kMixed = 0xFE;

struct DevTables {
    int32_t L, Z, ncell, pad0;
    float glo, inv_w, pad1, pad2;
    double price[kMaxLevels];
    uint8_t solid[kMaxLevels];
    uint8_t confl06[kMaxLevels];    // #{j : |price_j - price_k| <= 0.6} > 1   (M:3069-3070)
    uint8_t confl015[kMaxLevels];   // #{j : |price_j - price_k| <= 0.15} >= 2 (M:1885-1886)
    float zhi[kMaxZones];           // inclusive fp32 upper bound of zone z; zone z = (zhi[z-1], zhi[z]]
    uint8_t zlev[kMaxZones];        // level index touched in zone z, or kFar
    uint8_t cell[kMaxCells];        // level index / kFar / kMixed per grid cell
};

struct DevCombo {
    double prox, pmax, sp, tp, qmin, wr, wm, vstrong, vprox;
    int32_t cool_steps;             // ceil(cooldown_s / 60): now_ms < until_ms  <=>  step < exit_step + cool_steps
    int32_t overtouch;
    int32_t spd;
    int32_t table;
    uint32_t flags;
    uint32_t pad[3];
};

// grid cell of a price; identical IEEE single operations on host and device (no FMA contraction)
#ifdef __CUDACC__
__host__ __device__
#endif
inline int cell_of(float p, float glo, float inv_w, int ncell) {
#ifdef __CUDA_ARCH__
    const float t = __fmul_rn(__fsub_rn(p, glo), inv_w);
    int c = __float2int_rd(t);
    c = c < 0 ? 0 : c;
    return c > ncell - 1 ? ncell - 1 : c;
#else
    const volatile float d = p - glo;
    const volatile float t = d * inv_w;
    if (!(t > 0.0f)) return 0;                       // also NaN
    if (t >= (float)(ncell - 1)) return ncell - 1;
    return (int)t;                                   // t > 0: truncation == floor
#endif
}

}  // namespace qmmx
