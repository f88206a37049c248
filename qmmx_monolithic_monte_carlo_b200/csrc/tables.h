// Host/device table layouts shared by the table builder (host) and the kernels (device).
//
// Every per-tick decision of the reference's gate stack that depends only on the price p — which level is
// nearest and whether it is within CONTACT_PROX (`min(levels, key=|L.price - p|)`, first minimal element,
// qmmx_monolithic.py:1543-1545), whether p is still inside another level's contact window (latch release,
// M:1572-1576), whether `compute_confidence(...) < Q_MIN_PROB` (M:1415-1427, M:1627-1641) — is, for an
// fp32-representable p evaluated in float64, a monotone or piecewise-constant function of p.  The host
// evaluates those float64 predicates EXACTLY (same IEEE operations as CPython, no FMA contraction) and
// bisects over the ordered set of fp32 values, so the kernels replace the O(L) float64 work by fp32
// compares against exact thresholds and one shared-memory byte lookup.
#pragma once
#include <cstddef>
#include <cstdint>

namespace qmmx {

constexpr int kThreads = 256;           // threads per CTA; one thread walks one path
constexpr int kMaxLevels = 64;
constexpr int kMaxCells = 32768;
constexpr int kMaxZones = 640;          // runs of the (finer) class partition as well as of the level partition
constexpr int kMaxClasses = 254;        // class ids 0 .. ncls-1, single-boundary records ncls .. 0xFD, then kMixed, kFar
constexpr int kMaxTables = 8;           // distinct (CONTACT_PROX, Q_MIN_PROB) pairs across the combos of one run
constexpr int kMaxDays = 64;
constexpr uint8_t kFar = 0xFF;
constexpr uint8_t kMixed = 0xFE;          // cell straddles two or more zone boundaries: exact zone search
constexpr uint8_t kMixBase = 0x40;        // kMixBase + m: cell straddles exactly ONE boundary, record m resolves it
constexpr int kMaxMixed = kMixed - kMixBase;

constexpr uint32_t kLvSolid = 1u, kLvConfl06 = 2u, kLvConfl015 = 4u;

struct DevLevel {                       // 64 bytes
    double price;
    float wlo, whi;                     // |price - p| <= CONTACT_PROX  <=>  wlo <= p <= whi       (fp32 p)
    float cpass[3][2];                  // touch bucket b (tc<=1, tc==2, tc>=3): conf >= Q_MIN_PROB <=> [0] <= p <= [1]
    uint32_t cand_lo, cand_hi;          // levels j != k whose contact window overlaps this level's window
    uint32_t flags;                     // kLvSolid | kLvConfl06 (M:3069-3070) | kLvConfl015 (M:1885-1886)
    uint32_t pad[3];
};
static_assert(sizeof(DevLevel) == 64, "DevLevel layout");

struct alignas(16) DevTables {          // arrays of DevTables: cell[] is copied with 16-byte loads
    int32_t L, Z, ncell, shift;         // cell = clamp((bits(p) - obase) >> shift, 0, ncell-1)
    int32_t obase, pad0, pad1, pad2;
    DevLevel lev[kMaxLevels];
    float zhi[kMaxZones];               // inclusive fp32 upper bound of zone z; zone z = (zhi[z-1], zhi[z]]
    uint8_t zlev[kMaxZones];            // level touched in zone z, or kFar
    uint8_t cell[kMaxCells];            // level index / kFar / kMixBase + m / kMixed per grid cell
    // single-boundary cells: p <= mixb[m] ? mixlo[m] : mixhi[m]  (level index or kFar)
    int32_t nmixed, pad3[3];
    // far zones adjoining the grid (prices that leave the grid land in an end cell): guard[0] < p <= guard[1] or
    // guard[2] < p <= guard[3]  =>  far, without the zone search; an empty interval is (+inf, -inf]
    float guard[4];
    float mixb[kMaxMixed];
    uint8_t mixlo[kMaxMixed], mixhi[kMaxMixed];
    // ---- class partition (scan_lean.cuh / scan_tile.cuh): a class = (level touched, set of levels whose contact window
    // holds p).  The second member is what the latch release of M:1572-1576 asks per latched neighbour; as a function of
    // p it is piecewise constant, so one record per class replaces the per-neighbour window tests.  Same grid geometry as
    // cell[]; ids: class c < ncls, single-boundary record ncls + m (p <= fmixb[id] ? fmixlo[id] : fmixhi[id], class ids
    // or kFar), kMixed (zone search over fzhi / fzcls), kFar.  ncls < 0: the level set needs more than kMaxClasses
    // classes or kMaxZones runs and the first-generation kernels run instead.
    int32_t FZ, ncls, fnmixed, pad4;
    float fzhi[kMaxZones];
    uint8_t fzcls[kMaxZones];
    alignas(16) uint8_t fcell[kMaxCells];
    uint32_t cls_lev[256];              // level touched (class ids only)
    uint32_t cls_mask_lo[256], cls_mask_hi[256];   // levels whose window holds p, the touched one included
    float fmixb[256];
    uint8_t fmixlo[256], fmixhi[256];
    // prices that leave the grid (both end cells are kMixed when levels lie outside it): p < grid_lo or p > grid_hi is far
    // unless out_lo[j] <= p <= out_hi[j] for one of the n_out class runs outside the grid (class out_cls[j]).
    // n_out < 0: more than kMaxOut such runs -- the zone search answers.
    int32_t n_out, pad5[3];
    float grid_lo, grid_hi;             // first fp32 value of cell 1, last fp32 value of cell ncell - 2
    float out_lo[4], out_hi[4];
    uint32_t out_cls[4];
    uint32_t pad6[2];
};
constexpr int kMaxOut = 4;
static_assert(sizeof(DevTables) % 16 == 0, "DevTables stride");
static_assert(offsetof(DevTables, cell) % 16 == 0 && offsetof(DevTables, fcell) % 16 == 0, "cell tables are copied with 16-byte loads");

struct DevCombo {                       // one parameter combination (qmmx_params, preprocessed)
    double sp, tp, qmin, prox, pmax, wr, wm;
    double vstrong, vprox;              // VETO_VOL_STRONG, VETO_PROX (M:1381-1382)
    int32_t cool_steps;                 // max(1, ceil(cooldown_s / 60)): first eligible step after an exit at step i is i + cool_steps
    int32_t overtouch;
    int32_t spd;                        // steps per day
    int32_t table;
    uint32_t flags;
    uint32_t pad[3];
};

// grid cell of a price: integer arithmetic on the IEEE bit pattern (monotone for p >= 0; negative p -> cell 0)
#ifdef __CUDACC__
__host__ __device__
#endif
inline int cell_of_bits(int32_t bits, int32_t obase, int32_t shift, int32_t ncell) {
    const int32_t o = bits > obase ? bits : obase;
    const int32_t c = (o - obase) >> shift;
    return c < ncell - 1 ? c : ncell - 1;
}

}  // namespace qmmx
