// Host-side builder of the exact price -> {far | level k} classification tables (see tables.h).
// Compiled with -ffp-contract=off: every double/float operation below is a plain IEEE operation, i.e.
// the same arithmetic CPython performs in qmmx_monolithic.py:1543-1545.
#include "tables_host.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

namespace qmmx {

namespace {

// monotone bijection between finite floats and int32 ordinals (-0.0 -> -1, +0.0 -> 0)
inline int32_t f2o(float f) {
    int32_t i;
    std::memcpy(&i, &f, 4);
    return i ^ ((i >> 31) & 0x7fffffff);
}
inline float o2f(int32_t o) {
    int32_t i = o ^ ((o >> 31) & 0x7fffffff);
    float f;
    std::memcpy(&f, &i, 4);
    return f;
}

struct Classifier {
    int L;
    const double* price;
    double prox;
    // brute force, first minimal element wins (Python min, M:1543); returns level index or -1 (too far)
    int operator()(float pf) const {
        const double p = (double)pf;
        int k = 0;
        double dist = std::fabs(price[0] - p);
        for (int j = 1; j < L; ++j) {
            const double dj = std::fabs(price[j] - p);
            if (dj < dist) { dist = dj; k = j; }
        }
        return dist > prox ? -1 : k;
    }
};

struct Zone { int32_t hi; int lev; };

void refine(const Classifier& z, int64_t lo, int zlo, int64_t hi, int zhi, std::vector<Zone>& out) {
    // invariant: z(lo) = zlo, z(hi) = zhi, lo < hi.  Emits (last ordinal of a run, its value) for every
    // change found in (lo, hi]; runs are assumed constant between probes that agree.
    if (zlo == zhi) return;
    if (hi == lo + 1) { out.push_back({(int32_t)lo, zlo}); return; }
    const int64_t mid = lo + (hi - lo) / 2;
    const int zm = z(o2f((int32_t)mid));
    refine(z, lo, zlo, mid, zm, out);
    refine(z, mid, zm, hi, zhi, out);
}

}  // namespace

int build_tables(int L, const double* price, const uint8_t* solid, double prox, DevTables* T, std::string* err) {
    std::memset(T, 0, sizeof(*T));
    T->L = L;
    for (int j = 0; j < L; ++j) {
        T->price[j] = price[j];
        T->solid[j] = solid[j] ? 1 : 0;
        int c06 = 0, c015 = 0;
        for (int i = 0; i < L; ++i) {
            if (std::fabs(price[i] - price[j]) <= 0.6) ++c06;
            if (std::fabs(price[i] - price[j]) <= 0.15) ++c015;
        }
        T->confl06[j] = c06 > 1;
        T->confl015[j] = c015 >= 2;
    }
    if (L == 0) {
        T->Z = 1; T->zhi[0] = FLT_MAX; T->zlev[0] = kFar;
        T->ncell = 1; T->glo = 0.0f; T->inv_w = 1.0f; T->cell[0] = kFar;
        return 0;
    }
    const Classifier z{L, price, prox};
    const int64_t omin = f2o(-FLT_MAX), omax = f2o(FLT_MAX);

    // probe ordinals: a few fp32 neighbours of every real-valued change candidate
    std::vector<int64_t> probes{omin, omax};
    auto add = [&](double x) {
        if (!(std::fabs(x) < (double)FLT_MAX)) return;
        const int64_t o = f2o((float)x);
        for (int64_t d = -3; d <= 3; ++d) {
            const int64_t q = o + d;
            if (q >= omin && q <= omax) probes.push_back(q);
        }
    };
    for (int j = 0; j < L; ++j) {
        add(price[j]); add(price[j] - prox); add(price[j] + prox);
        for (int i = j + 1; i < L; ++i) add(0.5 * (price[i] + price[j]));
    }
    std::sort(probes.begin(), probes.end());
    probes.erase(std::unique(probes.begin(), probes.end()), probes.end());

    std::vector<Zone> zones;
    int64_t lo = probes[0];
    int zlo = z(o2f((int32_t)lo));
    for (size_t i = 1; i < probes.size(); ++i) {
        const int64_t hi = probes[i];
        const int zhi = z(o2f((int32_t)hi));
        refine(z, lo, zlo, hi, zhi, zones);
        lo = hi; zlo = zhi;
    }
    zones.push_back({(int32_t)omax, zlo});
    if ((int)zones.size() > kMaxZones) {
        if (err) *err = "level set produces more than 256 price zones";
        return -1;
    }
    T->Z = (int)zones.size();
    for (int i = 0; i < T->Z; ++i) {
        T->zhi[i] = o2f(zones[i].hi);
        T->zlev[i] = zones[i].lev < 0 ? kFar : (uint8_t)zones[i].lev;
    }
    // +inf and NaN never compare <= zhi: give the last zone an infinite bound so the device search ends
    T->zhi[T->Z - 1] = INFINITY;

    // uniform grid over [min level - prox - margin, max level + prox + margin)
    double pmin = price[0], pmax = price[0];
    for (int j = 1; j < L; ++j) { pmin = std::min(pmin, price[j]); pmax = std::max(pmax, price[j]); }
    const double target_w = 1.0 / 64.0;
    const double glo = pmin - prox - 4 * target_w, ghi = pmax + prox + 4 * target_w;
    int ncell = (int)std::ceil((ghi - glo) / target_w);
    ncell = std::max(1, std::min(kMaxCells, ncell));
    T->ncell = ncell;
    T->glo = (float)glo;
    T->inv_w = (float)((double)ncell / (ghi - glo));

    auto zone_index = [&](int64_t o) {   // first zone whose upper ordinal >= o
        int a = 0, b = (int)zones.size() - 1;
        while (a < b) {
            const int m = (a + b) / 2;
            if ((int64_t)zones[m].hi >= o) b = m; else a = m + 1;
        }
        return a;
    };
    int64_t cell_lo = omin;
    for (int c = 0; c < ncell; ++c) {
        // last ordinal whose cell index is <= c (cell_of is monotone non-decreasing in p)
        int64_t a = cell_lo, b = omax;
        if (cell_of(o2f((int32_t)cell_lo), T->glo, T->inv_w, ncell) > c) { T->cell[c] = kMixed; continue; }  // empty
        if (c == ncell - 1) a = omax;
        else {
            while (a < b) {
                const int64_t m = a + (b - a + 1) / 2;
                if (cell_of(o2f((int32_t)m), T->glo, T->inv_w, ncell) <= c) a = m; else b = m - 1;
            }
        }
        const int64_t cell_hi = a;
        const int z0 = zone_index(cell_lo), z1 = zone_index(cell_hi);
        T->cell[c] = (z0 == z1) ? T->zlev[z0] : kMixed;
        cell_lo = cell_hi + 1;
        if (cell_lo > omax) { for (int r = c + 1; r < ncell; ++r) T->cell[r] = kMixed; break; }
    }
    return 0;
}

int classify_bruteforce(int L, const double* price, double prox, float p) {
    if (L == 0) return -1;
    return Classifier{L, price, prox}(p);
}

int classify_tables(const DevTables* T, float p) {
    uint8_t v = T->cell[cell_of(p, T->glo, T->inv_w, T->ncell)];
    if (v == kMixed) {
        int a = 0, b = T->Z - 1;
        while (a < b) {
            const int m = (a + b) / 2;
            if (p <= T->zhi[m]) b = m; else a = m + 1;
        }
        v = T->zlev[a];
    }
    return v == kFar ? -1 : (int)v;
}

}  // namespace qmmx
