// Host-side builder of the exact price-classification tables (see tables.h).
// Compiled with -ffp-contract=off: every double/float operation below is a plain IEEE operation, i.e.
// the same arithmetic CPython performs in qmmx_monolithic.py:1415-1427 and M:1543-1576.
#include "tables_host.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <functional>
#include <vector>

namespace qmmx {

namespace {

// monotone bijection between finite floats and int32 ordinals (-0.0 -> -1, +0.0 -> 0); for p >= 0 the
// ordinal is the IEEE bit pattern, which is what the kernels use
inline int32_t f2o(float f) {
    int32_t i;
    std::memcpy(&i, &f, 4);
    return i ^ ((i >> 31) & 0x7fffffff);
}
inline float o2f(int64_t o64) {
    const int32_t o = (int32_t)o64;
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

using ClassFn = std::function<int(float)>;

void refine(const ClassFn& z, int64_t lo, int zlo, int64_t hi, int zhi, std::vector<Zone>& out) {
    // invariant: z(lo) = zlo, z(hi) = zhi, lo < hi.  Emits (last ordinal of a run, its value) for every
    // change found in (lo, hi]; runs are assumed constant between probes that agree.
    if (zlo == zhi) return;
    if (hi == lo + 1) { out.push_back({(int32_t)lo, zlo}); return; }
    const int64_t mid = lo + (hi - lo) / 2;
    const int zm = z(o2f(mid));
    refine(z, lo, zlo, mid, zm, out);
    refine(z, mid, zm, hi, zhi, out);
}

using Pred = std::function<bool(float)>;

// The set {p fp32 : ok(p)} is an interval around `center` (ok is monotone on each side of it).
// Returns false when it is empty.
bool true_interval(const Pred& ok, double center, float* lo_out, float* hi_out) {
    const int64_t omin = f2o(-FLT_MAX), omax = f2o(FLT_MAX);
    if (!(std::fabs(center) < (double)FLT_MAX)) return false;
    const int64_t oc0 = f2o((float)center);
    int64_t oc = INT64_MIN;
    for (int64_t d = -2; d <= 2; ++d) {
        const int64_t q = oc0 + d;
        if (q >= omin && q <= omax && ok(o2f(q))) { oc = q; break; }
    }
    if (oc == INT64_MIN) return false;
    int64_t a = omin, b = oc;            // first true in [omin, oc]
    while (a < b) {
        const int64_t m = a + (b - a) / 2;
        if (ok(o2f(m))) b = m; else a = m + 1;
    }
    *lo_out = o2f(a);
    a = oc; b = omax;                    // last true in [oc, omax]
    while (a < b) {
        const int64_t m = a + (b - a + 1) / 2;
        if (ok(o2f(m))) a = m; else b = m - 1;
    }
    *hi_out = o2f(a);
    return true;
}

}  // namespace

double confidence_f64(double dist, bool solid, int tc, double prox) {   // M:1415-1427
    double base = 1.0 - (dist / (0.0001 > prox ? 0.0001 : prox));
    if (!(base > 0.0)) base = 0.0;
    base += solid ? 0.08 : 0.02;
    if (tc <= 1) base += 0.10;
    else if (tc == 2) base -= 0.08;
    else base -= (0.08 * 2);
    base += 0.03;
    const double m = base < 1.0 ? base : 1.0;
    return m > 0.0 ? m : 0.0;
}

int build_tables(int L, const double* price, const uint8_t* solid, double prox, double qmin, DevTables* T,
                 std::string* err) {
    std::memset(T, 0, sizeof(*T));
    T->L = L;
    const float finf = INFINITY;
    for (int j = 0; j < L; ++j) {
        DevLevel& D = T->lev[j];
        D.price = price[j];
        int c06 = 0, c015 = 0;
        for (int i = 0; i < L; ++i) {
            if (std::fabs(price[i] - price[j]) <= 0.6) ++c06;
            if (std::fabs(price[i] - price[j]) <= 0.15) ++c015;
        }
        D.flags = (solid[j] ? kLvSolid : 0u) | (c06 > 1 ? kLvConfl06 : 0u) | (c015 >= 2 ? kLvConfl015 : 0u);
        const double lp = price[j];
        // contact window: a latched level stays latched while |lp - p| <= prox (M:1574 releases on >)
        D.wlo = finf; D.whi = -finf;
        true_interval([&](float pf) { return !(std::fabs(lp - (double)pf) > prox); }, lp, &D.wlo, &D.whi);
        // confidence gate per touch bucket: pass <=> !(conf < qmin)  (M:1628)
        for (int b = 0; b < 3; ++b) {
            D.cpass[b][0] = finf; D.cpass[b][1] = -finf;
            const bool sol = solid[j] != 0;
            true_interval([&](float pf) {
                return !(confidence_f64(std::fabs(lp - (double)pf), sol, b + 1, prox) < qmin);
            }, lp, &D.cpass[b][0], &D.cpass[b][1]);
        }
    }
    for (int k = 0; k < L; ++k) {
        uint64_t cand = 0;
        for (int j = 0; j < L; ++j) {
            if (j == k) continue;
            const DevLevel &A = T->lev[k], &B = T->lev[j];
            if (A.wlo <= A.whi && B.wlo <= B.whi && B.wlo <= A.whi && A.wlo <= B.whi) cand |= 1ull << j;
        }
        T->lev[k].cand_lo = (uint32_t)cand;
        T->lev[k].cand_hi = (uint32_t)(cand >> 32);
    }
    if (L == 0) {
        T->Z = 1; T->zhi[0] = finf; T->zlev[0] = kFar;
        T->ncell = 1; T->shift = 0; T->obase = 0; T->cell[0] = kFar;
        T->FZ = 1; T->fzhi[0] = finf; T->fzcls[0] = kFar; T->fcell[0] = kFar; T->ncls = 0;
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

    auto partition = [&](const ClassFn& cls, std::vector<Zone>& out) {
        int64_t lo = probes[0];
        int zlo = cls(o2f(lo));
        for (size_t i = 1; i < probes.size(); ++i) {
            const int64_t hi = probes[i];
            const int zhi = cls(o2f(hi));
            refine(cls, lo, zlo, hi, zhi, out);
            lo = hi; zlo = zhi;
        }
        out.push_back({(int32_t)omax, zlo});
    };
    std::vector<Zone> zones;
    partition([&](float pf) { return z(pf); }, zones);
    if ((int)zones.size() > kMaxZones) {
        if (err) *err = "level set produces too many price zones";
        return -1;
    }
    T->Z = (int)zones.size();
    for (int i = 0; i < T->Z; ++i) {
        T->zhi[i] = o2f(zones[i].hi);
        T->zlev[i] = zones[i].lev < 0 ? kFar : (uint8_t)zones[i].lev;
    }
    // +inf and NaN never compare <= a finite bound: the last zone is unbounded so the device search ends there
    T->zhi[T->Z - 1] = finf;

    // class partition: (level touched, levels whose window holds p).  Window membership is the device's own test
    // wlo <= p <= whi on the exact fp32 bounds found above, so a class mask IS what the per-neighbour loop would keep.
    std::vector<std::pair<int, uint64_t>> classes;
    auto class_of = [&](float pf) -> int {
        const int k = z(pf);
        if (k < 0) return -1;
        uint64_t mask = 0;
        for (int j = 0; j < L; ++j)
            if (pf >= T->lev[j].wlo && pf <= T->lev[j].whi) mask |= 1ull << j;
        const std::pair<int, uint64_t> key(k, mask);
        for (size_t c = 0; c < classes.size(); ++c) if (classes[c] == key) return (int)c;
        classes.push_back(key);
        return (int)classes.size() - 1;
    };
    std::vector<Zone> fzones;
    partition(class_of, fzones);
    const bool fine_ok = (int)classes.size() <= kMaxClasses && (int)fzones.size() <= kMaxZones;
    T->ncls = fine_ok ? (int)classes.size() : -1;
    T->FZ = fine_ok ? (int)fzones.size() : 0;
    if (fine_ok) {
        for (int i = 0; i < T->FZ; ++i) {
            T->fzhi[i] = o2f(fzones[i].hi);
            T->fzcls[i] = fzones[i].lev < 0 ? kFar : (uint8_t)fzones[i].lev;
        }
        T->fzhi[T->FZ - 1] = finf;
        for (size_t c = 0; c < classes.size(); ++c) {
            T->cls_lev[c] = (uint32_t)classes[c].first;
            T->cls_mask_lo[c] = (uint32_t)classes[c].second;
            T->cls_mask_hi[c] = (uint32_t)(classes[c].second >> 32);
        }
    }

    // ---- grid over fp32 ordinals: full coverage of all windows when it fits at <= 64 ulps per cell,
    //      else the densest cluster of levels at 16 ulps per cell (prices outside resolve through the zones)
    std::vector<int> order(L);
    for (int j = 0; j < L; ++j) order[j] = j;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return price[a] < price[b]; });
    auto ord_of = [&](double x) -> int64_t {
        if (x >= (double)FLT_MAX) return omax;
        if (x <= 0.0) return 0;
        return std::max<int64_t>(0, f2o((float)x));
    };
    const int64_t margin = 64;
    int64_t o_lo = std::max<int64_t>(0, ord_of(price[order[0]] - prox) - margin);
    int64_t o_hi = ord_of(price[order[L - 1]] + prox) + margin;
    int shift = -1;
    for (int s = 2; s <= 6; ++s)
        if (((o_hi - o_lo) >> s) + 1 <= kMaxCells) { shift = s; break; }
    if (shift < 0) {
        shift = 4;
        const int64_t cap = ((int64_t)kMaxCells << shift) - 2 * margin;
        int best = 0, best_n = 0;
        for (int i = 0; i < L; ++i) {
            const int64_t a = ord_of(price[order[i]] - prox);
            int n = 0;
            for (int j = i; j < L; ++j) { if (ord_of(price[order[j]] + prox) - a <= cap) ++n; else break; }
            if (n > best_n) { best_n = n; best = i; }
        }
        // centre the cluster in the grid: the spare cells on both sides absorb price excursions
        const int64_t c_lo = ord_of(price[order[best]] - prox), c_hi = ord_of(price[order[best + best_n - 1]] + prox);
        const int64_t spare = std::max<int64_t>(0, ((int64_t)kMaxCells << shift) - (c_hi - c_lo) - 2 * margin);
        o_lo = std::max<int64_t>(0, c_lo - margin - spare / 2);
        o_hi = std::min<int64_t>(o_lo + ((int64_t)kMaxCells << shift) - 1, omax);
    }
    const int ncell = (int)std::min<int64_t>(kMaxCells, ((o_hi - o_lo) >> shift) + 1);
    T->ncell = ncell; T->shift = shift; T->obase = (int32_t)o_lo;

    auto zone_index_in = [&](const std::vector<Zone>& zs, int64_t o) {   // first zone whose upper ordinal >= o
        int a = 0, b = (int)zs.size() - 1;
        while (a < b) {
            const int m = (a + b) / 2;
            if ((int64_t)zs[m].hi >= o) b = m; else a = m + 1;
        }
        return a;
    };
    auto zone_index = [&](int64_t o) { return zone_index_in(zones, o); };
    for (int c = 0; c < ncell; ++c) {
        const int64_t c_lo = (c == 0) ? omin : o_lo + ((int64_t)c << shift);
        const int64_t c_hi = (c == ncell - 1) ? omax : o_lo + (((int64_t)c + 1) << shift) - 1;
        const int z0 = zone_index(std::min(c_lo, omax)), z1 = zone_index(std::min(c_hi, omax));
        if (z0 == z1) T->cell[c] = T->zlev[z0];
        else if (z1 == z0 + 1 && T->nmixed < kMaxMixed) {
            const int m = T->nmixed++;
            T->mixb[m] = T->zhi[z0]; T->mixlo[m] = T->zlev[z0]; T->mixhi[m] = T->zlev[z1];
            T->cell[c] = (uint8_t)(kMixBase + m);
        } else T->cell[c] = kMixed;
        if (fine_ok) {
            const int f0 = zone_index_in(fzones, std::min(c_lo, omax)), f1 = zone_index_in(fzones, std::min(c_hi, omax));
            if (f0 == f1) T->fcell[c] = T->fzcls[f0];
            else if (f1 == f0 + 1 && T->ncls + T->fnmixed < (int)kMixed) {
                const int id = T->ncls + T->fnmixed++;
                T->fmixb[id] = T->fzhi[f0]; T->fmixlo[id] = T->fzcls[f0]; T->fmixhi[id] = T->fzcls[f1];
                T->fcell[c] = (uint8_t)id;
            } else T->fcell[c] = kMixed;
        }
    }
    // guard intervals: the far zones that contain the ordinals just outside the grid
    {
        const float finf2 = INFINITY;
        T->guard[0] = finf2; T->guard[1] = -finf2; T->guard[2] = finf2; T->guard[3] = -finf2;
        const int64_t below = o_lo - 1, above = o_lo + ((int64_t)ncell << shift);
        if (below >= omin) {
            const int z = zone_index(below);
            if (T->zlev[z] == kFar) { T->guard[0] = z > 0 ? T->zhi[z - 1] : -finf2; T->guard[1] = T->zhi[z]; }
        }
        if (above <= omax) {
            const int z = zone_index(above);
            if (T->zlev[z] == kFar) { T->guard[2] = z > 0 ? T->zhi[z - 1] : -finf2; T->guard[3] = T->zhi[z]; }
        }
    }
    // scan_lean.cuh clamps with one unsigned min, which sends prices below the grid to the LAST cell: both end cells
    // must then resolve the same way (both far in the full-coverage case; else through the exact zone search)
    if (T->cell[0] != T->cell[ncell - 1]) { T->cell[0] = kMixed; T->cell[ncell - 1] = kMixed; }
    if (fine_ok && T->fcell[0] != T->fcell[ncell - 1]) { T->fcell[0] = kMixed; T->fcell[ncell - 1] = kMixed; }
    // class runs outside the interior cells (levels the grid does not cover): a short list instead of the zone search
    T->n_out = -1;
    if (fine_ok && ncell >= 3 && o_lo >= 0) {
        const int64_t g_lo = o_lo + ((int64_t)1 << shift), g_hi = o_lo + (((int64_t)ncell - 1) << shift) - 1;   // interior ordinals
        int n = 0;
        bool fits = true;
        int64_t run_lo = omin;
        for (size_t zi = 0; zi < fzones.size() && fits; ++zi) {
            const int64_t run_hi = fzones[zi].hi;
            if (fzones[zi].lev >= 0) {
                const int64_t parts[2][2] = {{run_lo, std::min(run_hi, g_lo - 1)}, {std::max(run_lo, g_hi + 1), run_hi}};
                for (int s2 = 0; s2 < 2 && fits; ++s2) {
                    if (parts[s2][0] > parts[s2][1]) continue;
                    if (n == kMaxOut) { fits = false; break; }
                    T->out_lo[n] = o2f(parts[s2][0]); T->out_hi[n] = o2f(parts[s2][1]); T->out_cls[n] = (uint32_t)fzones[zi].lev;
                    ++n;
                }
            }
            run_lo = run_hi + 1;
        }
        if (fits) {
            T->n_out = n;
            T->grid_lo = o2f(g_lo); T->grid_hi = o2f(g_hi);
            for (int j = n; j < kMaxOut; ++j) { T->out_lo[j] = finf; T->out_hi[j] = -finf; T->out_cls[j] = kFar; }
        }
    }
    return 0;
}

// sklearn-form LR gate (M:1454-1466) evaluated on the host for every (day, level, side, touch count):
// x = [lvl_type(=1), |level - stop|, touch_count, dir_up];  prob = 1 / (1 + exp(-(b + w.x))), accumulated left
// to right in float64 exactly like oracle/headless.py:lr_prob.  mask bit (side*16 + tc) = gate FAILS.
void build_ml_tables(int L, const double* price, double sp, double qmin, const double* lr, int n_lr,
                     uint32_t* mask, double* prob) {
    for (int day = 0; day < n_lr; ++day) {
        const double* m = lr + 5 * day;
        for (int k = 0; k < kMaxLevels; ++k) {
            uint32_t bits = 0;
            for (int side = 0; side < 2; ++side) {
                for (int tc = 0; tc < 16; ++tc) {
                    double pr = 0.0;
                    if (k < L) {
                        const double lp = price[k];
                        const double stop = side == 0 ? lp - sp : lp + sp;      // M:1669-1674
                        const double distf = std::fabs(lp - stop);
                        double z = m[4];
                        z = z + m[0] * 1.0;
                        z = z + m[1] * distf;
                        z = z + m[2] * (double)tc;
                        z = z + m[3] * (side == 0 ? 1.0 : 0.0);
                        pr = 1.0 / (1.0 + std::exp(-z));
                        if (!(pr >= qmin)) bits |= 1u << (side * 16 + tc);
                    }
                    if (prob) prob[(((size_t)day * kMaxLevels + k) * 2 + side) * 16 + tc] = pr;
                }
            }
            mask[(size_t)day * kMaxLevels + k] = bits;
        }
    }
}

// smallest float64 z with sigmoid_policy(z) >= target (M:20-26, libm exp as CPython uses it); 0 < target < 1
double sigmoid_threshold(double target) {
    auto sig = [](double x) { return 1.0 / (1.0 + std::exp(-x)); };
    double lo = -50.0, hi = 50.0;              // sig(lo) < target <= sig(hi)
    for (int it = 0; it < 200; ++it) {
        const double mid = lo + (hi - lo) / 2;
        if (mid <= lo || mid >= hi) break;
        if (sig(mid) >= target) hi = mid; else lo = mid;
    }
    // walk to the exact boundary over adjacent doubles
    while (sig(std::nextafter(hi, -INFINITY)) >= target) hi = std::nextafter(hi, -INFINITY);
    return hi;
}

int classify_bruteforce(int L, const double* price, double prox, float p) {
    if (L == 0) return -1;
    return Classifier{L, price, prox}(p);
}

int classify_tables(const DevTables* T, float p) {
    int32_t bits;
    std::memcpy(&bits, &p, 4);
    uint8_t v = T->cell[cell_of_bits(bits, T->obase, T->shift, T->ncell)];
    if (v >= kMixBase && v < kMixed) {
        const int m = v - kMixBase;
        v = p <= T->mixb[m] ? T->mixlo[m] : T->mixhi[m];
    } else if (v == kMixed && ((p > T->guard[0] && p <= T->guard[1]) || (p > T->guard[2] && p <= T->guard[3]))) {
        v = kFar;
    } else if (v == kMixed) {
        int a = 0, b = T->Z - 1;
        while (a < b) {
            const int m = (a + b) >> 1;
            if (p <= T->zhi[m]) b = m; else a = m + 1;
        }
        v = T->zlev[a];
    }
    return v == kFar ? -1 : (int)v;
}

// class lookup exactly as the lean / tile kernels do it: level touched (-1 = far) and the window mask
int classify_tables_fine(const DevTables* T, float p, uint64_t* mask) {
    *mask = 0;
    if (T->ncls < 0) return -2;
    int32_t bits;
    std::memcpy(&bits, &p, 4);
    uint8_t v = T->fcell[cell_of_bits(bits, T->obase, T->shift, T->ncell)];
    if (v >= T->ncls && v < kMixed) v = p <= T->fmixb[v] ? T->fmixlo[v] : T->fmixhi[v];
    else if (v == kMixed && ((p > T->guard[0] && p <= T->guard[1]) || (p > T->guard[2] && p <= T->guard[3]))) v = kFar;
    else if (v == kMixed && T->n_out >= 0 && (p < T->grid_lo || p > T->grid_hi)) {
        v = kFar;
        for (int j = 0; j < T->n_out; ++j) if (p >= T->out_lo[j] && p <= T->out_hi[j]) v = (uint8_t)T->out_cls[j];
    } else if (v == kMixed) {
        int a = 0, b = T->FZ - 1;
        while (a < b) {
            const int m = (a + b) >> 1;
            if (p <= T->fzhi[m]) b = m; else a = m + 1;
        }
        v = T->fzcls[a];
    }
    if (v == kFar) return -1;
    *mask = ((uint64_t)T->cls_mask_hi[v] << 32) | T->cls_mask_lo[v];
    return (int)T->cls_lev[v];
}

// float64 brute force of the same: nearest level within prox (first wins) and every level with |L - p| <= prox (M:1574)
int classify_bruteforce_fine(int L, const double* price, double prox, float p, uint64_t* mask) {
    *mask = 0;
    if (L == 0) return -1;
    const int k = Classifier{L, price, prox}(p);
    if (k < 0) return -1;
    for (int j = 0; j < L; ++j)
        if (!(std::fabs(price[j] - (double)p) > prox)) *mask |= 1ull << j;
    return k;
}

}  // namespace qmmx
