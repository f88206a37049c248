/*
 * TEST INFRASTRUCTURE ONLY — plain-C restatement of the QMMX engine hot path, used as the
 * parity checker for the CUDA library at sizes the Python oracle cannot reach, and (bench.py
 * `cpu_baseline` / `--impl reference`) as the timed CPU port.  The product never links it.
 *
 * Parity status: PINNED — tests/test_c_oracle.py checks this file against oracle/headless.py
 * (itself pinned against the unmodified reference engine, see its header) and against the
 * reference-generated golden vectors in tests/golden/.
 *
 * M: = /root/reference/qmmx_monolithic.py.  Semantics = SURVEY.md Appendix A.
 * Arithmetic is float64 evaluated exactly like CPython: no FMA contraction (build with
 * -ffp-contract=off), brute-force first-wins nearest-level search (M:1543).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "qmmx_mc.h"

/* ---- OnlinePolicy._dot under CPython >= 3.12: builtin sum() of floats is Neumaier-compensated
 * (Python/bltinmodule.c); the first term is added to int 0 by the generic path. M:334-335 ---- */
static double dot_py312(const double* w, const double* x, int n) {
    double f = 0.0 + w[0] * x[0];
    double c = 0.0;
    for (int i = 1; i < n; ++i) {
        double v = w[i] * x[i];
        double t = f + v;
        if (fabs(f) >= fabs(v)) c += (f - t) + v;
        else c += (v - t) + f;
        f = t;
    }
    if (c != 0.0 && isfinite(c)) f += c;
    return f;
}

double oq_dot_py312(const double* w, const double* x, int n) { return dot_py312(w, x, n); }

/* builtin sum() of n floats, CPython >= 3.12 (same compensated loop as above with unit weights) */
static double sum_py312(const double* v, int n) {
    double f = 0.0 + v[0];
    double c = 0.0;
    for (int i = 1; i < n; ++i) {
        double t = f + v[i];
        if (fabs(f) >= fabs(v[i])) c += (f - t) + v[i];
        else c += (v[i] - t) + f;
        f = t;
    }
    if (c != 0.0 && isfinite(c)) f += c;
    return f;
}

/* _calc_volume_slope (M:1796-1811) at tick i of one path: recent_bars = the last min(i+1, 240) bars
 * (frozen in oracle/headless.py:run_path), window 6; volume is [step][paths] fp32 */
static double volume_slope_at(const float* volume, int64_t paths, int64_t path, int i) {
    const int len = i + 1 < 240 ? i + 1 : 240;
    if (!volume || len < 3) return 0.0;
    const int n = len < 6 ? len : 6;
    double v[6];
    for (int j = 0; j < n; ++j) {
        const double x = (double)volume[(int64_t)(i - n + 1 + j) * paths + path];
        v[j] = (x == 0.0) ? 0.0 : x;              /* `b.get("v") or ... or 0.0`: a falsy volume becomes +0.0 */
    }
    const int half = n / 2 > 2 ? n / 2 : 2;
    const double v1 = sum_py312(v, half) / (double)half;
    const double v2 = sum_py312(v + (n - half), half) / (double)half;
    if (v1 == 0 && v2 == 0) return 0.0;
    return (v2 - v1) / (fabs(v1) + 1e-9);
}
/* AccumulationBreakoutGuard (M:1241-1356) as the live loop feeds it (ingest_tick M:1857-1883 at M:2932): one-tick
 * minute bars, o = h = l = c; restated in oracle/headless.py:RegimeGuard.  Regimes 0 unknown, 1 accumulation,
 * 2 breakout_up, 3 breakout_down. */
typedef struct { int regime, inside, has_box; double box_low, box_high; } guard_t;

static double vol_at(const float* volume, int64_t paths, int64_t path, int j) {
    const double x = (double)volume[(int64_t)j * paths + path];
    return x == 0.0 ? 0.0 : x;                        /* `float(volume or 0.0)` */
}

static void guard_push(guard_t* G, const float* close, const float* volume, int64_t paths, int64_t path, int j) {
    const int n = j + 1 < 61 ? j + 1 : 61;            /* bars with t_end - t <= 60 min */
    if (n < 30) { G->regime = 0; G->inside = 0; G->has_box = 0; return; }
    double lo = INFINITY, hi = -INFINITY;
    for (int k = j - n + 1; k <= j; ++k) {
        const double c = (double)close[(int64_t)k * paths + path];
        if (c < lo) lo = c;
        if (c > hi) hi = c;
    }
    const double c = (double)close[(int64_t)j * paths + path];
    const double thresh = c * (18 / 10000.0);
    double v[20];
    for (int k = 0; k < 20; ++k) v[k] = vol_at(volume, paths, path, j - 19 + k);
    const double ma_s = sum_py312(v + 15, 5) / 5.0, ma_l = sum_py312(v, 20) / 20.0;
    const int breakout = G->regime == 2 || G->regime == 3;
    if (hi - lo <= (1e-6 > thresh ? 1e-6 : thresh)) {
        if (!breakout) G->regime = 1;
        G->box_low = lo; G->box_high = hi; G->has_box = 1;
    } else if (!breakout) G->regime = 0;
    if (G->has_box && ma_s != 0.0 && ma_l != 0.0) {
        if (c > G->box_high + 1e-6 && ma_s > 1.40 * ma_l) { G->regime = 2; G->inside = 0; }
        else if (c < G->box_low - 1e-6 && ma_s > 1.40 * ma_l) { G->regime = 3; G->inside = 0; }
    }
    if ((G->regime == 2 || G->regime == 3) && G->has_box) {
        if (G->box_low <= c && c <= G->box_high) { if (++G->inside >= 3) G->regime = 1; }
        else G->inside = 0;
    }
}

/* regime seen by evaluate_entry at every tick of one path (the guard has been fed bars 0..i-1) */
void oq_regime_series(const float* close, const float* volume, int64_t paths, int64_t path, int32_t steps, uint8_t* out) {
    guard_t G; memset(&G, 0, sizeof(G));
    for (int i = 0; i < steps; ++i) {
        if (i >= 1) guard_push(&G, close, volume, paths, path, i - 1);
        out[i] = (uint8_t)G.regime;
    }
}

double oq_volume_slope(const float* volume, int64_t paths, int64_t path, int i) { return volume_slope_at(volume, paths, path, i); }

/* M:20-26 */
static double* g_polw_out = NULL;

static double sigmoid_policy(double x);

/* OnlinePolicy.update_entry (M:357-363) -> _sgd_update (M:337-341): one logistic SGD step with L2, lr 0.03, l2 1e-6 (M:290) */
static void policy_sgd(double* w, const double* x, int label) {
    const double pred = sigmoid_policy(dot_py312(w, x, 7));
    double grad[7];
    for (int q = 0; q < 7; ++q) grad[q] = (pred - (double)label) * x[q] + 1e-6 * w[q];
    for (int q = 0; q < 7; ++q) w[q] -= 0.03 * grad[q];
}

static double sigmoid_policy(double x) {
    if (x < -50) return 0.0;
    if (x > 50) return 1.0;
    return 1.0 / (1.0 + exp(-x));
}

/* M:1415-1427 */
static double compute_confidence(double dist, int solid, int tc, double prox) {
    double base = 1.0 - (dist / (0.0001 > prox ? 0.0001 : prox));
    if (!(base > 0.0)) base = 0.0;       /* max(0.0, x) */
    base += solid ? 0.08 : 0.02;
    if (tc <= 1) base += 0.10;
    else if (tc == 2) base -= 0.08;
    else base -= (0.08 * 2);
    base += 0.03;
    double m = base < 1.0 ? base : 1.0;  /* min(1.0, base) */
    return m > 0.0 ? m : 0.0;            /* max(0.0, .)    */
}

double oq_confidence(double dist, int solid, int tc, double prox) {
    return compute_confidence(dist, solid, tc, prox);
}

typedef struct {
    const float *close, *high, *low, *volume;
    int64_t paths, p0, p1;
    int32_t steps;
    const double* lprice;
    const uint8_t* lsolid;
    int L;
    qmmx_params P;
    const double* lr; int n_lr;
    const double* pol; int n_pol;
    uint8_t* codes;
    qmmx_trade* trades;
    int32_t* tcounts;
    int32_t max_trades;
    double hist_lo, hist_hi;
    qmmx_agg agg;   /* per-thread partial */
} job_t;

static void run_range(job_t* J) {
    const qmmx_params* P = &J->P;
    const int L = J->L;
    const double prox = P->contact_prox;
    const double qmin = P->q_min_prob;
    double w_rules = P->w_rules;   /* M:1714-1719 (the `or 0.7` there guards the settings *string*) */
    double w_ml = P->w_ml;
    double s = w_rules + w_ml;
    if (s <= 0) { w_rules = 1.0; w_ml = 0.0; s = 1.0; }
    w_rules = w_rules / s; w_ml = w_ml / s;
    const int veto = (P->flags & QMMX_F_ENABLE_VETO) != 0;
    const int no_ml = (P->flags & QMMX_F_DISABLE_ML_GATE) != 0;
    const int blend = (P->flags & QMMX_F_USE_BLEND) != 0;
    const int no_pol = (P->flags & QMMX_F_DISABLE_POLICY_GATE) != 0;
    const int hl = (P->flags & QMMX_F_EXIT_ON_HL) != 0;
    /* the two labelled policy_events rows of a closed trade train the head of its action (oracle/headless.py run_path) */
    const int learn = !no_pol && (P->flags & QMMX_F_ONLINE_RETRAIN) != 0;
    const double inv_w = (double)QMMX_PNL_HIST_BINS / (J->hist_hi - J->hist_lo);
    qmmx_agg* A = &J->agg;
    memset(A, 0, sizeof(*A));
    A->max_path_pnl = -INFINITY; A->min_path_pnl = INFINITY;
    int touch[QMMX_MAX_LEVELS]; uint8_t latch[QMMX_MAX_LEVELS];

    for (int64_t path = J->p0; path < J->p1; ++path) {
        int have_prev = 0; double prev = 0.0;
        int last_dir = 0;
        int64_t cooldown_until = 0;
        int in_pos = 0, pside = 0, plevel = 0, pentry_bar = 0;
        double pentry = 0, pstop = 0, ptarget = 0;
        double path_pnl = 0.0;
        int ntr = 0;
        memset(touch, 0, sizeof(touch)); memset(latch, 0, sizeof(latch));
        double p = 0.0;
        guard_t G; memset(&G, 0, sizeof(G));
        double wl[14]; double pdist = 0.0;
        for (int q = 0; q < 14; ++q) wl[q] = (learn && J->n_pol > 0) ? J->pol[q] : 0.0;
        for (int i = 0; i < J->steps; ++i) {
            p = (double)J->close[(int64_t)i * J->paths + path];
            if (J->volume && i >= 1) guard_push(&G, J->close, J->volume, J->paths, path, i - 1);   /* ingest_tick, M:2932 */
            const int64_t now = 60000LL * i;
            const int day = i / P->steps_per_day;
            int code = -1, ok = 0, side = 0, k = 0;
            double stop = 0, target = 0, dist = 0;
            if (in_pos) code = QMMX_IN_POSITION;
            else if (now < cooldown_until) code = QMMX_COOLDOWN;
            else if (L == 0) code = QMMX_NOLEVELS;
            else {
                int d = 0;
                if (have_prev) {
                    if (p > prev + 1e-9) d = 1;
                    else if (p < prev - 1e-9) d = 2;
                    else d = last_dir;
                }
                if (d == 0) code = QMMX_DIR_UNKNOWN;
                else {
                    k = 0; dist = fabs(J->lprice[0] - p);
                    for (int j = 1; j < L; ++j) {
                        double dj = fabs(J->lprice[j] - p);
                        if (dj < dist) { dist = dj; k = j; }
                    }
                    if (dist > prox) code = QMMX_TOO_FAR;
                    else {
                        if (!latch[k]) { touch[k] += 1; latch[k] = 1; }
                        for (int j = 0; j < L; ++j)
                            if (j != k && latch[j] && fabs(J->lprice[j] - p) > prox) latch[j] = 0;
                        const int tc = touch[k];
                        if (tc >= P->overtouch_limit) code = QMMX_LEVEL_OVERTOUCHED;
                        else {
                            const double conf = compute_confidence(dist, J->lsolid[k], tc, prox);
                            if (conf < qmin) code = QMMX_CONF_LOW;
                            else {
                                const double lp = J->lprice[k];
                                if (d == 1) { stop = lp - P->stop_padding; target = lp + P->tp_padding; side = 0; }
                                else { stop = lp + P->stop_padding; target = lp - P->tp_padding; side = 1; }
                                code = QMMX_OK_ENTRY;
                                if (J->volume && ((G.regime == 2 && side == 1) || (G.regime == 3 && side == 0)))
                                    code = QMMX_ACC_BREAKOUT_GATE;         /* M:1652-1666 (int code 904) */
                                if (code == QMMX_OK_ENTRY && veto) {        /* M:1678-1705 + _soft_veto M:1773-1794; slope 0 without per-bar volume */
                                    const double vs = volume_slope_at(J->volume, J->paths, path, i);
                                    int confl = 0;
                                    for (int j = 0; j < L; ++j) if (fabs(J->lprice[j] - lp) <= 0.15) ++confl;
                                    if (!(fabs(vs) < 0.05 && !(confl >= 2))) {
                                        const double pw = prox * 0.12;
                                        const int near_ = dist <= (P->veto_prox > pw ? P->veto_prox : pw);
                                        /* the side follows the direction (M:1645), so only these two branches can fire */
                                        if (near_ && vs < -P->veto_vol_strong)
                                            code = (d == 1) ? QMMX_VETO_CONTRA_VOL_LONG : QMMX_VETO_CONTRA_VOL_SHORT;
                                    }
                                }
                                double mlp = conf; int ok_ml = 1;
                                if (code == QMMX_OK_ENTRY && !no_ml && J->n_lr > 0) {
                                    const double* m = J->lr + 5 * (day < J->n_lr ? day : J->n_lr - 1);
                                    const double distf = fabs(lp - stop);
                                    double z = m[4];
                                    z = z + m[0] * 1.0;
                                    z = z + m[1] * distf;
                                    z = z + m[2] * (double)tc;
                                    z = z + m[3] * (d == 1 ? 1.0 : 0.0);
                                    const double prob = 1.0 / (1.0 + exp(-z));
                                    ok_ml = prob >= qmin; mlp = prob;
                                }
                                if (code != QMMX_OK_ENTRY) { /* vetoed */ }
                                else if (blend) {
                                    const double blended = w_rules * conf + w_ml * mlp;
                                    if (blended < qmin) code = QMMX_COMBINED_LOW;
                                } else if (!no_ml && !ok_ml) code = QMMX_ML_CONF_LOW;
                                if (code == QMMX_OK_ENTRY) ok = 1;
                            }
                        }
                    }
                }
            }
            /* state update M:2952-2955 */
            const int had_prev = have_prev; const double pprev = prev;
            prev = p; have_prev = 1;
            if (had_prev && p != pprev) last_dir = (p > pprev) ? 1 : 2;
            /* exits M:2966-3014 */
            if (in_pos) {
                int reason = -1; double exit_px = p;
                if (hl) {
                    const double hh = (double)J->high[(int64_t)i * J->paths + path];
                    const double ll = (double)J->low[(int64_t)i * J->paths + path];
                    if (pside == 0) {
                        if (ll <= pstop) { reason = QMMX_EXIT_STOP; exit_px = pstop; }
                        else if (hh >= ptarget) { reason = QMMX_EXIT_TARGET; exit_px = ptarget; }
                    } else {
                        if (hh >= pstop) { reason = QMMX_EXIT_STOP; exit_px = pstop; }
                        else if (ll <= ptarget) { reason = QMMX_EXIT_TARGET; exit_px = ptarget; }
                    }
                } else {
                    if (pside == 0) {
                        if (p <= pstop) reason = QMMX_EXIT_STOP;
                        else if (p >= ptarget) reason = QMMX_EXIT_TARGET;
                    } else {
                        if (p >= pstop) reason = QMMX_EXIT_STOP;
                        else if (p <= ptarget) reason = QMMX_EXIT_TARGET;
                    }
                }
                if (reason >= 0) {
                    const double pnl = pside == 0 ? (exit_px - pentry) : (pentry - exit_px);
                    cooldown_until = now + (int64_t)P->cooldown_s * 1000;
                    if (J->trades) {
                        if (ntr < J->max_trades) {
                            qmmx_trade* t = &J->trades[path * J->max_trades + ntr];
                            t->entry_bar = pentry_bar; t->exit_bar = i; t->side = (uint8_t)pside;
                            t->reason = (uint8_t)reason; t->level_id = (uint16_t)plevel;
                            t->entry = (float)pentry; t->exit = (float)exit_px; t->pnl = (float)pnl;
                        } else A->n_trade_overflow++;
                    }
                    ntr++;
                    A->n_trades++; if (pnl > 0) A->n_wins++;
                    A->exit_hist[reason]++;
                    A->sum_pnl += pnl; A->sum_pnl2 += pnl * pnl; path_pnl += pnl;
                    if (learn) {      /* PRE_OPEN row (M:1757-1768), then the {"x": x} row whose keys the retrain cannot read */
                        const double xa[7] = {1.0, pdist < 1.0 ? pdist : 1.0, 0.0, pside == 1 ? 1.0 : 0.0, pside == 0 ? 1.0 : 0.0, 0.0, 0.0};
                        const double xb[7] = {1.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0};
                        policy_sgd(wl + 7 * pside, xa, pnl > 0 ? 1 : 0);
                        policy_sgd(wl + 7 * pside, xb, pnl > 0 ? 1 : 0);
                    }
                    in_pos = 0;
                }
            } else if (ok) {
                int pass = 1;
                if (!no_pol) {
                    const double lp = J->lprice[k];
                    static const double zero_w[21] = {0};
                    const double* w = J->n_pol > 0 ? J->pol + 21 * (day < J->n_pol ? day : J->n_pol - 1) : zero_w;
                    int cluster = 0;
                    for (int j = 0; j < L; ++j) if (fabs(J->lprice[j] - lp) <= 0.6) cluster++;
                    double x[7];
                    const double pa = fabs(p - lp);
                    const double tod = (double)(i % P->steps_per_day) / 390.0;
                    x[0] = 1.0; x[1] = pa < 1.0 ? pa : 1.0; x[2] = 0.0;
                    x[3] = side == 1 ? 1.0 : 0.0; x[4] = side == 0 ? 1.0 : 0.0;
                    x[5] = cluster > 1 ? 1.0 : 0.0; x[6] = tod < 1.0 ? tod : 1.0;
                    const double s_go = sigmoid_policy(dot_py312(learn ? wl + 7 * side : w + 7 * (side == 0 ? 0 : 1), x, 7));
                    const double s_skip = sigmoid_policy(dot_py312(w + 14, x, 7));
                    pass = (s_go >= 0.60) && (s_skip < 0.55);
                }
                if (!pass) code = QMMX_ONLINE_POLICY;
                else {
                    in_pos = 1; pside = side; plevel = k; pentry = p; pstop = stop; ptarget = target;
                    pentry_bar = i; pdist = fabs(J->lprice[k] - p);
                }
            }
            A->entry_hist[code]++;
            if (J->codes) J->codes[(int64_t)i * J->paths + path] = (uint8_t)code;
        }
        if (in_pos) {
            const double mtm = pside == 0 ? (p - pentry) : (pentry - p);
            if (J->trades) {
                if (ntr < J->max_trades) {
                    qmmx_trade* t = &J->trades[path * J->max_trades + ntr];
                    t->entry_bar = pentry_bar; t->exit_bar = -1; t->side = (uint8_t)pside;
                    t->reason = QMMX_EXIT_OPEN; t->level_id = (uint16_t)plevel;
                    t->entry = (float)pentry; t->exit = (float)p; t->pnl = 0.0f;
                } else A->n_trade_overflow++;
            }
            ntr++;
            A->n_open++; A->exit_hist[QMMX_EXIT_OPEN]++; A->sum_open_mtm += mtm;
        }
        if (J->tcounts) J->tcounts[path] = ntr;
        if (learn && g_polw_out) memcpy(g_polw_out + path * 14, wl, sizeof(wl));
        A->n_paths++;
        A->sum_path_pnl2 += path_pnl * path_pnl;
        if (path_pnl > A->max_path_pnl) A->max_path_pnl = path_pnl;
        if (path_pnl < A->min_path_pnl) A->min_path_pnl = path_pnl;
        double fb = floor((path_pnl - J->hist_lo) * inv_w);
        int bin = fb < 0 ? 0 : (fb > QMMX_PNL_HIST_BINS - 1 ? QMMX_PNL_HIST_BINS - 1 : (int)fb);
        A->pnl_hist[bin]++;
    }
    A->n_steps = 0;
    for (int c = 0; c < QMMX_N_ENTRY_CODES; ++c) A->n_steps += A->entry_hist[c];
}

static void* thread_main(void* arg) { run_range((job_t*)arg); return NULL; }

/* QMMX_F_ONLINE_RETRAIN: where the next oq_run_paths* call leaves the learnt go_long / go_short heads, [path][14] (NULL = nowhere) */
void oq_set_policy_weights_out(double* out) { g_polw_out = out; }

/* Batch run over `[step][path]` fp32 bars; n_threads pthreads each take a contiguous path range. */
int oq_run_paths_v(const float* close, const float* high, const float* low, const float* volume, int64_t paths, int32_t steps,
                   const double* lprice, const uint8_t* lsolid, int L, const qmmx_params* P,
                   const double* lr, int n_lr, const double* pol, int n_pol,
                   double hist_lo, double hist_hi,
                   uint8_t* codes, qmmx_trade* trades, int32_t* tcounts, int32_t max_trades,
                   qmmx_agg* agg, int n_threads);

int oq_run_paths(const float* close, const float* high, const float* low, int64_t paths, int32_t steps,
                 const double* lprice, const uint8_t* lsolid, int L, const qmmx_params* P,
                 const double* lr, int n_lr, const double* pol, int n_pol,
                 double hist_lo, double hist_hi,
                 uint8_t* codes, qmmx_trade* trades, int32_t* tcounts, int32_t max_trades,
                 qmmx_agg* agg, int n_threads) {
    return oq_run_paths_v(close, high, low, NULL, paths, steps, lprice, lsolid, L, P, lr, n_lr, pol, n_pol, hist_lo, hist_hi,
                          codes, trades, tcounts, max_trades, agg, n_threads);
}

/* same with per-bar volume [step][path] (NULL = none): only the soft veto reads it */
int oq_run_paths_v(const float* close, const float* high, const float* low, const float* volume, int64_t paths, int32_t steps,
                   const double* lprice, const uint8_t* lsolid, int L, const qmmx_params* P,
                   const double* lr, int n_lr, const double* pol, int n_pol,
                   double hist_lo, double hist_hi,
                   uint8_t* codes, qmmx_trade* trades, int32_t* tcounts, int32_t max_trades,
                   qmmx_agg* agg, int n_threads) {
    if (L > QMMX_MAX_LEVELS || paths < 0 || steps < 0) return -1;
    if ((P->flags & QMMX_F_EXIT_ON_HL) && (!high || !low)) return -1;
    if (n_threads < 1) n_threads = 1;
    if ((int64_t)n_threads > paths) n_threads = paths > 0 ? (int)paths : 1;
    job_t* jobs = (job_t*)calloc((size_t)n_threads, sizeof(job_t));
    pthread_t* th = (pthread_t*)calloc((size_t)n_threads, sizeof(pthread_t));
    for (int t = 0; t < n_threads; ++t) {
        job_t* J = &jobs[t];
        J->close = close; J->high = high; J->low = low; J->volume = volume; J->paths = paths; J->steps = steps;
        J->p0 = paths * t / n_threads; J->p1 = paths * (t + 1) / n_threads;
        J->lprice = lprice; J->lsolid = lsolid; J->L = L; J->P = *P;
        J->lr = lr; J->n_lr = n_lr; J->pol = pol; J->n_pol = n_pol;
        J->codes = codes; J->trades = trades; J->tcounts = tcounts; J->max_trades = max_trades;
        J->hist_lo = hist_lo; J->hist_hi = hist_hi;
        if (n_threads > 1) pthread_create(&th[t], NULL, thread_main, J);
        else run_range(J);
    }
    qmmx_agg out; memset(&out, 0, sizeof(out));
    out.struct_size = sizeof(out);
    out.max_path_pnl = -INFINITY; out.min_path_pnl = INFINITY;
    out.pnl_hist_lo = hist_lo; out.pnl_hist_hi = hist_hi;
    for (int t = 0; t < n_threads; ++t) {
        if (n_threads > 1) pthread_join(th[t], NULL);
        const qmmx_agg* a = &jobs[t].agg;
        out.n_paths += a->n_paths; out.n_steps += a->n_steps; out.n_trades += a->n_trades;
        out.n_wins += a->n_wins; out.n_open += a->n_open; out.n_trade_overflow += a->n_trade_overflow;
        for (int c = 0; c < QMMX_N_ENTRY_CODES; ++c) out.entry_hist[c] += a->entry_hist[c];
        for (int c = 0; c < QMMX_N_EXIT_CODES; ++c) out.exit_hist[c] += a->exit_hist[c];
        for (int c = 0; c < QMMX_PNL_HIST_BINS; ++c) out.pnl_hist[c] += a->pnl_hist[c];
        out.sum_pnl += a->sum_pnl; out.sum_pnl2 += a->sum_pnl2; out.sum_path_pnl2 += a->sum_path_pnl2;
        out.sum_open_mtm += a->sum_open_mtm;
        if (a->max_path_pnl > out.max_path_pnl) out.max_path_pnl = a->max_path_pnl;
        if (a->min_path_pnl < out.min_path_pnl) out.min_path_pnl = a->min_path_pnl;
    }
    if (agg) *agg = out;
    free(jobs); free(th);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11) and the
 * path generators.  No reference equivalent: the reference simulates on real Polygon bars.
 * ---------------------------------------------------------------------------------------- */
void oq_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Box-Muller on two 32-bit draws; same operation order as the device code, libm transcendentals */
static void box_muller(uint32_t a, uint32_t b, float* z0, float* z1) {
    const float u1 = fmaf((float)a, 2.3283064365386963e-10f, 1.16415321826934814453125e-10f);          /* (a + 0.5) 2^-32, (0, 1] */
    const float r = sqrtf(log2f(u1) * -1.38629436111989061883f);           /* -2 ln u1 = (-2 ln 2) lg2 u1 */
    /* angle 2 pi u2 - pi, u2 = (b' + 0.5) 2^-24 in (0, 1): the half turn is taken as a sign, as the device does */
    const float ang = fmaf((float)(b >> 8), 3.7450702829239286e-07f, 1.8725351414619643e-07f);
    *z0 = -r * cosf(ang); *z1 = -r * sinf(ang);
}

static float round_cents(float s) { return rintf(s * 100.0f) / 100.0f; }

/* GBM closes (and OHLC when sub_steps == 4) for global path ids [path_offset, path_offset+paths) */
int oq_gen_gbm(uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps, const qmmx_gbm* g,
               float* close, float* high, float* low) {
    const float drift = (float)(((double)g->mu_annual - 0.5 * (double)g->sigma_annual * (double)g->sigma_annual)
                                * (double)g->dt_years / (double)g->sub_steps);
    const float vol = (float)((double)g->sigma_annual * sqrt((double)g->dt_years / (double)g->sub_steps));
    /* the step in base 2, as the device takes it: S *= 2^(vol2 * z + drift2), one fused multiply-add */
    const float drift2 = (float)((double)drift * 1.4426950408889634074), vol2 = (float)((double)vol * 1.4426950408889634074);
    const uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    const int cents = (g->flags & QMMX_GBM_ROUND_CENTS) != 0;
    if (g->sub_steps != 1 && g->sub_steps != 4) return -1;
    for (int64_t q = 0; q < paths; ++q) {
        const uint64_t gid = path_offset + (uint64_t)q;
        float S = g->s0;
        float z[4] = {0, 0, 0, 0};
        for (int i = 0; i < steps; ++i) {
            if (g->sub_steps == 4 || (i & 3) == 0) {
                const uint32_t ctr[4] = {(uint32_t)gid, (uint32_t)(gid >> 32),
                                         (uint32_t)(g->sub_steps == 4 ? i : i >> 2), 0u};
                uint32_t x[4];
                oq_philox4x32_10(ctr, key, x);
                box_muller(x[0], x[1], &z[0], &z[1]);
                box_muller(x[2], x[3], &z[2], &z[3]);
            }
            if (g->sub_steps == 1) {
                S = S * exp2f(fmaf(vol2, z[i & 3], drift2));
                close[(int64_t)i * paths + q] = cents ? round_cents(S) : S;
            } else {
                float hi = S, lo = S;
                for (int s = 0; s < 4; ++s) {
                    S = S * exp2f(fmaf(vol2, z[s], drift2));
                    hi = fmaxf(hi, S); lo = fminf(lo, S);
                }
                close[(int64_t)i * paths + q] = cents ? round_cents(S) : S;
                if (high) high[(int64_t)i * paths + q] = cents ? round_cents(hi) : hi;
                if (low) low[(int64_t)i * paths + q] = cents ? round_cents(lo) : lo;
            }
        }
    }
    return 0;
}

/* Circular block bootstrap of the window's bar returns (relative to the previous close). */
int oq_gen_bootstrap_v(uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                       const float* window_ohlc, const float* window_volume, int32_t window_len, int32_t block_len,
                       float* close, float* high, float* low, float* volume);

int oq_gen_bootstrap(uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                     const float* window_ohlc, int32_t window_len, int32_t block_len,
                     float* close, float* high, float* low) {
    return oq_gen_bootstrap_v(seed, path_offset, paths, steps, window_ohlc, NULL, window_len, block_len, close, high, low, NULL);
}

/* with per-bar volume: a resampled bar carries the volume of the window bar it was drawn from */
int oq_gen_bootstrap_v(uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                       const float* window_ohlc, const float* window_volume, int32_t window_len, int32_t block_len,
                       float* close, float* high, float* low, float* volume) {
    const int nret = window_len - 1;
    if (nret < 1 || block_len < 1 || nret > 4096) return -1;
    float* rh = (float*)malloc(sizeof(float) * 3 * (size_t)nret);
    float* rl = rh + nret; float* rc = rl + nret;
    for (int t = 1; t < window_len; ++t) {
        const float pc = window_ohlc[4 * (t - 1) + 3];
        rh[t - 1] = window_ohlc[4 * t + 1] / pc;
        rl[t - 1] = window_ohlc[4 * t + 2] / pc;
        rc[t - 1] = window_ohlc[4 * t + 3] / pc;
    }
    const uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    const float s0 = window_ohlc[4 * (window_len - 1) + 3];
    for (int64_t q = 0; q < paths; ++q) {
        const uint64_t gid = path_offset + (uint64_t)q;
        float S = s0;
        uint32_t x[4] = {0, 0, 0, 0};
        int start = 0;
        for (int i = 0; i < steps; ++i) {
            const int blk = i / block_len, off = i % block_len;
            if (off == 0) {
                if ((blk & 3) == 0) {
                    const uint32_t ctr[4] = {(uint32_t)gid, (uint32_t)(gid >> 32), (uint32_t)(blk >> 2), 1u};
                    oq_philox4x32_10(ctr, key, x);
                }
                start = (int)(((uint64_t)x[blk & 3] * (uint64_t)nret) >> 32);
            }
            int idx = start + off; idx %= nret;
            const int64_t o = (int64_t)i * paths + q;
            if (high) high[o] = S * rh[idx];
            if (low) low[o] = S * rl[idx];
            S = S * rc[idx];
            close[o] = S;
            if (volume && window_volume) volume[o] = window_volume[idx + 1];
        }
    }
    free(rh);
    return 0;
}
