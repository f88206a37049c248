/*
 * qmmx_mc.h — C ABI of libqmmx_mc.so: the B200 (sm_100a) Monte Carlo evaluator of the QMMX
 * level-touch trade engine.
 *
 * The reference (RustyWill/QMMX_Monolithic_Monte_Carlo, `M` = qmmx_monolithic.py) has no FFI;
 * its de-facto interface for this path is object state + three Python call shapes.  Each entry
 * point below names the reference code it stands in for:
 *
 *   qmmx_set_levels      engine.levels_cache = load_levels(conn)                       M:140-144, M:1368
 *   qmmx_set_params      engine.CONTACT_PROX / STOP_PADDING / TP_PADDING / Q_MIN_PROB /
 *                        Q_SIGNAL_COOLDOWN_S / ENABLE_VETO / DISABLE_ML_GATE and the
 *                        per-call settings_get of USE_BLEND / W_RULES / W_ML /
 *                        DISABLE_POLICY_GATE                                           M:1371-1386, M:1711-1715, M:3090
 *   qmmx_set_lr          engine.model (sklearn LogisticRegression, 4 coef + intercept)  M:1401-1407, M:1454-1466
 *   qmmx_set_policy      app.policy.w_entry (OnlinePolicy go_long/go_short/skip, dim 7) M:296-331, M:2096-2104
 *   qmmx_upload_bars     engine.recent_bars = [{t,o,h,l,c}, ...]                        M:220-240, M:2430
 *   qmmx_upload_volume   the "v" of those bars (read by _calc_volume_slope only)        M:1796-1811, M:1880
 *   qmmx_run_scan        evaluate_entry per tick + the live loop's exit block,
 *                        open_trade / close_trade pnl, set_cooldown                     M:1492-1771, M:2927-3016, M:3044-3115,
 *                                                                                       M:1888-1948, M:259-270
 *   qmmx_lr_score        OnlinePolicy.score_entry / predict_proba as a batch            M:334-355, M:20-26, M:1463
 *   qmmx_run_jitter      QMMXApp.simulate_monte_carlo trial loop (walk_outcome)         M:3449-3510
 *   qmmx_run_last_bars   QMMXApp.simulate_last_bars forward walk (stop before target)    M:3619-3639
 *   qmmx_gen_gbm / qmmx_gen_bootstrap / qmmx_run_fused_gbm / qmmx_run_fused_bootstrap
 *                        no reference equivalent (the reference simulates on real Polygon bars);
 *                        synthetic 1-minute paths demanded by BASELINE.json north_star.
 *
 * Conventions: every function returns an int status (0 = QMMX_OK, negative = error, see
 * qmmx_strerror); nothing throws or aborts across the ABI.  The caller owns every host buffer;
 * the library owns device buffers inside the ctx; no pointer handed out by the library outlives
 * qmmx_destroy.  One ctx = one GPU = one CUDA stream; a ctx is not thread-safe, distinct ctxs are
 * independent.  Calls are synchronous (they return after the stream has drained) unless the
 * name ends in _async.  There is NO CPU fallback: without a CUDA device qmmx_create fails.
 */
#ifndef QMMX_MC_H
#define QMMX_MC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QMMX_ABI_VERSION 1

#define QMMX_MAX_LEVELS      64
#define QMMX_N_ENTRY_CODES   20
#define QMMX_N_EXIT_CODES    4
#define QMMX_PNL_HIST_BINS   256
#define QMMX_MAX_COMBOS      1024
#define QMMX_LR_DIM          5   /* coef[4] + intercept, x = [lvl_type(=1), |level-stop|, touch, dir_up] */
#define QMMX_POLICY_HEADS    3   /* go_long, go_short, skip */
#define QMMX_POLICY_DIM      7

/* status codes */
enum {
    QMMX_OK            = 0,
    QMMX_E_BADARG      = -1,
    QMMX_E_CUDA        = -2,
    QMMX_E_NOMEM       = -3,
    QMMX_E_STATE       = -4,   /* call order: levels/params/bars not set */
    QMMX_E_UNSUPPORTED = -5,
    QMMX_E_NCCL        = -6    /* libnccl.so.2 missing, or an NCCL call failed (qmmx_last_error has the text) */
};

/* entry decision codes: reason strings of M:246-257 plus the sub-reasons the reference writes
 * into policy_events (M:1600, M:1659, M:1695, M:1734, M:1747, M:3098).  Stable integer ids. */
enum {
    QMMX_OK_ENTRY = 0, QMMX_IN_POSITION, QMMX_COOLDOWN, QMMX_NOLEVELS, QMMX_DIR_UNKNOWN,
    QMMX_TOO_FAR, QMMX_LEVEL_OVERTOUCHED, QMMX_CONF_LOW, QMMX_ML_CONF_LOW, QMMX_COMBINED_LOW,
    QMMX_VETO_CONTRA_VOL_LONG, QMMX_VETO_CONTRA_VOL_SHORT, QMMX_VETO_EDGE_FATIGUE,
    QMMX_VETO_BOUNCE_BUDGET, QMMX_VETO_LEVEL_COOLDOWN, QMMX_ACC_BREAKOUT_GATE,
    QMMX_ONLINE_POLICY, QMMX_PRICE_STALE, QMMX_MISSING_API_KEY, QMMX_RISK_INVALID
};
/* exit codes: M:2979-3012 ("STOP", "TARGET"), M:1998 ("ESCALATE"), end-of-path open (M:3630) */
enum { QMMX_EXIT_STOP = 0, QMMX_EXIT_TARGET, QMMX_EXIT_ESCALATE, QMMX_EXIT_OPEN };

/* qmmx_params.flags */
#define QMMX_F_ENABLE_VETO          (1u << 0)  /* ENABLE_VETO (needs qmmx_upload_volume to ever fire) */
#define QMMX_F_DISABLE_ML_GATE      (1u << 1)  /* DISABLE_ML_GATE                                  */
#define QMMX_F_USE_BLEND            (1u << 2)  /* USE_BLEND                                        */
#define QMMX_F_DISABLE_POLICY_GATE  (1u << 3)  /* DISABLE_POLICY_GATE (1 = OnlinePolicy bypassed)  */
#define QMMX_F_EXIT_ON_HL           (1u << 4)  /* exits tested on bar low/high, stop first (M:3621-3636) */
#define QMMX_F_TINY_PRICES          (1u << 5)  /* prices below 2^-5 may occur: evaluate the +-1e-9 direction guard
                                                  (M:1529-1537) in float64 on every tick instead of the fp32 shortcut */
#define QMMX_F_ONLINE_RETRAIN       (1u << 6)  /* the OnlinePolicy go_long / go_short heads learn from the path's own closed
                                                 * trades (update_entry M:357-363 through _retrain_from_labeled_events
                                                 * M:3753-3803, lr 0.03, l2 1e-6 of M:290): qmmx_set_policy gives ONE table, the
                                                 * initial weights; needs the policy gate on, one combo, the lean engine (<= 32 levels for the fused kernels) */

typedef struct qmmx_params {
    uint32_t struct_size;      /* = sizeof(qmmx_params) */
    uint32_t flags;
    double   contact_prox;     /* CONTACT_PROX       default 0.05 (M:1371) */
    double   stop_padding;     /* STOP_PADDING       default 0.35 (M:1374) */
    double   tp_padding;       /* TP_PADDING         default 0.25 (M:1375) */
    double   q_min_prob;       /* Q_MIN_PROB         default 0.60 (M:1429-1447) */
    double   w_rules;          /* W_RULES            default 0.7  (M:1714), normalised like M:1716-1719 */
    double   w_ml;             /* W_ML               default 0.3  (M:1715) */
    double   veto_vol_strong;  /* VETO_VOL_STRONG    default 0.25 (M:1381) */
    double   veto_prox;        /* VETO_PROX          default 0.06 (M:1382) */
    int32_t  cooldown_s;       /* Q_SIGNAL_COOLDOWN  default 8 s  (M:1372) */
    int32_t  overtouch_limit;  /* hard-coded 4 in the reference (M:1579); 1..15 here */
    int32_t  steps_per_day;    /* 390; day = step / steps_per_day selects the LR / policy row */
    int32_t  reserved;
} qmmx_params;

/* GBM path recipe: S *= exp((mu - sigma^2/2) dt + sigma sqrt(dt) z), fp32 */
#define QMMX_GBM_ROUND_CENTS (1u << 0)  /* stored bars rounded to $0.01 (recurrence stays unrounded) */
typedef struct qmmx_gbm {
    uint32_t struct_size;
    uint32_t flags;
    float    s0;
    float    mu_annual;
    float    sigma_annual;
    float    dt_years;         /* 1 / (252 * 390) */
    int32_t  sub_steps;        /* 1 = close-only bars; 4 = OHLC from 4 sub-steps per bar */
    int32_t  reserved;
} qmmx_gbm;

/* one closed (or end-of-path open) trade; parity-mode output, [path][max_trades_per_path] */
typedef struct qmmx_trade {
    int32_t  entry_bar;
    int32_t  exit_bar;         /* -1 for QMMX_EXIT_OPEN */
    uint8_t  side;             /* 0 long, 1 short */
    uint8_t  reason;           /* QMMX_EXIT_* */
    uint16_t level_id;         /* index into the level list in the order given to qmmx_set_levels */
    float    entry;
    float    exit;             /* exit price; last close for OPEN */
    float    pnl;              /* realised $/share (M:1926); 0 for OPEN; fp32 cast of the fp64 value */
} qmmx_trade;

/* aggregates of one parameter combo over all paths of a run (or of all ranks after reduction) */
typedef struct qmmx_agg {
    uint32_t struct_size;
    uint32_t combo;
    /* integer counters — exactly reproducible for any GPU count */
    uint64_t n_paths;
    uint64_t n_steps;          /* path-steps evaluated = sum(entry_hist) */
    uint64_t n_trades;         /* closed trades (STOP + TARGET) */
    uint64_t n_wins;           /* closed trades with pnl > 0 (label of M:1936) */
    uint64_t n_open;           /* paths ending with an open position */
    uint64_t n_trade_overflow; /* trades dropped from the parity-mode record buffer */
    uint64_t entry_hist[QMMX_N_ENTRY_CODES];
    uint64_t exit_hist[QMMX_N_EXIT_CODES];
    uint64_t pnl_hist[QMMX_PNL_HIST_BINS]; /* per-path realised PnL, fixed bins over [pnl_hist_lo, pnl_hist_hi) */
    /* fp64 sums (per-trade pnl is exact in fp64; sums are fp64, block-ordered) */
    double   sum_pnl;          /* sum of realised pnl over closed trades */
    double   sum_pnl2;         /* sum of pnl^2 over closed trades */
    double   sum_path_pnl2;    /* sum over paths of (path realised pnl)^2 */
    double   sum_open_mtm;     /* sum of mark-to-market of end-of-path open positions (M:3266-3270) */
    double   max_path_pnl;
    double   min_path_pnl;
    double   pnl_hist_lo;
    double   pnl_hist_hi;
} qmmx_agg;

/* flat reduction buffer layout per combo (what lives on the device and what NCCL reduces):
 *   uint64 counters [QMMX_AGG_NU]  -> SUM
 *   double sums     [QMMX_AGG_ND]  -> SUM
 *   double extrema  [2] = {max_path_pnl, -min_path_pnl} -> MAX                                  */
#define QMMX_AGG_NU (6 + QMMX_N_ENTRY_CODES + QMMX_N_EXIT_CODES + QMMX_PNL_HIST_BINS)
#define QMMX_AGG_ND 4
#define QMMX_AGG_NX 2
#define QMMX_AGG_SLOTS (QMMX_AGG_NU + QMMX_AGG_ND + QMMX_AGG_NX)   /* 8-byte slots per combo */

typedef struct qmmx_ctx qmmx_ctx;

/* ---- lifetime ---- */
int  qmmx_abi_version(void);
const char* qmmx_strerror(int status);
const char* qmmx_last_error(const qmmx_ctx* ctx);          /* detail text of the last failure */
int  qmmx_create(int device, qmmx_ctx** out);
int  qmmx_destroy(qmmx_ctx* ctx);
int  qmmx_set_stream(qmmx_ctx* ctx, void* cuda_stream);    /* run on a caller stream (e.g. torch's); NULL = own */

/* ---- engine configuration ---- */
/* levels in the reference's load_levels order (ORDER BY color, level_type, level_index): ties in the
 * nearest-level search go to the FIRST level in this order (Python min, M:1543). */
int  qmmx_set_levels(qmmx_ctx* ctx, const double* price, const uint8_t* solid, int n_levels);
/* one qmmx_params per parameter combination; every run reports one qmmx_agg per combo, in this order.  Combos that
 * differ only in STOP_PADDING / TP_PADDING / cooldown (a sweep, 20 or more of them, no blend / policy / veto) are
 * evaluated 32 to a warp over the same path by one launch; the results are identical to one launch per combo
 * (environment QMMX_NO_SWEEP=1 forces the latter). */
int  qmmx_set_params(qmmx_ctx* ctx, const qmmx_params* params, int n_combos);
int  qmmx_set_lr(qmmx_ctx* ctx, const double* coef_intercept /* [n_days][5] */, int n_days);   /* n_days 0 = no model */
int  qmmx_set_policy(qmmx_ctx* ctx, const double* w /* [n_days][3][7] */, int n_days);        /* n_days 0 = zero weights */
int  qmmx_set_pnl_hist_range(qmmx_ctx* ctx, double lo, double hi);

/* ---- bars: device-resident, reusable across runs (parameter sweeps) ---- */
/* host arrays, fp32, layout [step][path]; high/low may be NULL (close-only) */
int  qmmx_upload_bars(qmmx_ctx* ctx, const float* close, const float* high, const float* low,
                      int64_t paths, int32_t steps);
/* same, but the arrays already live on this ctx's device (e.g. torch tensors); not copied.
 * pitch = row stride in elements (>= paths) */
int  qmmx_bind_device_bars(qmmx_ctx* ctx, const float* d_close, const float* d_high, const float* d_low,
                           int64_t paths, int32_t steps, int64_t pitch);
/* device pointers of the resident bars (NULL when a field is absent) */
int  qmmx_device_bars(qmmx_ctx* ctx, float** d_close, float** d_high, float** d_low, int64_t* pitch);
/* page-lock / unlock a caller-owned host buffer so uploads run at full PCIe rate */
int  qmmx_pin_host(qmmx_ctx* ctx, void* ptr, size_t bytes);
int  qmmx_unpin_host(qmmx_ctx* ctx, void* ptr);
int  qmmx_download_bars(qmmx_ctx* ctx, float* close, float* high, float* low);
/* per-bar volume of the resident bars ([step][path] fp32, same shape; NULL drops it).  Two gates read it:
 *  - the soft veto (QMMX_F_ENABLE_VETO: _calc_volume_slope over the last 6 bars, M:1796-1811, _soft_veto M:1773-1794);
 *  - the AccumulationBreakoutGuard regime gate (always on in the reference: 60-minute box, 5/20-bar volume means,
 *    M:1241-1356, fed by ingest_tick M:1857-1883; counter-trend entries after a confirmed break get
 *    QMMX_ACC_BREAKOUT_GATE, M:1652-1666).  The regime plane is computed once, when the volume arrives.
 * Without volume the slope is 0 and no break is ever confirmed: both gates are inert, which is the reference on
 * its own data. */
int  qmmx_upload_volume(qmmx_ctx* ctx, const float* volume, int64_t paths, int32_t steps);
int  qmmx_download_volume(qmmx_ctx* ctx, float* volume);
/* QMMX_F_ONLINE_RETRAIN: the learnt go_long / go_short heads of every path after the last qmmx_run_scan,
 * out[path][14] float64 (go_long[7], go_short[7]); the skip head never learns (its rows carry no label) */
int  qmmx_download_policy_weights(qmmx_ctx* ctx, double* out, int64_t paths);
/* volume of the window bars ([window_len] fp32, NULL clears): the next qmmx_gen_bootstrap with the same
 * window_len also writes a volume plane (a resampled bar carries the volume of the window bar it came from) */
int  qmmx_set_bootstrap_volume(qmmx_ctx* ctx, const float* window_volume, int32_t window_len);
/* Philox4x32-10 generators, counter = (global path id, step group), key = seed: results do not
 * depend on how paths are split across GPUs.  path_offset = first global path id of this shard. */
int  qmmx_gen_gbm(qmmx_ctx* ctx, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                  const qmmx_gbm* gbm);
/* block bootstrap of a candle window: window_ohlc = [window_len][4] (o,h,l,c); blocks of block_len
 * consecutive bar returns are drawn with replacement and chained from the window's last close */
int  qmmx_gen_bootstrap(qmmx_ctx* ctx, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                        const float* window_ohlc, int32_t window_len, int32_t block_len);

/* ---- the hot path ---- */
/* Scan every resident path through the engine for every combo.
 *   agg_out            [n_combos] or NULL
 *   trades_out         [paths][max_trades_per_path] or NULL   (parity mode, combo 0 only)
 *   trade_counts_out   [paths] or NULL
 *   codes_out          [steps][paths] entry code per step or NULL (parity mode, combo 0 only) */
int  qmmx_run_scan(qmmx_ctx* ctx, qmmx_agg* agg_out, qmmx_trade* trades_out, int32_t* trade_counts_out,
                   int32_t max_trades_per_path, uint8_t* codes_out);
/* End to end from HOST bars: path chunks are uploaded on a copy stream while the previous chunk is
 * scanned (double-buffered device staging); aggregates only.  chunk_paths <= 0 picks a default. */
int  qmmx_run_scan_host(qmmx_ctx* ctx, const float* close, const float* high, const float* low,
                        int64_t paths, int32_t steps, int64_t chunk_paths, qmmx_agg* agg_out);
/* generate GBM closes in registers and scan them in the same kernel (no bar traffic to HBM) */
int  qmmx_run_fused_gbm(qmmx_ctx* ctx, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                        const qmmx_gbm* gbm, qmmx_agg* agg_out);
/* block-bootstrap the window (as qmmx_gen_bootstrap: same draws, same bars) and scan in the same kernel; the bars
 * are never written, so they carry no volume plane (the soft veto is inert on this path) */
int  qmmx_run_fused_bootstrap(qmmx_ctx* ctx, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                              const float* window_ohlc, int32_t window_len, int32_t block_len, qmmx_agg* agg_out);
int  qmmx_run_fused_bootstrap_async(qmmx_ctx* ctx, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                                    const float* window_ohlc, int32_t window_len, int32_t block_len);
/* launch only; results stay on the device until qmmx_fetch_agg / qmmx_agg_device */
int  qmmx_run_scan_async(qmmx_ctx* ctx);
int  qmmx_run_fused_gbm_async(qmmx_ctx* ctx, uint64_t seed, uint64_t path_offset, int64_t paths, int32_t steps,
                              const qmmx_gbm* gbm);
int  qmmx_fetch_agg(qmmx_ctx* ctx, qmmx_agg* agg_out /* [n_combos] */);

/* flat device reduction buffer of the last run: n_combos * QMMX_AGG_SLOTS 8-byte slots, for the
 * caller's collective (torch.distributed / NCCL all-reduce); unpack turns a flat HOST copy into structs */
int  qmmx_agg_device(qmmx_ctx* ctx, void** d_flat, size_t* bytes);
int  qmmx_agg_unpack(qmmx_ctx* ctx, const void* flat_host, int n_combos, qmmx_agg* out);
/* multi-GPU merge with ONE collective: all-gather every rank's flat buffer into d_parts
 * ([n_parts][n_combos * QMMX_AGG_SLOTS] 8-byte slots, device memory), then this call reduces them into the ctx's
 * own buffer on the ctx stream (SUM counters, SUM sums, MAX extrema).  Asynchronous; qmmx_fetch_agg reads it. */
int  qmmx_merge_flat(qmmx_ctx* ctx, const void* d_parts, int n_parts);

/* Multi-GPU merge without torch (SURVEY 8b `qmmx_allreduce(ctx, nccl_comm)`; the reference itself is one process,
 * 8e): ONE ncclAllGather of the flat buffer on the ctx stream, then qmmx_merge_flat, so that after the call every
 * rank's buffer holds the whole job's aggregates (SUM counters, SUM sums, MAX extrema).  libnccl.so.2 is dlopen'ed on
 * first use (the copy already loaded in the process, e.g. torch's, else the system one); nothing links against it.
 * nccl_comm: the caller's ncclComm_t, or NULL to use the communicator made by qmmx_nccl_init.  Asynchronous on the
 * ctx stream; qmmx_fetch_agg reads the result.
 * qmmx_nccl_unique_id: rank 0 fills 128 bytes (ncclGetUniqueId) which the caller ships to the other ranks by any
 * means (file, socket, MPI, torch.distributed); qmmx_nccl_init: every rank joins with the same id (ncclCommInitRank);
 * the communicator lives until qmmx_destroy. */
int  qmmx_nccl_unique_id(qmmx_ctx* ctx, void* id128_out);
int  qmmx_nccl_init(qmmx_ctx* ctx, const void* id128, int world_size, int rank);
int  qmmx_allreduce(qmmx_ctx* ctx, void* nccl_comm);
/* the CUDA stream the ctx launches on (its own, or the one given to qmmx_set_stream): for callers that order their
 * own work against it */
int  qmmx_get_stream(qmmx_ctx* ctx, void** cuda_stream_out);

/* batch logistic scoring, fp32 FMA: out[i][h] = sigmoid(w[h] . x[i]) (+ bias[h] if bias != NULL) */
int  qmmx_lr_score(qmmx_ctx* ctx, const float* x /* [n][dim] */, int64_t n, int32_t dim,
                   const float* w /* [heads][dim] */, const float* bias /* [heads] or NULL */,
                   int32_t heads, float* out /* [n][heads] */);

/* reference jitter Monte Carlo (M:3449-3510): candidates come from the host (cheap, O(n L));
 * trials x candidates walks run on the device over shared high/low arrays.
 *   cand_*            [n_cand]   bar index, side (0 long/1 short), level price, entry price
 *   sigmas            [4]        level, entry, stop, target jitter std (M:3354-3355)
 *   replay_normals    NULL, or [trials][n_cand][4] N(0,1) draws   } test mode: host-drawn numbers,
 *   replay_uniforms   NULL, or [trials][n_cand]    U[0,1) draws   } exact fp64 replay of the reference
 *   out_*             [trials]   total R, max drawdown, wins, losses, opens                       */
int  qmmx_run_jitter(qmmx_ctx* ctx, const double* high, const double* low, int32_t n_bars,
                     const int32_t* cand_bar, const uint8_t* cand_side, const double* cand_level,
                     const double* cand_entry, int32_t n_cand,
                     double stop_padding, double tp_padding, const double* sigmas,
                     int64_t trials, uint64_t seed, uint64_t trial_offset,
                     const double* replay_normals, const double* replay_uniforms,
                     double* out_total_r, double* out_max_dd, int32_t* out_wins, int32_t* out_losses,
                     int32_t* out_opens);

/* deterministic walk of QMMXApp.simulate_last_bars (M:3619-3639): for each candidate the first bar after cand_bar
 * whose low/high reaches cand_stop or cand_target, the stop tested first on a bar that reaches both.
 *   cand_*            [n_cand]   bar index, side (0 long/1 short), stop and target prices (M:3610-3615)
 *   out_outcome       [n_cand]   0 = "tp", 1 = "stop", 2 = "open" (never hit)
 *   out_exit_bar      [n_cand]   index of the hit bar, -1 when open                                  */
int  qmmx_run_last_bars(qmmx_ctx* ctx, const double* high, const double* low, int32_t n_bars,
                        const int32_t* cand_bar, const uint8_t* cand_side, const double* cand_stop,
                        const double* cand_target, int32_t n_cand, int32_t* out_outcome, int32_t* out_exit_bar);

/* qmmx_run_last_bars without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same results and the same kernel as qmmx_run_last_bars; the device copy of
 * the bars and candidates and the two result arrays live in ONE file-scope `static __device__` array of 1 MiB inside the
 * library (every GPU that loads the library holds it until the process exits) instead of a cudaMalloc'ed arena.
 * Limits: n_bars <= QMMX_STATIC_MAX_BARS and n_cand <= QMMX_STATIC_MAX_CAND, beyond them QMMX_E_UNSUPPORTED with the
 * limits in qmmx_last_error (use qmmx_run_last_bars).  The call makes no cudaMalloc / cudaFree / allocating library
 * call; the ctx (qmmx_create resolves the array's device address once per device) must exist beforehand.
 * The array is shared by EVERY ctx and call on a device and keeps the last call's contents; nothing in it has to start
 * at zero (each call overwrites its inputs and writes both outputs of every candidate).  Calls that overlap on one GPU
 * (two ctxs, two threads, two streams) race on it: serialise them. */
#define QMMX_STATIC_MAX_BARS 12000
#define QMMX_STATIC_MAX_CAND 16384
int  qmmx_run_last_bars_static(qmmx_ctx* ctx, const double* high, const double* low, int32_t n_bars,
                               const int32_t* cand_bar, const uint8_t* cand_side, const double* cand_stop,
                               const double* cand_target, int32_t n_cand, int32_t* out_outcome, int32_t* out_exit_bar);

/* Host-side check of the exact price-classification tables the kernels use (no device needed):
 * for each fp32 price, level_table/level_brute = touched level by table lookup / by the float64 brute force
 * of M:1543-1545 (-1 = TOO_FAR); conf_mask bit b = confidence gate passes for touch bucket b (tc<=1, 2, >=3)
 * by threshold compare, conf_brute = the same by float64 compute_confidence; near_table/near_brute = bit j set
 * when p is inside level j's contact window (|price_j - p| <= prox). Any output may be NULL.
 * near_table comes from the CLASS partition the lean / tile kernels read (one record per price class = touched level +
 * the levels whose window holds the price: the latch release of M:1572-1576 as one AND) whenever the level set has one;
 * a class level that differs from level_table is reported as level_table = -100. */
int  qmmx_debug_classify(const double* price, const uint8_t* solid, int n_levels, double prox, double qmin,
                         const float* p, int64_t n, int32_t* level_table, int32_t* level_brute,
                         uint8_t* conf_mask, uint8_t* conf_brute, uint64_t* near_table, uint64_t* near_brute);

/* Sizes of the class partition of a level set (host only): classes, runs along the fp32 line, single-boundary records.
 * n_classes = -1: the set needs more than 254 classes or 640 runs and the first-generation kernels run instead. */
int  qmmx_debug_class_partition(const double* price, const uint8_t* solid, int n_levels, double prox, double qmin,
                                int32_t* n_classes, int32_t* n_runs, int32_t* n_records);

/* device time of the last run's kernels in ms (CUDA events on the ctx stream) */
int  qmmx_last_timing(qmmx_ctx* ctx, float* ms);
/* FP32 FFMA throughput of this device in TFLOP/s, measured now with an FMA-only kernel (best of 3 after a warm-up).
 * The denominator of the FP32/issue roofline that SURVEY 8(d) asks for beside the HBM one; no reference counterpart. */
int  qmmx_measure_fp32_peak(qmmx_ctx* ctx, double* tflops);
/* Issue-rate probes of this device, measured now (best of 3 after a warm-up), in 1e9 warp-instructions per second over the
 * whole device: *alu = a kernel of independent LOP3 chains (the ALU pipe: integer, logic, compare, select, shift -- the pipe
 * the scan kernels keep busiest); *mixed = the same chains interleaved one-to-one with FFMA chains (ALU + FMA pipes fed
 * together).  The ceilings `bench.py` quotes next to the scan's own instruction rates.  No reference counterpart. */
int  qmmx_measure_issue_peak(qmmx_ctx* ctx, double* alu_gwips, double* mixed_gwips);
/* number of kernel launches issued by this ctx since creation */
int  qmmx_launch_count(qmmx_ctx* ctx, uint64_t* n);

#ifdef __cplusplus
}
#endif
#endif /* QMMX_MC_H */
