"""Drop-in call shapes of the reference's simulators, running on the GPU library.

`monte_carlo(...)`  ~ `QMMXApp.simulate_monte_carlo` (qmmx_monolithic.py:3353-3538): candidate pass on the host
                      (O(n L), cheap), trials x candidates walk-forward on the device (`qmmx_run_jitter`),
                      summary with the reference's index rules and message text.
`last_bars(...)`    ~ `QMMXApp.simulate_last_bars` (M:3540-3699): same candidate pass, deterministic forward walks on
                      the device (`qmmx_run_last_bars`), the reference's rounding and summary rules.
`run_paths(...)`    ~ the live engine (`evaluate_entry` + exit block) over many price paths.

`levels` is the reference's list of dicts {color, type, index, price}; `bars` is the reference's list of
{t, o, h, l, c} dicts (M:220-240) or a dict of arrays.  This module never touches oracle/.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from statistics import mean, median, pstdev
from typing import Optional, Sequence

import numpy as np

from . import abi
from .engine import MCEngine, sort_levels


def _bar_arrays(bars, n):
    if isinstance(bars, dict):
        c, h, l = (np.asarray(bars[k], dtype=np.float64) for k in ("c", "h", "l"))
    else:
        c = np.array([float(b["c"]) for b in bars]); h = np.array([float(b["h"]) for b in bars])
        l = np.array([float(b["l"]) for b in bars])
    return c[-n:], h[-n:], l[-n:]


def _confidence(dist, solid, touch_count, prox):
    """`compute_confidence` (M:1415-1427) with the direction known (+0.03 always)."""
    base = max(0.0, 1.0 - (dist / max(1e-4, prox)))
    base += 0.08 if solid else 0.02
    if touch_count <= 1:
        base += 0.10
    elif touch_count == 2:
        base -= 0.08
    else:
        base -= 0.08 * 2
    base += 0.03
    return max(0.0, min(1.0, base))


class EntryGates:
    """The engine gate as the simulators call it (`evaluate_entry`, M:1492-1771, from M:3414-3426 and M:3582-3607).

    The reference runs it on the live engine object; the drop-in runs it on a fresh one: no open trade, no cooldown
    (`cooldown_until_ms` = 0 against a wall-clock `now_ms`), `last_direction` unknown and never updated (only the live
    loop writes it, M:2952-2955), accumulation guard never fed ("unknown": M:1589-1621 and M:1652-1666 inert), bars
    without volume (soft veto inert, M:1773-1794).  What carries over from candidate to candidate is the engine's
    contact latch and per-level touch count (M:1557-1587).  O(L) per call, a few hundred calls per run: host work.
    """

    def __init__(self, rows, prices, solid, *, contact_prox, stop_padding, q_min_prob=0.60, overtouch_limit=4, lr=None,
                 disable_ml_gate=False, use_blend=False, w_rules=0.7, w_ml=0.3):
        self.rows, self.price, self.solid = rows, [float(x) for x in prices], [bool(x) for x in solid]
        self.prox, self.sp, self.qmin = float(contact_prox), float(stop_padding), float(q_min_prob or 0.0)
        self.overtouch, self.lr = int(overtouch_limit), (None if lr is None else [float(x) for x in np.ravel(lr)])
        self.disable_ml_gate, self.use_blend = bool(disable_ml_gate), bool(use_blend)
        s = float(w_rules) + float(w_ml)                                  # M:1714-1719
        self.w_rules, self.w_ml = (1.0, 0.0) if s <= 0 else (float(w_rules) / s, float(w_ml) / s)
        self.touch = [0] * len(self.price)
        self.latch = [False] * len(self.price)

    def evaluate(self, p, prev):
        """-> (allowed, reason, message, side, k); `message` is the reference's own text."""
        lp, prox, qmin = self.price, self.prox, self.qmin
        if prev is not None and p > prev + 1e-9:                          # M:1529-1537
            up = True
        elif prev is not None and p < prev - 1e-9:
            up = False
        else:
            return False, "DIR_UNKNOWN", "Flat tick; cannot infer approach.", None, None
        k = min(range(len(lp)), key=lambda j: abs(lp[j] - p))             # M:1543, first minimum wins
        dist = abs(lp[k] - p)
        if dist > prox:
            r = self.rows[k]
            return False, "TOO_FAR", (f"Nearest level {r['color']}/{r['type']}[{r['index']}] @{lp[k]:.2f} "
                                      f"too far ({dist:.2f})."), None, k
        if not self.latch[k]:                                             # M:1562-1566
            self.touch[k] += 1
            self.latch[k] = True
        for j in range(len(lp)):                                          # M:1572-1576
            if j != k and self.latch[j] and abs(lp[j] - p) > prox:
                self.latch[j] = False
        tc = self.touch[k]
        if tc >= self.overtouch:                                          # M:1579
            return False, "LEVEL_OVERTOUCHED", f"Level over-touched (#{tc}).", None, k
        conf = _confidence(dist, self.solid[k], tc, prox)
        if conf < qmin:                                                   # M:1628
            return False, "CONF_LOW", f"Confidence {conf:.2f} < min {qmin:.2f}.", None, k
        side = 0 if up else 1
        mlp, ok_ml = conf, True
        if not self.disable_ml_gate and self.lr is not None:              # M:1454-1466: x = [1, |L - stop|, touches, up]
            stop = lp[k] - self.sp if side == 0 else lp[k] + self.sp
            w = self.lr
            z = w[4]
            for wi, xi in zip(w[:4], (1.0, abs(lp[k] - stop), float(tc), 1.0 if up else 0.0)):
                z = z + wi * xi
            mlp = 1.0 / (1.0 + math.exp(-z))                             # float64, left-to-right sum, libm exp (as build_ml_tables)
            ok_ml = mlp >= qmin
        if self.use_blend:                                                # M:1730-1742
            blended = self.w_rules * conf + self.w_ml * mlp
            if blended < qmin:
                return False, "COMBINED_LOW", f"Blended {blended:.2f} < min {qmin:.2f}", side, k
        elif not self.disable_ml_gate and not ok_ml:                      # M:1745-1755
            return False, "ML_CONF_LOW", f"ML prob {mlp:.2f} < min {qmin:.2f}", side, k
        return True, "OK", "Entry allowed.", side, k


def find_candidates(closes, level_prices, contact_prox, stop_padding, tp_padding, touch_limit, *, gates=None,
                    gate_order="mc", skips=None):
    """Candidate pass of the two simulators: nearest level to the close (first minimum wins), within CONTACT_PROX,
    side by close-to-close move (flat => short), per-level touch limit keyed on round(level, 4).
    gates = an EntryGates (with_gates=True) or None.  gate_order "mc" = `simulate_monte_carlo` (M:3385-3442): the touch
    limit comes BEFORE the gate and the candidate takes level and side from the gate's extras; "sim" =
    `simulate_last_bars` (M:3566-3612): the gate comes first and only the bars it lets through count against the limit.
    `skips` collects (i, reason, message, prox) of refused bars (the sim's GATE_SKIP audit rows, M:3594-3597)."""
    out, seen, prev = [], {}, None
    lp = [float(x) for x in level_prices]
    for i, c in enumerate(closes):
        c = float(c)
        if prev is not None:
            k = min(range(len(lp)), key=lambda j: abs(lp[j] - c))
            dist = abs(lp[k] - c)
            if dist <= contact_prox:
                side = 0 if c > prev else 1
                key = round(lp[k], 4)
                tc = seen.get(key, 0) + 1
                ok = True
                if gate_order == "mc":
                    ok = tc <= touch_limit
                    if ok:
                        seen[key] = tc
                if ok and gates is not None:
                    allowed, reason, msg, g_side, g_k = gates.evaluate(c, prev)
                    if not allowed:
                        ok = False
                        if skips is not None:
                            skips.append((i, reason, msg, dist))
                    elif gate_order == "mc":
                        side, k = g_side, g_k                     # M:3425-3426
                if ok and gate_order != "mc":
                    ok = tc <= touch_limit
                    if ok:
                        seen[key] = tc
                if ok:
                    out.append(dict(i=i, side=side, k=k, lvl=lp[k], entry=c, prox=dist, tc=tc,
                                    stop=lp[k] - stop_padding if side == 0 else lp[k] + stop_padding,
                                    target=lp[k] + tp_padding if side == 0 else lp[k] - tp_padding))
        prev = c
    return out


@dataclass
class MCSummary:
    """Fields of the reference's MC/SUMMARY message (M:3513-3526)."""
    candidates: int
    trials: int
    mean: float
    median: float
    stdev: float
    best: float
    worst: float
    var05: float
    cvar05: float
    dd_mean: float
    dd_worst: float
    wins: float
    losses: float
    opens: float
    totals: np.ndarray = field(repr=False, default=None)
    drawdowns: np.ndarray = field(repr=False, default=None)

    def message(self) -> str:
        return (f"Candidates: {self.candidates} | Trials: {self.trials}\n"
                f"Total R — mean {self.mean:+.2f}, median {self.median:+.2f}, stdev {self.stdev:.2f}\n"
                f"Best {self.best:+.2f} | Worst {self.worst:+.2f}\n"
                f"VaR(5%) {self.var05:+.2f} | CVaR(5%) {self.cvar05:+.2f}\n"
                f"Max drawdown (R) — mean {self.dd_mean:+.2f}, worst {self.dd_worst:+.2f}\n"
                f"Avg counts — wins {self.wins:.1f}, losses {self.losses:.1f}, open {self.opens:.1f}")


def summarize(totals, drawdowns, wins, losses, opens, n_cand) -> MCSummary:
    """M:3513-3526: p05 index = max(0, int(0.05 N) - 1); CVaR = mean of the sorted totals up to that index."""
    t = [float(x) for x in totals]
    dd = [float(x) for x in drawdowns]
    ts = sorted(t)
    p05 = max(0, int(0.05 * len(ts)) - 1)
    return MCSummary(n_cand, len(t), mean(t), median(t), pstdev(t), max(t), min(t), ts[p05], mean(ts[:p05 + 1]),
                     mean(dd), min(dd), mean(int(x) for x in wins), mean(int(x) for x in losses),
                     mean(int(x) for x in opens), np.asarray(totals), np.asarray(drawdowns))


def _gates(with_gates, rows, price, solid, contact_prox, stop_padding, gate_settings):
    return EntryGates(rows, price, solid, contact_prox=contact_prox, stop_padding=stop_padding, **gate_settings) if with_gates else None


def monte_carlo(eng: MCEngine, bars, levels: Sequence[dict], *, contact_prox=0.05, stop_padding=0.35, tp_padding=0.25,
                n=600, touch_limit=1, trials=500, with_gates=True, entry_slip_std=0.01, level_jitter_std=0.02,
                stop_slip_std=0.0, target_slip_std=0.0, seed: Optional[int] = None, replay=None,
                **gate_settings) -> Optional[MCSummary]:
    """Same arguments and defaults as `simulate_monte_carlo` (M:3353-3355), `with_gates=True` included: every bar that
    survives the touch limit goes through the engine gate (EntryGates; `gate_settings` = q_min_prob, overtouch_limit,
    lr = coef[4] + intercept, disable_ml_gate, use_blend, w_rules, w_ml: the engine settings the reference reads from
    its object).  The random source is the device's Philox stream (key = seed), or `replay` = (normals[trials][K][4],
    uniforms[trials][K]) to reproduce host-drawn numbers exactly.  Returns None when no candidate is found (the
    reference shows "No candidates found", M:3444-3446)."""
    rows, price, solid = sort_levels(levels)
    c, h, l = _bar_arrays(bars, n)
    cands = find_candidates(c, price, contact_prox, stop_padding, tp_padding, touch_limit, gate_order="mc",
                            gates=_gates(with_gates, rows, price, solid, contact_prox, stop_padding, gate_settings))
    if not cands:
        return None
    out = eng.run_jitter(h, l, cands, stop_padding, tp_padding,
                         (level_jitter_std, entry_slip_std, stop_slip_std, target_slip_std), trials,
                         seed=0 if seed is None else int(seed), replay=replay)
    return summarize(out["totals"], out["drawdowns"], out["wins"], out["losses"], out["opens"], len(cands))


@dataclass
class SimSummary:
    """Fields of the reference's SIM/SUMMARY line and "Sim complete" box (M:3650-3693)."""
    trades: int
    wins: int
    loss: int
    open: int
    total_R: float
    avg_R: float
    max_win_R: float
    max_loss_R: float
    max_profit_usd: float
    max_loss_usd: float

    def message(self) -> str:
        return (f"Trades: {self.trades} | Wins: {self.wins} Loss: {self.loss} Open: {self.open}\n"
                f"Total R: {self.total_R} | Avg R (closed): {self.avg_R}\n"
                f"Max Profit: ${self.max_profit_usd} | Max Loss: ${self.max_loss_usd}\n"
                f"Max Win (R): {self.max_win_R} | Max Loss (R): {self.max_loss_R}")


def last_bars(eng: MCEngine, bars, levels: Sequence[dict], *, contact_prox=0.05, stop_padding=0.35, tp_padding=0.25,
              n=1200, touch_limit=2, with_gates=False, gate_skips=None, **gate_settings):
    """Same arguments and defaults as `simulate_last_bars` (M:3540-3541; `with_gates` defaults to False there, the GUI
    button passes True, M:2655-2657): candidate pass on the host (the gate runs BEFORE the touch limit, M:3582-3607),
    the forward walks on the device (`qmmx_run_last_bars`, stop tested before target), then the reference's per-trade
    `round(., 2)` and summary rules (M:3641-3666).  Returns (trades, SimSummary); the trade dicts carry the reference's
    keys plus `j` (exit bar, -1 when open).  `gate_skips` (a list) receives the SIM audit rows of the refused bars,
    ("SIM", "GATE_SKIP:<code>", "<reason> prox=<d>") as M:3596-3597 writes them; `gate_settings` as in monte_carlo."""
    rows, price, solid = sort_levels(levels)
    c, h, l = _bar_arrays(bars, n)
    skips = []
    cands = find_candidates(c, price, contact_prox, stop_padding, tp_padding, touch_limit, gate_order="sim", skips=skips,
                            gates=_gates(with_gates, rows, price, solid, contact_prox, stop_padding, gate_settings))
    if gate_skips is not None:       # the reference's code for the two sub-reasons is CONF_LOW (M:1742, M:1755)
        gate_skips.extend(("SIM", "GATE_SKIP:" + ("CONF_LOW" if r in ("ML_CONF_LOW", "COMBINED_LOW") else r),
                           f"{msg} prox={d:.03f}") for _, r, msg, d in skips)
    outcome, exit_bar = eng.run_last_bars(h, l, cands) if cands else ((), ())
    trades = []
    for cand, oc, xb in zip(cands, outcome, exit_bar):
        entry, stop, target = cand["entry"], cand["stop"], cand["target"]
        if oc == 2:
            name, exit_px, r = "open", float(c[-1]), 0.0
        else:
            name = "tp" if oc == 0 else "stop"
            exit_px = target if oc == 0 else stop
            risk = abs(entry - stop) or 1e-9
            r = (abs(target - entry) / risk) if oc == 0 else -1.0
        trades.append({"i": cand["i"], "j": int(xb), "side": "long" if cand["side"] == 0 else "short",
                       "lvl": round(cand["lvl"], 2), "entry": round(entry, 2), "stop": round(stop, 2),
                       "target": round(target, 2), "exit": round(exit_px, 2), "outcome": name, "R": round(r, 2),
                       "prox": round(cand["prox"], 3), "touch#": cand["tc"]})
    closed = [t for t in trades if t["outcome"] in ("tp", "stop")]

    def usd(t):
        return (t["exit"] - t["entry"]) * (1.0 if t["side"] == "long" else -1.0)
    s = SimSummary(
        trades=len(trades), wins=sum(1 for t in trades if t["outcome"] == "tp"),
        loss=sum(1 for t in trades if t["outcome"] == "stop"), open=sum(1 for t in trades if t["outcome"] == "open"),
        total_R=round(sum(t["R"] for t in trades), 2),
        avg_R=round(sum(t["R"] for t in closed) / max(1, len(closed)), 2),
        max_win_R=round(max((t["R"] for t in closed), default=0.0), 2),
        max_loss_R=round(min((t["R"] for t in closed), default=0.0), 2),
        max_profit_usd=round(max((usd(t) for t in closed), default=0.0), 2),
        max_loss_usd=round(min((usd(t) for t in closed), default=0.0), 2))
    return trades, s


def run_paths(eng: MCEngine, close, levels: Sequence[dict], params, *, high=None, low=None, lr=None, policy=None):
    """Engine over host price paths `[steps][paths]` (fp32): configure, upload, scan; one MCResult per combo."""
    eng.set_levels(levels)
    eng.set_params(params)
    eng.set_lr(lr)
    eng.set_policy(policy)
    eng.upload_bars(close, high, low)
    return eng.run_scan()


# ---- results sink (second half): reason-code histograms instead of policy_events rows ----------
# The reference writes one `policy_events` row per evaluated tick (M:1468-1490) and `log_analyzer.py:23-73` folds the
# `phase='entry' AND action='skip'` rows into counts per `reason`.  The device already produces those counts
# (qmmx_agg.entry_hist), so the sink is columnar: one row per parameter combo, one column per reason.

# reasons the reference files under action='policy_skip' (M:1658, M:1694, M:3097): log_analyzer's query does not see them
POLICY_SKIP_REASONS = ("ACC_BREAKOUT_GATE", "ONLINE_POLICY", "VETO_CONTRA_VOL_LONG", "VETO_CONTRA_VOL_SHORT",
                       "VETO_EDGE_FATIGUE", "VETO_BOUNCE_BUDGET", "VETO_LEVEL_COOLDOWN")


def skip_reason_counts(result, policy_skips: bool = False) -> dict:
    """reason -> count of the entry-phase skips, the fold `log_analyzer.py:14-31` computes from the action='skip'
    rows; `policy_skips=True` adds the reasons the reference writes as action='policy_skip' rows."""
    return {name: n for name, n in result.entry_hist.items()
            if name != "OK" and n > 0 and (policy_skips or name not in POLICY_SKIP_REASONS)}


def skip_report(result) -> str:
    """The report text of `log_analyzer.py:17-73` (header, then reasons by descending count).  The per-reason
    detail lines of the reference need the per-event extras, which the aggregate does not keep."""
    counts = skip_reason_counts(result)
    total = sum(counts.values())
    if not total:
        return "No 'entry' phase skip events found in the policy_events table."
    lines = [f"Found {total} 'entry' phase skip events.\n", "--- Summary of Skip Reasons ---"]
    for reason, n in sorted(counts.items(), key=lambda kv: kv[1], reverse=True):
        lines.append(f"Reason: {reason} (Count: {n})")
        lines.append("")
    return "\n".join(lines)


def results_columns(results, params=None) -> dict:
    """Column arrays over the combos of one run: the qmmx_agg counters, PnL moments and both reason histograms
    (`entry_<REASON>`, `exit_<REASON>`), plus the swept settings when `params` (the list given to set_params) is passed."""
    results = list(results)
    cols = {k: np.array([getattr(r, k) for r in results], dtype=np.int64)
            for k in ("n_paths", "n_steps", "n_trades", "n_wins", "n_open")}
    for k in ("sum_pnl", "sum_pnl2", "sum_path_pnl2", "sum_open_mtm", "max_path_pnl", "min_path_pnl"):
        cols[k] = np.array([float(getattr(r.agg, k)) for r in results], dtype=np.float64)
    for i, name in enumerate(abi.ENTRY_CODES):
        cols["entry_" + name] = np.array([int(r.agg.entry_hist[i]) for r in results], dtype=np.int64)
    for i, name in enumerate(abi.EXIT_CODES):
        cols["exit_" + name] = np.array([int(r.agg.exit_hist[i]) for r in results], dtype=np.int64)
    if params is not None:
        plist = list(params) if isinstance(params, (list, tuple)) else [params]
        if len(plist) != len(results):
            raise ValueError("params and results differ in length")
        for k in ("contact_prox", "stop_padding", "tp_padding", "q_min_prob", "cooldown_s"):
            cols[k.upper()] = np.array([float(getattr(p, k)) for p in plist], dtype=np.float64)
    return cols


def write_results(path: str, results, params=None) -> str:
    """Write `results_columns` as one Parquet file (pyarrow).  Replaces the audit_log / policy_events tables for runs
    whose tick count (10^8 .. 10^11) rules out a row per tick."""
    import pyarrow as pa
    import pyarrow.parquet as pq
    cols = results_columns(results, params)
    pq.write_table(pa.table({k: pa.array(v) for k, v in cols.items()}), path)
    return path
