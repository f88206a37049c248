"""TEST INFRASTRUCTURE ONLY — headless scalar restatement of the QMMX trade engine.

This is the parity oracle for the CUDA path (`qmmx_monolithic_monte_carlo_b200`).
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl
reference` legs may import it; the product never does.

Parity status: PINNED — `tests/test_oracle_vs_reference.py` drives the unmodified
reference `MonolithicEngine.evaluate_entry` (stub-imported, see `oracle/ref_harness.py`)
step by step against this file on random paths, and `tests/golden/` holds vectors
produced by that reference run (`oracle/make_golden.py`) plus the known answers
recovered from the reference's `qmmx.db-wal`.
Exception: the sklearn-form LR gate has no reference artefact (no `qmmx_lr.joblib`
ships) — pinned only against sklearn's `predict_proba` with synthetic coefficients.

Pure Python, float64, no GUI / network / SQLite.  `M:` = /root/reference/qmmx_monolithic.py.
Frozen composition: tick i = bar i's close, now_ms = 60000*i,
evaluate_entry once per tick, escalation off, acc-guard regime "unknown".
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

# ---- reason codes (M:246-257 plus the sub-reasons written into policy_events) ----
ENTRY_CODES = [
    "OK", "IN_POSITION", "COOLDOWN", "NOLEVELS", "DIR_UNKNOWN", "TOO_FAR",
    "LEVEL_OVERTOUCHED", "CONF_LOW", "ML_CONF_LOW", "COMBINED_LOW",
    "VETO_CONTRA_VOL_LONG", "VETO_CONTRA_VOL_SHORT", "VETO_EDGE_FATIGUE",
    "VETO_BOUNCE_BUDGET", "VETO_LEVEL_COOLDOWN", "ACC_BREAKOUT_GATE",
    "ONLINE_POLICY", "PRICE_STALE", "MISSING_API_KEY", "RISK_INVALID",
]
EXIT_CODES = ["STOP", "TARGET", "ESCALATE", "OPEN"]
E = {n: i for i, n in enumerate(ENTRY_CODES)}
X = {n: i for i, n in enumerate(EXIT_CODES)}
N_ENTRY_CODES = len(ENTRY_CODES)
N_EXIT_CODES = len(EXIT_CODES)

COLOR_ORDER = {"black": 0, "blue": 1, "orange": 2, "teal": 3}


@dataclass
class Params:
    """Engine settings (defaults: M:1371-1383, M:1429-1447, M:1711-1715, M:3090)."""
    contact_prox: float = 0.05        # CONTACT_PROX
    stop_padding: float = 0.35        # STOP_PADDING
    tp_padding: float = 0.25          # TP_PADDING
    q_min_prob: float = 0.60          # Q_MIN_PROB
    cooldown_s: int = 8               # Q_SIGNAL_COOLDOWN (seconds)
    enable_veto: bool = False         # ENABLE_VETO (inert without volume, M:1773-1811)
    veto_vol_strong: float = 0.25
    veto_prox: float = 0.06
    disable_ml_gate: bool = False     # DISABLE_ML_GATE
    use_blend: bool = False           # USE_BLEND
    w_rules: float = 0.7
    w_ml: float = 0.3
    disable_policy_gate: bool = True  # DISABLE_POLICY_GATE=1 (recorded run: policy=False)
    exit_on_hl: bool = False          # [CHOICE] False: live exits on close (M:2966-3014)
    overtouch_limit: int = 4          # hard-coded M:1579
    steps_per_day: int = 390
    online_retrain: bool = False      # OnlinePolicy learns from its own closed trades (M:3753-3803), see run_path
    policy_lr: float = 0.03           # OnlinePolicy(lr=0.03, l2=1e-6), M:290
    policy_l2: float = 1e-6


@dataclass
class Level:
    price: float
    solid: bool
    key: tuple = ()


def sort_levels(levels: Sequence[dict]) -> List[Level]:
    """`load_levels` order: ORDER BY color, level_type, level_index (M:140-144)."""
    rows = sorted(levels, key=lambda lv: (str(lv["color"]), str(lv["type"]), int(lv["index"])))
    return [Level(float(lv["price"]), str(lv["type"]) == "solid",
                  (str(lv["color"]), str(lv["type"]), int(lv["index"]))) for lv in rows]


@dataclass
class LRModel:
    """sklearn-form logistic regression, inference side of M:1454-1466: 4 coef + intercept."""
    coef: Sequence[float]
    intercept: float


@dataclass
class Trade:
    entry_bar: int
    exit_bar: int
    side: int          # 0 long, 1 short
    level: int         # index into the sorted level list
    reason: int        # EXIT_CODES id
    entry: float
    exit: float
    pnl: float         # realised; 0 for OPEN (M:3630-3633)
    mtm: float = 0.0   # mark-to-market for OPEN (M:3266-3270)


@dataclass
class PathResult:
    codes: List[int] = field(default_factory=list)
    trades: List[Trade] = field(default_factory=list)
    entry_hist: List[int] = field(default_factory=lambda: [0] * N_ENTRY_CODES)
    exit_hist: List[int] = field(default_factory=lambda: [0] * N_EXIT_CODES)
    touch_counts: List[int] = field(default_factory=list)
    policy_weights: Optional[dict] = None      # online_retrain: the OnlinePolicy entry heads at the end of the path


def dot_py312(w: Sequence[float], x: Sequence[float]) -> float:
    """`OnlinePolicy._dot` = builtin `sum(wi*xi ...)` (M:334-335) as evaluated by CPython >= 3.12,
    whose float `sum` is Neumaier-compensated (Python/bltinmodule.c).  Spelled out so the contract
    does not depend on the interpreter running the oracle; tests assert it equals `sum` here."""
    f = 0 + w[0] * x[0]
    c = 0.0
    for wi, xi in zip(w[1:], x[1:]):
        v = wi * xi
        t = f + v
        if abs(f) >= abs(v):
            c += (f - t) + v
        else:
            c += (v - t) + f
        f = t
    if c and math.isfinite(c):
        f += c
    return f


def sigmoid_policy(x: float) -> float:
    """M:20-26."""
    if x < -50:
        return 0.0
    if x > 50:
        return 1.0
    return 1.0 / (1.0 + math.exp(-x))


def compute_confidence(dist: float, solid: bool, touch_count: int, prox: float) -> float:
    """M:1415-1427 (direction is always known where this is called, so +0.03 applies)."""
    base = max(0.0, 1.0 - (dist / max(0.0001, prox)))
    base += (0.08 if solid else 0.02)
    if touch_count <= 1:
        base += 0.10
    elif touch_count == 2:
        base -= 0.08
    else:
        base -= (0.08 * 2)
    base += 0.03
    return float(max(0.0, min(1.0, base)))


def lr_prob(model: LRModel, lvl_type: float, distf: float, touch: float, direc: float) -> float:
    """sigma(b + w.x), x = [lvl_type, |level-stop|, touch_count, dir_up] (M:1457-1463).

    Contract: z accumulated left to right in float64 without fused multiply-add.
    """
    z = float(model.intercept)
    z = z + float(model.coef[0]) * lvl_type
    z = z + float(model.coef[1]) * distf
    z = z + float(model.coef[2]) * touch
    z = z + float(model.coef[3]) * direc
    return 1.0 / (1.0 + math.exp(-z))


def volume_slope(vols: Sequence[float], window: int = 6) -> float:
    """M:1796-1811 on a plain list of bar volumes (oldest first)."""
    if not vols or len(vols) < 3:
        return 0.0
    last = [float(v or 0.0) for v in vols[-min(window, len(vols)):]]
    half = max(2, len(last) // 2)
    v1 = sum(last[:half]) / half
    v2 = sum(last[-half:]) / half
    if v1 == 0 and v2 == 0:
        return 0.0
    return (v2 - v1) / (abs(v1) + 1e-9)


class RegimeGuard:
    """AccumulationBreakoutGuard (M:1241-1356) as the live loop feeds it: `ingest_tick` (M:1857-1883, called at
    M:2932 before `evaluate_entry`) rolls one-tick minute bars, so at tick i the guard has seen bars 0..i-1 with
    o = h = l = c = that tick's price and v = that tick's volume [CHOICE: one tick per minute, Appendix A].

    Regimes: 0 unknown, 1 accumulation, 2 breakout_up, 3 breakout_down.  `allow_trade` (M:1338-1349) blocks SHORTs
    in breakout_up and LONGs in breakout_down (code 904, M:1652-1666).

    LevelTouchMemory (M:1112-1239) is consulted in the accumulation regime (M:1589-1621) but is inert in the
    reference: `_on_minute_close` passes `levels=self.levels_cache` (dicts) to `register_touch_if_any`, whose
    `round(float(L), 2)` raises TypeError on the first level inside a `try/except: pass` (M:1838-1851), so no touch
    and no edge tap is ever recorded; `edge_fatigued` returns None and `allow_trade_at` returns (True, "ok", 1.0).
    tests/test_oracle_vs_reference.py::test_touch_memory_is_inert_in_the_reference pins that.
    """
    UNKNOWN, ACCUMULATION, BREAKOUT_UP, BREAKOUT_DOWN = 0, 1, 2, 3

    def __init__(self):
        self.bars: List[tuple] = []          # (ts_ms, high, low, close, volume); deque(maxlen=600) >= the 61 ever read
        self.regime = 0
        self.box_low: Optional[float] = None
        self.box_high: Optional[float] = None
        self.inside = 0

    def push(self, ts_ms: int, h: float, l: float, c: float, v: float) -> None:
        self.bars.append((ts_ms, h, l, c, v))
        if len(self.bars) > 600:
            self.bars = self.bars[-600:]
        window = [b for b in self.bars if ts_ms - b[0] <= 60 * 60 * 1000]          # M:1273-1278
        if len(window) < 30:                                                         # M:1288-1292
            self.regime, self.box_low, self.box_high, self.inside = 0, None, None, 0
            return
        vols = [b[4] for b in window]
        close = window[-1][3]
        box_low, box_high = min(b[2] for b in window), max(b[1] for b in window)
        thresh = close * (18 / 10000.0)                                              # compression_bp = 18
        ma_s = sum(vols[-5:]) / float(5)                                             # _ma, M:1280-1284 (len >= 30)
        ma_l = sum(vols[-20:]) / float(20)
        breakout = self.regime in (2, 3)
        if box_high - box_low <= max(1e-6, thresh):                                  # M:1311-1322
            if not breakout:
                self.regime = 1
            self.box_low, self.box_high = box_low, box_high
        elif not breakout:
            self.regime = 0
        if self.box_low is not None and ma_s and ma_l:                               # M:1325-1334
            if close > self.box_high + 1e-6 and ma_s > 1.40 * ma_l:
                self.regime, self.inside = 2, 0
            elif close < self.box_low - 1e-6 and ma_s > 1.40 * ma_l:
                self.regime, self.inside = 3, 0
        if self.regime in (2, 3) and self.box_low is not None:                       # M:1337-1344
            if self.box_low <= close <= self.box_high:
                self.inside += 1
                if self.inside >= 3:
                    self.regime = 1
            else:
                self.inside = 0


def regime_series(close: Sequence[float], volume: Sequence[float]) -> List[int]:
    """Regime seen by `evaluate_entry` at every tick (the guard has been fed bars 0..i-1)."""
    g = RegimeGuard()
    out = []
    for i in range(len(close)):
        if i >= 1:
            c = float(close[i - 1])
            g.push(60000 * (i - 1), c, c, c, float(volume[i - 1] or 0.0))
        out.append(g.regime)
    return out


def run_path(close: Sequence[float], levels: Sequence[Level], params: Params, *,
             lr: Optional[Sequence[Optional[LRModel]]] = None,
             policy: Optional[Sequence[dict]] = None,
             high: Optional[Sequence[float]] = None, low: Optional[Sequence[float]] = None,
             volume: Optional[Sequence[float]] = None) -> PathResult:
    """One path through the engine: gates M:1492-1771, tick order/exits M:2927-3014,
    policy gate M:3046-3115, pnl M:1926.

    lr: per-day list of LRModel (or None = no model, M:1455); day = i // steps_per_day,
        clamped to the last entry.
    policy: per-day list of {"go_long","go_short","skip"} weight vectors (7 each).

    online_retrain: every trade leaves TWO policy_events rows carrying its id: the PRE_OPEN row that
    `evaluate_entry` writes on OK (M:1757-1768: proximity_abs, approach, touch_count, conf, ml_prob), which
    `open_trade` attaches to the trade (M:1899-1911), and the {"x": x} row of the live loop (M:3139-3150).
    `close_trade` labels both pnl > 0 (M:1936-1941) and `_retrain_from_labeled_events` (M:3753-3803, Tk timer
    M:3824-3828) feeds the labelled rows, once each and in id order, to `update_entry` (M:357-363: one logistic SGD
    step with L2 on the head of the action taken).  The retrain rebuilds the features from the row's keys: the first
    row gives [1, min(1, |p - L|), 0, from_above, from_below, 0, 0] (no confluence, no minutes key), the second has
    none of the keys (only "x"), so it trains on the defaults [1, 0, 0, 1, 0, 0, 0] -- reproduced here, it is what
    the reference computes.  [CHOICE] the 2-minute timer is modelled as "before every tick": a label written at
    tick i is learnt before tick i+1 is evaluated.  `policy` must then hold ONE
    table: the initial weights; what is learnt carries over the day boundaries.
    """
    P = params
    L = len(levels)
    res = PathResult()
    prev: Optional[float] = None
    last_dir = 0            # 0 unknown, 1 up, 2 down   (EngineState.last_direction)
    cooldown_until = 0      # ms
    pos = None
    touch = [0] * L
    latch = [False] * L
    prox = float(P.contact_prox)
    qmin = float(P.q_min_prob or 0.0)
    # M:1714-1719 weight normalisation
    w_rules = float(P.w_rules)   # `float(settings_get(..) or 0.7)`: the `or` guards the *string*, "0" stays 0
    w_ml = float(P.w_ml)
    s = w_rules + w_ml
    if s <= 0:
        w_rules, w_ml, s = 1.0, 0.0, 1.0
    w_rules, w_ml = w_rules / s, w_ml / s

    guard = RegimeGuard() if volume is not None else None    # always on in the reference; inert without volume
    learn = bool(P.online_retrain) and not P.disable_policy_gate
    if learn:
        if policy is not None and len(policy) != 1:
            raise ValueError("online_retrain starts from ONE weight table")
        w0 = policy[0] if policy else {}
        wlearn = {a: [float(v) for v in w0.get(a, [0.0] * 7)] for a in ("go_long", "go_short", "skip")}
        policy = [wlearn]
    pending = []            # closed trades whose two labelled rows are not trained yet: (action, label, |p - L| at entry, side)
    for i, p in enumerate(close):
        p = float(p)
        now = 60000 * i
        day = i // P.steps_per_day
        for action, label, dist0, side0 in pending:                   # M:3771-3790 -> update_entry M:357-363
            w = wlearn[action]
            for xr in ([1.0, min(1.0, dist0), 0.0, 1.0 if side0 == 1 else 0.0, 1.0 if side0 == 0 else 0.0, 0.0, 0.0],
                       [1.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0]):          # PRE_OPEN row, then the {"x": ...} row
                pred = sigmoid_policy(dot_py312(w, xr))
                grad = [(pred - label) * xi + P.policy_l2 * wi for wi, xi in zip(w, xr)]     # M:339
                for q in range(7):
                    w[q] -= P.policy_lr * grad[q]
        pending = []
        if guard is not None and i >= 1:                          # ingest_tick closes bar i-1 (M:2932, M:1864-1870)
            cb = float(close[i - 1])
            guard.push(60000 * (i - 1), cb, cb, cb, float(volume[i - 1] or 0.0))
        code = None
        ok_extras = None
        # ---- evaluate_entry (first matching rule returns) ----
        if pos is not None:
            code = E["IN_POSITION"]                                   # M:1508
        elif now < cooldown_until:
            code = E["COOLDOWN"]                                      # M:1516
        elif L == 0:
            code = E["NOLEVELS"]                                      # M:1524
        else:
            d = 0
            if prev is not None:                                      # M:1529-1537
                if p > prev + 1e-9:
                    d = 1
                elif p < prev - 1e-9:
                    d = 2
                else:
                    d = last_dir
            if d == 0:
                code = E["DIR_UNKNOWN"]                               # M:1538
            else:
                k = 0
                dist = abs(levels[0].price - p)
                for j in range(1, L):                                 # M:1543 first-wins min
                    dj = abs(levels[j].price - p)
                    if dj < dist:
                        dist, k = dj, j
                if dist > prox:
                    code = E["TOO_FAR"]                               # M:1545
                else:
                    if not latch[k]:                                  # M:1562-1566
                        touch[k] += 1
                        latch[k] = True
                    for j in range(L):                                # M:1572-1576
                        if j != k and latch[j] and abs(levels[j].price - p) > prox:
                            latch[j] = False
                    tc = touch[k]
                    if tc >= P.overtouch_limit:
                        code = E["LEVEL_OVERTOUCHED"]                 # M:1579
                    else:
                        conf = compute_confidence(dist, levels[k].solid, tc, prox)
                        if conf < qmin:
                            code = E["CONF_LOW"]                      # M:1628
                        else:
                            lp = levels[k].price
                            if d == 1:                                # M:1669-1674
                                stop, target, side = lp - P.stop_padding, lp + P.tp_padding, 0
                            else:
                                stop, target, side = lp + P.stop_padding, lp - P.tp_padding, 1
                            code = E["OK"]
                            if guard is not None and ((guard.regime == 2 and side == 1) or
                                                      (guard.regime == 3 and side == 0)):
                                code = E["ACC_BREAKOUT_GATE"]         # M:1652-1666 (int code 904)
                            if code == E["OK"] and P.enable_veto:     # M:1678-1705
                                vs = volume_slope(volume[max(0, i - 239):i + 1]) if volume is not None else 0.0
                                confl = sum(1 for lv in levels if abs(lv.price - lp) <= 0.15) >= 2
                                if not (abs(vs) < 0.05 and not confl):
                                    near = dist <= max(P.veto_prox, prox * 0.12)
                                    strong = P.veto_vol_strong
                                    if d == 1:
                                        if near and side == 0 and vs < -strong:
                                            code = E["VETO_CONTRA_VOL_LONG"]
                                        elif near and side == 1 and vs > strong:
                                            code = E["VETO_CONTRA_VOL_SHORT"]
                                    else:
                                        if near and side == 0 and vs > strong:
                                            code = E["VETO_CONTRA_VOL_LONG"]
                                        elif near and side == 1 and vs < -strong:
                                            code = E["VETO_CONTRA_VOL_SHORT"]
                            if code == E["OK"]:                       # M:1707-1756
                                model = None
                                if lr:
                                    model = lr[min(day, len(lr) - 1)]
                                mlp = conf
                                ok_ml = True
                                if not P.disable_ml_gate and model is not None:
                                    distf = abs(lp - stop)
                                    prob = lr_prob(model, 1.0, distf, float(tc), 1.0 if d == 1 else 0.0)
                                    ok_ml = prob >= qmin
                                    mlp = prob
                                if P.use_blend:
                                    blended = w_rules * conf + w_ml * mlp
                                    if blended < qmin:
                                        code = E["COMBINED_LOW"]
                                elif (not P.disable_ml_gate) and not ok_ml:
                                    code = E["ML_CONF_LOW"]
                            if code == E["OK"]:
                                ok_extras = (side, k, stop, target, dist)
        # ---- state update (M:2952-2955) ----
        pprev = prev
        prev = p
        if pprev is not None and p != pprev:
            last_dir = 1 if p > pprev else 2
        # ---- exits for a position opened on an earlier tick (M:2966-3014) ----
        if pos is not None:
            side, stop, target = pos["side"], pos["stop"], pos["target"]
            reason = None
            exit_px = p
            if P.exit_on_hl:   # [CHOICE] mirrors simulate_last_bars M:3621-3636
                hh, ll = float(high[i]), float(low[i])
                if side == 0:
                    if ll <= stop:
                        reason, exit_px = X["STOP"], stop
                    elif hh >= target:
                        reason, exit_px = X["TARGET"], target
                else:
                    if hh >= stop:
                        reason, exit_px = X["STOP"], stop
                    elif ll <= target:
                        reason, exit_px = X["TARGET"], target
            else:
                if side == 0:
                    if p <= stop:
                        reason = X["STOP"]
                    elif p >= target:
                        reason = X["TARGET"]
                else:
                    if p >= stop:
                        reason = X["STOP"]
                    elif p <= target:
                        reason = X["TARGET"]
            if reason is not None:
                entry = pos["entry"]
                pnl = (exit_px - entry) if side == 0 else (entry - exit_px)   # M:1926
                cooldown_until = now + P.cooldown_s * 1000                    # M:269-270
                res.trades.append(Trade(pos["entry_bar"], i, side, pos["level"], reason,
                                        entry, exit_px, pnl))
                res.exit_hist[reason] += 1
                if learn:
                    pending.append(("go_long" if side == 0 else "go_short", 1 if pnl > 0 else 0, pos["dist"], side))   # M:1936
                pos = None
        elif ok_extras is not None:
            side, k, stop, target, dist = ok_extras
            pass_gate = True
            if not P.disable_policy_gate:                                     # M:3046-3093
                lp = levels[k].price
                w = policy[min(day, len(policy) - 1)] if policy else None
                confl = sum(1 for lv in levels if abs(lv.price - lp) <= 0.6) > 1
                xf = [1.0, min(1.0, abs(p - lp)), 0.0,
                      1.0 if side == 1 else 0.0, 1.0 if side == 0 else 0.0,
                      1.0 if confl else 0.0, min(1.0, (i % P.steps_per_day) / 390.0)]
                if w is None:
                    w = {"go_long": [0.0] * 7, "go_short": [0.0] * 7, "skip": [0.0] * 7}
                s_go = sigmoid_policy(dot_py312(w["go_long" if side == 0 else "go_short"], xf))
                s_skip = sigmoid_policy(dot_py312(w["skip"], xf))
                pass_gate = (s_go >= 0.60) and (s_skip < 0.55)
            if not pass_gate:
                code = E["ONLINE_POLICY"]
            else:
                pos = dict(side=side, entry=p, stop=stop, target=target, level=k, entry_bar=i, dist=dist)
        res.codes.append(code)
        res.entry_hist[code] += 1

    if pos is not None:   # end of path: OPEN, realised 0, MTM separately  [CHOICE]
        last = float(close[-1])
        mtm = (last - pos["entry"]) if pos["side"] == 0 else (pos["entry"] - last)
        res.trades.append(Trade(pos["entry_bar"], -1, pos["side"], pos["level"], X["OPEN"],
                                pos["entry"], last, 0.0, mtm))
        res.exit_hist[X["OPEN"]] += 1
    res.touch_counts = touch
    res.policy_weights = wlearn if learn else None
    return res


# ---------------------------------------------------------------------------------
# Reference jitter Monte Carlo (M:3353-3538) and deterministic sim (M:3540-3699)
# ---------------------------------------------------------------------------------

def nearest_level_first(levels: Sequence[Level], px: float):
    k = 0
    dist = abs(levels[0].price - px)
    for j in range(1, len(levels)):
        dj = abs(levels[j].price - px)
        if dj < dist:
            dist, k = dj, j
    return k, dist


class SimGates:
    """`MonolithicEngine.evaluate_entry` (M:1492-1771) as the two simulators drive it (M:3414-3426, M:3582-3607).

    The simulators call the gate on a FRESH engine: no open trade (M:1508 never fires), `cooldown_until_ms`
    = 0 against a wall-clock `now_ms` (M:1516 never fires), `last_ts_ms = now_ms` (fresh), `state.last_direction` = None
    and never updated (only the live loop writes it, M:2952-2955), so a flat close is DIR_UNKNOWN (M:1538).  What DOES
    persist from candidate to candidate is the engine's own contact latch and per-level touch count (M:1557-1587):
    they are advanced only by the candidates that reach the gate.  The accumulation guard is never fed by the simulators
    (`ingest_tick` is the live loop's, M:2932), so its regime stays "unknown" and M:1589-1621 / M:1652-1666 are inert; the
    bars carry no volume (M:234), so the soft veto never fires (M:1773-1794 with slope 0).
    evaluate() returns (allowed, reason, message, side, k): `reason` is the step code (sub-reasons ML_CONF_LOW /
    COMBINED_LOW kept apart), `message` the reference's own text (the SIM audit row GATE_SKIP quotes it, M:3596-3597).
    """

    def __init__(self, levels: Sequence[Level], params: Params, lr: Optional[LRModel] = None):
        self.levels, self.P, self.lr = levels, params, lr
        self.touch = [0] * len(levels)
        self.latch = [False] * len(levels)
        w_rules, w_ml = float(params.w_rules), float(params.w_ml)
        s = w_rules + w_ml
        if s <= 0:
            w_rules, w_ml, s = 1.0, 0.0, 1.0
        self.w_rules, self.w_ml = w_rules / s, w_ml / s

    def evaluate(self, p: float, prev: Optional[float]):
        P, levels = self.P, self.levels
        prox, qmin = float(P.contact_prox), float(P.q_min_prob or 0.0)
        d = 0
        if prev is not None:                                          # M:1529-1537, last_direction is None
            if p > prev + 1e-9:
                d = 1
            elif p < prev - 1e-9:
                d = 2
        if d == 0:
            return False, "DIR_UNKNOWN", "Flat tick; cannot infer approach.", None, None
        k, dist = nearest_level_first(levels, p)
        lv = levels[k]
        if dist > prox:                                               # unreachable from the simulators (same test, M:3398)
            return False, "TOO_FAR", (f"Nearest level {lv.key[0]}/{lv.key[1]}[{lv.key[2]}] @{lv.price:.2f} "
                                      f"too far ({dist:.2f})."), None, k
        if not self.latch[k]:                                         # M:1562-1566
            self.touch[k] += 1
            self.latch[k] = True
        for j in range(len(levels)):                                  # M:1572-1576
            if j != k and self.latch[j] and abs(levels[j].price - p) > prox:
                self.latch[j] = False
        tc = self.touch[k]
        if tc >= P.overtouch_limit:                                   # M:1579
            return False, "LEVEL_OVERTOUCHED", f"Level over-touched (#{tc}).", None, k
        conf = compute_confidence(dist, lv.solid, tc, prox)
        if conf < qmin:                                               # M:1628
            return False, "CONF_LOW", f"Confidence {conf:.2f} < min {qmin:.2f}.", None, k
        side = 0 if d == 1 else 1
        stop = lv.price - P.stop_padding if side == 0 else lv.price + P.stop_padding
        mlp, ok_ml = conf, True
        if not P.disable_ml_gate and self.lr is not None:             # M:1454-1466
            prob = lr_prob(self.lr, 1.0, abs(lv.price - stop), float(tc), 1.0 if d == 1 else 0.0)
            ok_ml, mlp = prob >= qmin, prob
        if P.use_blend:                                               # M:1730-1742
            blended = self.w_rules * conf + self.w_ml * mlp
            if blended < qmin:
                return False, "COMBINED_LOW", f"Blended {blended:.2f} < min {qmin:.2f}", side, k
        elif (not P.disable_ml_gate) and not ok_ml:                   # M:1745-1755
            return False, "ML_CONF_LOW", f"ML prob {mlp:.2f} < min {qmin:.2f}", side, k
        return True, "OK", "Entry allowed.", side, k


def find_candidates(bars_c: Sequence[float], levels: Sequence[Level], params: Params, *,
                    touch_limit: int, gate_order: str = "mc", with_gates: bool = False,
                    lr: Optional[LRModel] = None, skips: Optional[list] = None):
    """Candidate pass of `simulate_monte_carlo` (M:3385-3442, gate_order="mc": the per-level touch limit is applied
    BEFORE the gate, so a bar the limit drops never reaches the engine and a bar the gate drops still uses up the
    limit) and of `simulate_last_bars` (M:3566-3612, gate_order="sim": the gate runs FIRST, the touch limit counts only
    the bars it let through).  with_gates=True runs `evaluate_entry` per surviving bar on one fresh engine (SimGates);
    the Monte Carlo then takes level and side from the gate's extras (M:3425-3426), the deterministic sim keeps its
    own.  `skips` collects (i, reason, message, prox) of the bars the gate refused (the sim's GATE_SKIP audit rows).
    Returns list of dict(i, side, k, lvl, entry, stop, target, prox, tc).
    """
    prox, sp, tp = float(params.contact_prox), float(params.stop_padding), float(params.tp_padding)
    gates = SimGates(levels, params, lr) if with_gates else None
    cands = []
    touch_counts = {}
    prev_c = None
    for i, c in enumerate(bars_c):
        c = float(c)
        if prev_c is None:
            prev_c = c
            continue
        k, proximity_abs = nearest_level_first(levels, c)
        lvl = levels[k].price
        if proximity_abs > prox:
            prev_c = c
            continue
        side = 0 if c > prev_c else 1
        key = round(lvl, 4)
        if gate_order == "mc":
            tc = touch_counts.get(key, 0) + 1
            if tc > touch_limit:
                prev_c = c
                continue
            touch_counts[key] = tc
        if gates is not None:
            ok, reason, msg, g_side, g_k = gates.evaluate(c, prev_c)
            if not ok:
                if skips is not None:
                    skips.append((i, reason, msg, proximity_abs))
                prev_c = c
                continue
            if gate_order == "mc":                                    # M:3425-3426 (same level: same first-wins argmin)
                side, k, lvl = g_side, g_k, levels[g_k].price
        if gate_order != "mc":
            tc = touch_counts.get(key, 0) + 1
            if tc > touch_limit:
                prev_c = c
                continue
            touch_counts[key] = tc
        if side == 0:
            stop, target = lvl - sp, lvl + tp
        else:
            stop, target = lvl + sp, lvl - tp
        cands.append(dict(i=i, side=side, k=k, lvl=lvl, entry=c, stop=stop, target=target,
                          prox=proximity_abs, tc=tc))
        prev_c = c
    return cands


def walk_outcome(cand, highs, lows, sp, tp, normals, uniforms):
    """M:3449-3486 with the random draws supplied: normals = (level, entry, stop, target)
    already scaled by their sigmas; uniforms = iterator of U[0,1) used one per tied bar.
    Returns (R, outcome) with outcome 0 tp, 1 stop, 2 open.
    """
    side = cand["side"]
    lvl_j = cand["lvl"] + normals[0]
    entry = cand["entry"] + normals[1]
    if side == 0:
        stop = lvl_j - sp + normals[2]
        target = lvl_j + tp + normals[3]
    else:
        stop = lvl_j + sp + normals[2]
        target = lvl_j - tp + normals[3]
    risk = abs(entry - stop) or 1e-9
    reward = abs(target - entry)
    for j in range(cand["i"] + 1, len(highs)):
        hh, ll = float(highs[j]), float(lows[j])
        stop_hit = (ll <= stop) if side == 0 else (hh >= stop)
        target_hit = (hh >= target) if side == 0 else (ll <= target)
        if stop_hit and target_hit:
            up_span = max(0.0, hh - entry)
            down_span = max(0.0, entry - ll)
            p_target_first = up_span / (up_span + down_span + 1e-9)
            if next(uniforms) < p_target_first:
                return reward / risk, 0
            return -1.0, 1
        elif target_hit:
            return reward / risk, 0
        elif stop_hit:
            return -1.0, 1
    return 0.0, 2


def mc_summary(totals, worst_dds, win_counts, loss_counts, open_counts):
    """M:3513-3526 index rules, returned as numbers (the reference formats them)."""
    from statistics import mean, median, pstdev
    ts = sorted(totals)
    p05_idx = max(0, int(0.05 * len(ts)) - 1)
    return dict(
        mean=mean(totals), median=median(totals), stdev=pstdev(totals),
        best=max(totals), worst=min(totals),
        var05=ts[p05_idx], cvar05=mean(ts[:p05_idx + 1]),
        dd_mean=mean(worst_dds), dd_worst=min(worst_dds),
        wins=mean(win_counts), losses=mean(loss_counts), opens=mean(open_counts),
    )


def jitter_monte_carlo(bars, levels, params: Params, *, n=600, touch_limit=1, trials=500,
                       entry_slip_std=0.01, level_jitter_std=0.02, stop_slip_std=0.0,
                       target_slip_std=0.0, seed=None, draws=None, with_gates=False, lr=None):
    """`simulate_monte_carlo` (M:3353-3538) on bars = dict(o,h,l,c lists); with_gates as the reference's argument
    (its default is True and its only call site passes True, M:2662).

    Random source: Python `random.Random(seed)` exactly as the reference (M:3489), or
    `draws` = (normals[trials][K][4] unscaled N(0,1), uniforms[trials][K]) to replay
    host-drawn numbers (the device test mode).
    """
    import random
    c = list(bars["c"])[-n:]
    h = list(bars["h"])[-n:]
    l = list(bars["l"])[-n:]
    cands = find_candidates(c, levels, params, touch_limit=touch_limit, gate_order="mc", with_gates=with_gates, lr=lr)
    if not cands:
        return None
    sp, tp = float(params.stop_padding), float(params.tp_padding)
    rng = random.Random(seed)
    sig = (level_jitter_std, entry_slip_std, stop_slip_std, target_slip_std)
    totals, dds, wins_l, loss_l, open_l = [], [], [], [], []
    for t in range(trials):
        eq = peak = max_dd = 0.0
        wins = losses = opens = 0
        for ci, cand in enumerate(cands):
            if draws is None:
                normals = tuple(rng.gauss(0.0, s) for s in sig)

                def _u():
                    while True:
                        yield rng.random()
                uniforms = _u()
            else:
                normals = tuple(float(draws[0][t][ci][q]) * sig[q] for q in range(4))
                uniforms = iter([float(draws[1][t][ci])])
            r, outcome = walk_outcome(cand, h, l, sp, tp, normals, uniforms)
            eq += r
            peak = max(peak, eq)
            max_dd = min(max_dd, eq - peak)
            if outcome == 0:
                wins += 1
            elif outcome == 1:
                losses += 1
            else:
                opens += 1
        totals.append(eq)
        dds.append(max_dd)
        wins_l.append(wins)
        loss_l.append(losses)
        open_l.append(opens)
    out = mc_summary(totals, dds, wins_l, loss_l, open_l)
    out.update(candidates=len(cands), trials=trials, totals=totals, drawdowns=dds)
    return out


def simulate_last_bars(bars, levels, params: Params, *, n=1200, touch_limit=2, with_gates=False, lr=None, skips=None):
    """`simulate_last_bars` (M:3540-3699): stop tested before target; with_gates as the reference's argument (default
    False; the GUI button passes True, M:2655-2657).  `skips` collects the GATE_SKIP rows (M:3594-3597)."""
    c = list(bars["c"])[-n:]
    h = list(bars["h"])[-n:]
    l = list(bars["l"])[-n:]
    cands = find_candidates(c, levels, params, touch_limit=touch_limit, gate_order="sim", with_gates=with_gates, lr=lr,
                            skips=skips)
    trades = []
    for cand in cands:
        side, stop, target, entry = cand["side"], cand["stop"], cand["target"], cand["entry"]
        hit = None
        for j in range(cand["i"] + 1, len(c)):
            hh, ll = float(h[j]), float(l[j])
            if side == 0:
                if ll <= stop:
                    hit = ("stop", j); break
                if hh >= target:
                    hit = ("tp", j); break
            else:
                if hh >= stop:
                    hit = ("stop", j); break
                if ll <= target:
                    hit = ("tp", j); break
        if hit is None:
            outcome, exit_px, r = "open", float(c[-1]), 0.0
        else:
            outcome = hit[0]
            exit_px = target if outcome == "tp" else stop
            risk = abs(entry - stop) or 1e-9
            reward = abs(target - entry)
            r = (reward / risk) if outcome == "tp" else -1.0
        trades.append(dict(i=cand["i"], j=(hit[1] if hit else -1), side=side, lvl=round(cand["lvl"], 2),
                           entry=round(entry, 2), stop=round(stop, 2), target=round(target, 2),
                           exit=round(exit_px, 2), outcome=outcome, R=round(r, 2),
                           prox=round(cand["prox"], 3), touch=cand["tc"], R_raw=r))
    closed = [t for t in trades if t["outcome"] in ("tp", "stop")]

    def usd(t):
        return (t["exit"] - t["entry"]) * (1.0 if t["side"] == 0 else -1.0)
    summary = dict(
        trades=len(trades),
        wins=sum(1 for t in trades if t["outcome"] == "tp"),
        loss=sum(1 for t in trades if t["outcome"] == "stop"),
        open=sum(1 for t in trades if t["outcome"] == "open"),
        total_R=round(sum(t["R"] for t in trades), 2),
        avg_R=round(sum(t["R"] for t in closed) / max(1, len(closed)), 2),
        max_win_R=round(max((t["R"] for t in closed), default=0.0), 2),
        max_loss_R=round(min((t["R"] for t in closed), default=0.0), 2),
        max_profit_usd=round(max((usd(t) for t in closed), default=0.0), 2),
        max_loss_usd=round(min((usd(t) for t in closed), default=0.0), 2),
    )
    return trades, summary
