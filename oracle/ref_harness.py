"""TEST INFRASTRUCTURE ONLY — drives the *unmodified* reference engine headlessly.

This file exists to pin `oracle/headless.py` against the reference's own code and to
generate the golden vectors under `tests/golden/`. It only works in the build
container, where `/root/reference` is mounted; nothing on the GPU box imports it
(tests that need it skip when the reference is absent).

How: the monolith imports tkinter/matplotlib at module top
(`qmmx_monolithic.py:5-13`), neither of which is installed.  We register dummy modules
for those names in `sys.modules`, import `qmmx_monolithic`, build a
`MonolithicEngine` on an in-memory SQLite DB and drive `evaluate_entry`
(`qmmx_monolithic.py:1492-1771`) with a transcription of the live loop's tick order
and exit block (`qmmx_monolithic.py:2927-3016`, `3044-3115`).
"""
from __future__ import annotations

import json
import os
import sqlite3
import sys
import types
from types import SimpleNamespace

REFERENCE_DIR = os.environ.get("QMMX_REFERENCE_DIR", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "qmmx_monolithic.py"))


_q = None
_captured_messages: list = []


def _stub_gui_modules() -> None:
    class _Any:
        def __init__(self, *a, **k):
            pass

        def __getattr__(self, name):
            return _Any()

        def __call__(self, *a, **k):
            return _Any()

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)

        def _module_getattr(attr):
            if attr.startswith("__"):      # __file__, __path__ ...: tools that walk sys.modules must see a plain module
                raise AttributeError(attr)
            return _Any
        m.__getattr__ = _module_getattr  # type: ignore[attr-defined]
        sys.modules[name] = m
        return m

    class Tk:  # `class QMMXApp(tk.Tk)` executes at import time
        def __init__(self, *a, **k):
            pass

    def showinfo(title, msg=None, *a, **k):
        _captured_messages.append(("info", title, msg))

    def showerror(title, msg=None, *a, **k):
        _captured_messages.append(("error", title, msg))

    tk = mod("tkinter", Tk=Tk, END="end", Text=_Any, StringVar=_Any, BooleanVar=_Any)
    ttk = mod("tkinter.ttk")
    mb = mod("tkinter.messagebox", showinfo=showinfo, showerror=showerror)
    st = mod("tkinter.scrolledtext")
    tk.ttk, tk.messagebox, tk.scrolledtext = ttk, mb, st
    mpl = mod("matplotlib")
    be = mod("matplotlib.backends")
    bt = mod("matplotlib.backends.backend_tkagg", FigureCanvasTkAgg=_Any)
    pp = mod("matplotlib.pyplot")
    mpl.backends, be.backend_tkagg, mpl.pyplot = be, bt, pp


def load_reference():
    """Import the reference monolith with GUI modules stubbed; returns the module."""
    global _q
    if _q is not None:
        return _q
    if not reference_available():
        raise RuntimeError(f"reference not present under {REFERENCE_DIR}")
    for name in ("tkinter", "matplotlib"):
        if name not in sys.modules:
            _stub_gui_modules()
            break
    if REFERENCE_DIR not in sys.path:
        sys.path.insert(0, REFERENCE_DIR)
    import qmmx_monolithic as q  # noqa: E402

    _q = q
    return q


def make_engine(levels, settings=None):
    """Fresh reference `MonolithicEngine` on sqlite :memory: with `levels` and `settings`.

    levels: list of {"color","type","index","price"} dicts (any order; the engine reads
    them back through `load_levels`, i.e. ORDER BY color, level_type, level_index).
    settings: {key: str(value)} written with `settings_set` before the engine is built.
    """
    q = load_reference()
    conn = sqlite3.connect(":memory:")
    q.db_init(conn)
    q.replace_levels(conn, levels)
    for k, v in (settings or {}).items():
        q.settings_set(conn, k, str(v))
    eng = q.MonolithicEngine(conn)
    eng.model = None
    return q, conn, eng


class _CoefModel:
    """Stands in for a fitted sklearn LogisticRegression: predict_proba = sigma(b + w.x).

    Used only when sklearn itself is not wanted; `make_sklearn_model` builds a real one.
    """

    def __init__(self, coef, intercept):
        self.coef = [float(c) for c in coef]
        self.intercept = float(intercept)

    def predict_proba(self, X):
        import math

        out = []
        for x in X:
            z = self.intercept
            for w, xi in zip(self.coef, x):
                z = z + w * float(xi)
            p = 1.0 / (1.0 + math.exp(-z))
            out.append([1.0 - p, p])
        return out


def make_sklearn_model(coef, intercept):
    """A real sklearn LogisticRegression with hand-set coefficients (no fit)."""
    import numpy as np
    from sklearn.linear_model import LogisticRegression

    m = LogisticRegression()
    m.classes_ = np.array([0, 1])
    m.coef_ = np.asarray(coef, dtype=np.float64).reshape(1, -1)
    m.intercept_ = np.asarray([intercept], dtype=np.float64)
    m.n_features_in_ = m.coef_.shape[1]
    return m


def drive_reference_path(closes, levels, settings=None, *, lr_model=None, policy_w=None,
                         disable_policy_gate=True, steps_per_day=390, volume=None, online_retrain=False):
    """Run the real `evaluate_entry` over one close series with the live loop's order.

    Transcribes `_engine_loop` (`qmmx_monolithic.py:2927-3016`, `3044-3115`) with the
    frozen choices of SURVEY.md Appendix A: tick i = close i, now_ms = 60000*i,
    evaluate_entry called once, escalation off, policy minutes = i mod steps_per_day.

    lr_model: None, or an object with predict_proba (set as engine.model).
    policy_w: None or {"go_long": [7], "go_short": [7], "skip": [7]} or a per-day list of
              such dicts; the reference `OnlinePolicy` scores them.
    volume:   None or one per-bar volume per close; at tick i `engine.recent_bars` holds the last
              min(i+1, 240) bars `{t,o,h,l,c,v}` (Appendix A), which is what `_calc_volume_slope` reads.
              When the soft veto fires the reference raises NameError (`VETO` is undefined, M:1705,
              Appendix C) AFTER writing its policy_events row; the step code is taken from that row.
    online_retrain: the live loop's learning path run for real: an entry that passes the policy gate files the
              policy_events row of M:3139-3150 (features_json {"x": x}, trade id), the engine's own `close_trade` labels it
              (M:1936-1941) and the unbound `QMMXApp._retrain_from_labeled_events` (M:3753-3803) is called before every
              tick on the reference's own `OnlinePolicy`; `policy_w` is then ONE weight table, the initial weights.
    Returns dict(codes=[...str per step], trades=[...], touch_counts=...).
    """
    q, conn, eng = make_engine(levels, settings)
    eng.model = lr_model if not isinstance(lr_model, list) else None
    policy = q.OnlinePolicy()
    if online_retrain:
        policy.build_features(proximity_abs=0.0, volume_trend=0.0, approach="from_above", confluence=False,
                              minutes_since_open=0)               # fixes the dimension and zero-fills the heads (M:321-327)
        if policy_w is not None:
            w0 = policy_w[0] if isinstance(policy_w, list) else policy_w
            policy.w_entry = {k: list(map(float, v)) for k, v in w0.items()}
        app = types.SimpleNamespace(conn=conn, policy=policy, _log_ui=lambda *a, **k: None, _save_policy=lambda: None)
    cool_s = eng.Q_SIGNAL_COOLDOWN_S
    codes, trades = [], []
    pos = None
    for i, price in enumerate(closes):
        price = float(price)
        day = i // steps_per_day
        if isinstance(lr_model, list):
            eng.model = lr_model[min(day, len(lr_model) - 1)]
        prev_price = eng.state.last_price
        now_ms = 60000 * i
        eng.state.last_ts_ms = now_ms  # always fresh (Appendix A [CHOICE])
        if online_retrain:
            q.QMMXApp._retrain_from_labeled_events(app, max_rows=1000)      # Tk timer body, M:3824-3828
        if volume is not None:
            # the live loop feeds every tick to ingest_tick before evaluate_entry (M:2932): it rolls the one-tick minute
            # bar i-1 into acc_guard / touchmem (M:1857-1883, M:1813-1855)
            eng.ingest_tick(now_ms, price, float(volume[i]))
            lo = max(0, i - 239)
            eng.recent_bars = [dict(t=60000 * j, o=float(closes[j]), h=float(closes[j]), l=float(closes[j]),
                                    c=float(closes[j]), v=float(volume[j])) for j in range(lo, i + 1)]
        try:
            ok, code, msg, extras = eng.evaluate_entry(price, prev_price, now_ms, True)
        except NameError:
            row = conn.execute("SELECT features_json FROM policy_events ORDER BY id DESC LIMIT 1").fetchone()
            ok, code, msg, extras = False, "VETO_" + json.loads(row[0])["reason"], "veto", {}
        eng.state.last_price = price
        if prev_price is not None and price != prev_price:
            eng.state.last_direction = "up" if price > prev_price else "down"
        step_code = code if not isinstance(code, int) else ("ACC_BREAKOUT_GATE" if code == 904 else f"INT_{code}")
        if step_code == "CONF_LOW":
            # sub-reason lives in the policy_events row the gate just wrote (M:1629, 1733, 1746)
            row = conn.execute("SELECT features_json FROM policy_events ORDER BY id DESC LIMIT 1").fetchone()
            sub = json.loads(row[0]).get("reason", "CONF_LOW")
            step_code = sub if sub in ("ML_CONF_LOW", "COMBINED_LOW") else "CONF_LOW"
        if eng.state.open_trade_id is not None:
            side, stop, target = pos["side"], pos["stop"], pos["target"]
            reason = None
            if side == "long":
                if price <= stop:
                    reason = "STOP"
                elif price >= target:
                    reason = "TARGET"
            else:
                if price >= stop:
                    reason = "STOP"
                elif price <= target:
                    reason = "TARGET"
            if reason is not None:
                tid = eng.state.open_trade_id
                eng.close_trade(tid, price, reason)
                row = conn.execute("SELECT pnl FROM trades WHERE id=?", (tid,)).fetchone()
                eng.state.set_cooldown(now_ms, cool_s)
                trades.append(dict(entry_bar=pos["entry_bar"], exit_bar=i, side=side,
                                   level=pos["level"], reason=reason, entry=pos["entry"],
                                   exit=price, pnl=float(row[0])))
                pos = None
        elif ok:
            pass_gate = True
            if not disable_policy_gate:
                w = policy_w
                if isinstance(w, list):
                    w = w[min(day, len(w) - 1)]
                lvl_price = float(extras["level_price"])
                side = extras["side"]
                approach = "from_below" if side == "long" else "from_above"
                cluster = [lv for lv in eng.levels_cache
                           if abs(float(lv["price"]) - lvl_price) <= 0.6]
                x = policy.build_features(proximity_abs=abs(price - lvl_price), volume_trend=0.0,
                                          approach=approach, confluence=(len(cluster) > 1),
                                          minutes_since_open=i % steps_per_day)
                if w is not None and not online_retrain:
                    policy.w_entry = {k: list(v) for k, v in w.items()}
                scores = policy.score_entry(x)
                chosen = "go_long" if side == "long" else "go_short"
                pass_gate = (scores.get(chosen, 0.5) >= 0.60) and (scores.get("skip", 0.0) < 0.55)
            if not pass_gate:
                step_code = "ONLINE_POLICY"
            else:
                tid = eng.open_trade(extras["side"], price, extras["stop"], extras["target"], "sim")
                if online_retrain and not disable_policy_gate:
                    conn.execute("INSERT INTO policy_events (ts, phase, action, features_json, trade_id, notes) "
                                 "VALUES (?,?,?,?,?,?)",
                                 (q.utcnow(), "entry", "go_long" if extras["side"] == "long" else "go_short",
                                  json.dumps({"x": x}), tid, "sim"))                       # M:3139-3150
                    conn.commit()
                pos = dict(side=extras["side"], stop=extras["stop"], target=extras["target"],
                           entry=price, entry_bar=i, level=tuple(extras["level"]))
        codes.append(step_code)
    if pos is not None:
        last = float(closes[-1])
        mtm = (last - pos["entry"]) if pos["side"] == "long" else (pos["entry"] - last)
        trades.append(dict(entry_bar=pos["entry_bar"], exit_bar=-1, side=pos["side"],
                           level=pos["level"], reason="OPEN", entry=pos["entry"], exit=last,
                           pnl=0.0, mtm=mtm))
    # the fold of the reference's log_analyzer.py:14-31 over the rows this run wrote
    skip_rows, policy_skip_rows = {}, {}
    for action, fold in (("skip", skip_rows), ("policy_skip", policy_skip_rows)):
        for (fj,) in conn.execute("SELECT features_json FROM policy_events WHERE phase='entry' AND action=?", (action,)):
            r = json.loads(fj).get("reason", "UNKNOWN_REASON")
            fold[r] = fold.get(r, 0) + 1
    return dict(codes=codes, trades=trades, skip_rows=skip_rows, policy_skip_rows=policy_skip_rows,
                policy_weights={k: list(v) for k, v in policy.w_entry.items()},
                touch_counts=dict(eng.state.level_touch_counts),
                touchmem_touches=dict(eng.touchmem.touch), touchmem_edge_taps=len(eng.touchmem.edge_taps),
                regime=eng.acc_guard.regime)


def run_reference_monte_carlo(bars, levels, settings=None, lr_model=None, **kw):
    """Unbound `QMMXApp.simulate_monte_carlo` (`qmmx_monolithic.py:3353-3538`); returns text.  A FRESH engine per run
    (with_gates=True leaves touch counts and latches dirty); lr_model: None or an object with predict_proba."""
    q, conn, eng = make_engine(levels, settings)
    eng.model = lr_model
    eng.recent_bars = [dict(b) for b in bars]
    fake = SimpleNamespace(engine=eng, conn=conn, _log_ui=lambda *a, **k: None)
    _captured_messages.clear()
    q.QMMXApp.simulate_monte_carlo(fake, **kw)
    return _captured_messages[-1] if _captured_messages else None


def run_reference_last_bars(bars, levels, settings=None, lr_model=None, **kw):
    """Unbound `QMMXApp.simulate_last_bars` (`qmmx_monolithic.py:3540-3699`).

    Returns (messagebox tuple, list of audit rows (phase, code, message)).
    """
    q, conn, eng = make_engine(levels, settings)
    eng.model = lr_model
    eng.recent_bars = [dict(b) for b in bars]
    fake = SimpleNamespace(engine=eng, conn=conn, _log_ui=lambda *a, **k: None)
    _captured_messages.clear()
    q.QMMXApp.simulate_last_bars(fake, **kw)
    rows = conn.execute("SELECT phase, code, message FROM audit_log WHERE phase='SIM' ORDER BY id").fetchall()
    return (_captured_messages[-1] if _captured_messages else None), rows
