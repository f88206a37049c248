"""Differential test: oracle/headless.py vs the UNMODIFIED reference engine, step by step.
Runs only where /root/reference is mounted (the build container); skipped on the GPU box."""
import numpy as np
import pytest

from oracle import headless as H
from oracle import ref_harness as rh
from oracle.make_golden import settings_for
from qmmx_monolithic_monte_carlo_b200 import abi, engine as E, fixtures as F, montecarlo as MC

pytestmark = pytest.mark.skipif(not rh.reference_available(), reason="reference not mounted")


def _agg_from_codes(codes):
    a = abi.Agg()
    for c in codes:
        a.entry_hist[c] += 1
    return E.MCResult(a)


def _check(levels, params, closes, lr=None, pol=None, volume=None):
    lv = H.sort_levels(levels)
    P = H.Params(**params)
    models = [rh.make_sklearn_model(m.coef, m.intercept) for m in lr] if lr else None
    for j in range(closes.shape[1]):
        vj = None if volume is None else [float(x) for x in volume[:, j]]
        ref = rh.drive_reference_path(closes[:, j], levels, settings_for(params), lr_model=models,
                                      policy_w=pol, disable_policy_gate=P.disable_policy_gate, volume=vj,
                                      online_retrain=P.online_retrain, steps_per_day=P.steps_per_day)
        o = H.run_path(closes[:, j], lv, P, lr=lr, policy=pol, volume=vj)
        assert [H.ENTRY_CODES[c] for c in o.codes] == ref["codes"]
        assert len(o.trades) == len(ref["trades"])
        for a, b in zip(ref["trades"], o.trades):
            assert (a["entry_bar"], a["exit_bar"], a["reason"], a["level"], a["side"]) == \
                   (b.entry_bar, b.exit_bar, H.EXIT_CODES[b.reason], lv[b.level].key, ("long", "short")[b.side])
            assert a["pnl"] == b.pnl and a["entry"] == b.entry and a["exit"] == b.exit
        assert {lv[i].key: t for i, t in enumerate(o.touch_counts) if t} == ref["touch_counts"]
        if P.online_retrain:
            assert o.policy_weights == ref["policy_weights"]        # every float64 weight, bit for bit
        # the reason histogram (what the product's results sink keeps) == log_analyzer's fold of the policy_events rows:
        # action='skip' rows for the gate reasons, action='policy_skip' rows for 904 / soft veto (M:1658, M:1694; the veto
        # row carries the bare veto code).  ONLINE_POLICY rows are written by the GUI loop (M:3097), not by the engine.
        a = _agg_from_codes(o.codes)
        assert MC.skip_reason_counts(a) == ref["skip_rows"]
        pol_skips = {k: v for k, v in MC.skip_reason_counts(a, policy_skips=True).items()
                     if k not in ref["skip_rows"] and k != "ONLINE_POLICY"}
        assert {(k[5:] if k.startswith("VETO_") else k): v for k, v in pol_skips.items()} == ref["policy_skip_rows"]


@pytest.mark.parametrize("seed", [101, 102, 103])
def test_random_params_against_reference(seed):
    rng = np.random.default_rng(seed)
    params = dict(contact_prox=float(rng.choice([0.05, 0.1, 0.2])),
                  stop_padding=float(rng.choice([0.1, 0.35, 0.85])),
                  tp_padding=float(rng.choice([0.15, 0.25, 0.5])),
                  cooldown_s=int(rng.choice([0, 8, 60, 300, 900])),
                  q_min_prob=float(rng.choice([0.5, 0.6, 0.75])),
                  use_blend=bool(rng.integers(2)), enable_veto=bool(rng.integers(2)),
                  disable_policy_gate=bool(rng.integers(2)))
    lr = [H.LRModel(rng.normal(0, 0.5, 4).tolist(), float(rng.normal(0.5, 0.5)))]
    pol = [{k: rng.normal(0.3, 0.8, 7).tolist() for k in ("go_long", "go_short", "skip")}]
    closes = F.host_gbm_closes(6, 390, seed=seed, round_cents=bool(rng.integers(2)),
                               sigma=float(rng.choice([0.05, 0.15, 0.4])))
    _check(F.levels_30(), params, closes, lr=lr, pol=pol)


def test_volume_soft_veto_against_reference():
    """Row a7: per-bar volume in engine.recent_bars makes the reference's soft veto fire (M:1678-1705, M:1773-1811)."""
    rng = np.random.default_rng(7)
    closes = F.host_gbm_closes(4, 390, seed=77)
    vol = (1e5 * np.exp(0.5 * np.cumsum(rng.normal(0, 0.35, size=closes.shape), axis=0))).astype(np.float32)
    vol[rng.random(size=vol.shape) < 0.01] = 0.0
    params = dict(contact_prox=0.1, cooldown_s=60, enable_veto=True, veto_prox=0.08)
    _check(F.levels_30(), params, closes, volume=vol)
    lv = H.sort_levels(F.levels_30())
    codes = [c for j in range(4) for c in H.run_path(closes[:, j], lv, H.Params(**params),
                                                      volume=[float(x) for x in vol[:, j]]).codes]
    assert sum(1 for c in codes if H.ENTRY_CODES[c].startswith("VETO_CONTRA")) >= 3


def _breakout_paths(n, seed):
    """Quiet stretches (60-minute range under 18 bp) broken by jumps on a volume spike: every regime occurs."""
    rng = np.random.default_rng(seed)
    closes = np.empty((390, n), dtype=np.float32)
    vol = np.empty((390, n), dtype=np.float32)
    for j in range(n):
        c = 646.0 + rng.uniform(0, 5) + np.cumsum(rng.normal(0, 0.02, 390))
        v = rng.uniform(1e4, 2e4, 390)
        for t0 in sorted(rng.integers(60, 360, size=3)):
            c[t0:] += rng.choice([-1.0, 1.0]) * rng.uniform(1.3, 2.5)
            if rng.random() < 0.8:
                v[max(0, t0 - 2):t0 + 6] *= rng.uniform(3, 8)
        closes[:, j], vol[:, j] = c, v
    return closes, vol


def test_regime_guard_against_reference():
    """AccumulationBreakoutGuard fed through ingest_tick; code 904 blocks counter-trend entries."""
    closes, vol = _breakout_paths(12, 21)
    params = dict(contact_prox=0.1, cooldown_s=60, enable_veto=True)
    _check(F.levels_30(), params, closes, volume=vol)
    lv = H.sort_levels(F.levels_30())
    n904, regimes = 0, set()
    for j in range(closes.shape[1]):
        o = H.run_path(closes[:, j], lv, H.Params(**params), volume=[float(x) for x in vol[:, j]])
        n904 += sum(1 for c in o.codes if H.ENTRY_CODES[c] == "ACC_BREAKOUT_GATE")
        regimes |= set(H.regime_series(closes[:, j], vol[:, j]))
    assert n904 >= 3 and regimes == {0, 1, 2, 3}


def test_online_policy_retrain_against_reference():
    """the reference's OnlinePolicy learning from its own closed trades, driven for real (rows, labels,
    `_retrain_from_labeled_events`), against the oracle: per-tick codes, trades and every learnt weight identical.
    Three days per path, so that the learning (a bias step per trade, whatever the entry looked like -- the retrain
    reads feature keys the rows do not have) turns the gate against one side and ONLINE_POLICY skips appear."""
    closes = F.host_gbm_closes(4, 3 * 390, seed=404)
    pol = [{"go_long": [0.55, -0.3, 0.0, 0.0, 0.0, 0.05, -0.1], "go_short": [0.5, -0.2, 0.0, 0.0, 0.0, 0.05, 0.05],
            "skip": [-0.6, 0.2, 0.0, 0.0, 0.1, -0.1, 0.1]}]
    params = dict(contact_prox=0.1, cooldown_s=60, disable_policy_gate=False, online_retrain=True)
    _check(F.levels_30(), params, closes, pol=pol)
    lv = H.sort_levels(F.levels_30())
    n_pol, moved, changed = 0, 0.0, 0
    for j in range(closes.shape[1]):
        o = H.run_path(closes[:, j], lv, H.Params(**params), policy=pol)
        n_pol += sum(1 for c in o.codes if H.ENTRY_CODES[c] == "ONLINE_POLICY")
        moved += abs(o.policy_weights["go_long"][0] - 0.55) + abs(o.policy_weights["go_short"][0] - 0.5)
        assert o.policy_weights["skip"] == pol[0]["skip"]            # only the head of the action taken learns
        assert o.policy_weights["go_long"][1] != -0.3                # L2 decay and the proximity feature move it
        frozen = H.run_path(closes[:, j], lv, H.Params(**dict(params, online_retrain=False)), policy=pol)
        changed += frozen.codes != o.codes                           # the learning changes decisions
    assert n_pol > 0 and moved > 0.1 and changed >= 2


def test_touch_memory_is_inert_in_the_reference():
    """LevelTouchMemory never records a touch or an edge tap: register_touch_if_any gets dict levels and fails inside
    its try/except (M:1838-1851), so the accumulation-regime gates of M:1589-1621 cannot fire."""
    closes, vol = _breakout_paths(2, 22)
    for j in range(2):
        ref = rh.drive_reference_path(closes[:, j], F.levels_30(), settings_for(dict(contact_prox=0.1)),
                                      volume=[float(x) for x in vol[:, j]])
        assert ref["touchmem_edge_taps"] == 0
        assert all(r["count"] == 0 for r in ref["touchmem_touches"].values())
        assert not any(c.startswith("VETO_EDGE") or c.startswith("VETO_BOUNCE") or c.startswith("VETO_LEVEL") for c in ref["codes"])


def test_target_escalation_is_inert_in_the_reference():
    """at a target touch the live loop asks maybe_escalate_on_target (M:2988, M:3007) with the bars of
    PriceFeed.get_minute_bars (dicts, M:2972).  ExitStrategy unpacks every bar as a (price, volume, ts) tuple
    (M:986, M:1009), which raises on dicts inside its try/except, so it never escalates: the exit is always TARGET."""
    q, conn, eng = rh.make_engine(F.levels_30(), settings_for(dict(contact_prox=0.1)))
    lp = 648.76
    for side, entry, stop, target in (("long", lp - 0.02, lp - 0.35, lp + 0.25), ("short", lp + 0.02, lp + 0.35, lp - 0.25)):
        eng.open_trade(side, entry, stop, target, "sim")
        rising = [dict(t=60000 * j, o=target - 0.5 + 0.05 * j, h=target, l=target - 0.6, c=target - 0.5 + 0.05 * j, v=1e4 * (j + 1))
                  for j in range(10)]
        for bars in (None, [], rising, [{k: b[k] for k in "tohlc"} for b in rising],
                     [dict(price=b["c"], volume=b["v"], ts=b["t"]) for b in rising]):
            escalated, meta = eng.maybe_escalate_on_target(target, bars)
            assert escalated is False
        eng.close_trade(eng.state.open_trade_id, target, "TARGET")


def test_lr_contract_matches_sklearn_predict_proba():
    rng = np.random.default_rng(5)
    for _ in range(50):
        coef = rng.normal(0, 1, 4)
        b = float(rng.normal())
        m = rh.make_sklearn_model(coef, b)
        x = [1.0, abs(648.76 - (648.76 - 0.35)), float(rng.integers(1, 4)), float(rng.integers(2))]
        ours = H.lr_prob(H.LRModel(coef.tolist(), b), *x)
        assert ours == pytest.approx(float(m.predict_proba([x])[0][1]), abs=1e-15)  # scipy expit vs 1/(1+exp(-z)): last-bit differences only


def _random_levels(rng):
    """4..22 levels around the start price: duplicate prices (first-wins ties across colours), pairs closer than the
    contact window (overlapping windows: latch release cases), a far outlier; both types, shuffled colours/indices."""
    n = int(rng.integers(4, 23))
    base = F.GBM_S0 + np.round(rng.normal(0, 2.5, n), 2)
    if n >= 6:
        base[1] = base[0]                                  # exact duplicate
        base[3] = base[2] + float(rng.choice([0.03, 0.09, 0.11]))     # inside / at the edge of a 0.1 window
        base[5] = base[4] - 0.2                            # windows touch at prox 0.1, overlap at 0.2
    base[-1] = float(rng.choice([500.0, 700.0]))
    colours = ["black", "blue", "orange", "teal"]
    seen, out = {}, []
    for p in base:
        c, t = str(rng.choice(colours)), str(rng.choice(["solid", "dashed"]))
        i = seen.get((c, t), 0)
        seen[(c, t)] = i + 1
        out.append({"color": c, "type": t, "index": i, "price": float(p)})
    return out


def _shaped_closes(rng, n, steps):
    """GBM with flat stretches, exact returns to an earlier price (zero net move) and cent rounding on half the paths."""
    sigma = float(rng.choice([0.03, 0.1, 0.15, 0.4]))
    c = F.host_gbm_closes(n, steps, seed=int(rng.integers(1 << 30)), sigma=sigma, round_cents=bool(rng.integers(2))).copy()
    for j in range(n):
        for _ in range(int(rng.integers(0, 4))):
            a = int(rng.integers(0, steps - 2)); b = min(steps, a + int(rng.integers(2, 30)))
            c[a:b, j] = c[a, j]                            # flat stretch (DIR from last_direction, M:1536)
        if rng.random() < 0.3:
            c[: int(rng.integers(2, 12)), j] = c[0, j]     # flat start: DIR_UNKNOWN prefix
    return c


@pytest.mark.parametrize("seed", range(2000, 2040))
def test_wide_seeded_differential_against_reference(seed):
    """40 seeds x 8 paths = 320 paths through the UNMODIFIED reference engine and the oracle, step by step: random level
    sets (duplicates, overlapping windows, outliers), every flag combination, flat stretches, 1-3 day logistic and
    policy tables, short and long cooldowns.  The reference costs ~15 ms per path; the C oracle and the GPU are pinned
    to headless.py on tens of thousands of paths elsewhere, this test is the independent anchor."""
    rng = np.random.default_rng(seed)
    days = int(rng.choice([1, 1, 2, 3]))
    steps = int(rng.choice([60, 150, 390])) if days == 1 else 130 * days
    params = dict(contact_prox=float(rng.choice([0.05, 0.1, 0.2])), stop_padding=float(rng.choice([0.1, 0.35, 0.85])),
                  tp_padding=float(rng.choice([0.15, 0.25, 0.5])), cooldown_s=int(rng.choice([0, 8, 60, 120, 900])),
                  q_min_prob=float(rng.choice([0.3, 0.6, 0.75, 0.9])), use_blend=bool(rng.integers(2)),
                  enable_veto=bool(rng.integers(2)), disable_policy_gate=bool(rng.integers(2)),
                  disable_ml_gate=bool(rng.random() < 0.25))
    if params["use_blend"]:
        params.update(w_rules=float(rng.choice([0.7, 0.5, 1.0])), w_ml=float(rng.choice([0.3, 0.5, 0.0])))
    if days > 1:
        params["steps_per_day"] = 130
    lr = [H.LRModel(rng.normal(0, 0.6, 4).tolist(), float(rng.normal(0.5, 0.6))) for _ in range(days)] if rng.random() < 0.8 else None
    pol = [{k: rng.normal(0.3, 0.8, 7).tolist() for k in ("go_long", "go_short", "skip")} for _ in range(days)]
    _check(_random_levels(rng), params, _shaped_closes(rng, 8, steps), lr=lr, pol=pol)
