"""The oracle (oracle/headless.py) against the reference-generated golden vectors and the
known answers recovered from the reference's qmmx.db-wal."""
import pytest

from oracle import headless as H
from tests.golden_util import (format_mc_message, levels_by_name, load_engine_cases,
                               load_jitter_cases, oracle_inputs, jitter_case_setup)

ENGINE_CASES = load_engine_cases()
JITTER_CASES = load_jitter_cases()


@pytest.mark.parametrize("case", ENGINE_CASES, ids=[c["name"] for c in ENGINE_CASES])
def test_headless_matches_reference_golden(case):
    lv, P, lr, pol = oracle_inputs(case)
    for j in range(case["paths"]):
        r = H.run_path(case["close"][:, j], lv, P, lr=lr, policy=pol)
        assert r.codes == case["codes"][:, j].tolist()
        gt = case["trades"][j]
        assert len(gt) == len(r.trades)
        for g, t in zip(gt, r.trades):
            assert (g["entry_bar"], g["exit_bar"], g["side"], g["level"], g["reason"]) == \
                   (t.entry_bar, t.exit_bar, t.side, t.level, t.reason)
            assert g["entry"] == t.entry and g["exit"] == t.exit and g["pnl"] == t.pnl
            assert g["mtm"] == t.mtm


# ---- Appendix B.3: compute_confidence to the last float64 bit (prox 0.1, touch 1, solid) ----
WAL_CONF = [
    (0.0, True, 1, 1.0),
    (0.009999999999990905, True, 1, 1.0),
    (0.029999999999972715, True, 1, 0.9100000000002728),
    (0.029999999999972715, False, 1, 0.8500000000002729),
    (0.03999999999996362, True, 1, 0.8100000000003638),
    (0.049999999999954525, True, 1, 0.7100000000004547),
    (0.049999999999954525, False, 1, 0.6500000000004548),
    (0.05999999999994543, True, 1, 0.6100000000005458),
    (0.06999999999993634, True, 1, 0.5100000000006367),
    (0.07000000000005002, True, 1, 0.5099999999994999),
    (0.0764000000000351, True, 1, 0.445999999999649),
    (0.08000000000004093, True, 1, 0.4099999999995908),
    (0.09000000000003183, True, 1, 0.30999999999968175),
]


@pytest.mark.parametrize("dist,solid,tc,conf", WAL_CONF)
def test_confidence_wal_known_answers(dist, solid, tc, conf):
    assert H.compute_confidence(dist, solid, tc, 0.1) == conf


def test_confidence_touch_branches():
    # survey probe on the live engine: touch 2 / 3 at dist 0.01 solid
    d = abs(648.76 - 648.77)
    assert H.compute_confidence(d, True, 2, 0.1) == pytest.approx(0.930000000000091, abs=1e-15)
    assert H.compute_confidence(d, True, 3, 0.1) == pytest.approx(0.8500000000000909, abs=1e-15)
    # WAL "Confidence 0.05 < min 0.60. prox=0.090": touch >= 3 branch
    assert f"{H.compute_confidence(0.09, True, 3, 0.1):.2f}" == "0.05"
    # fp64 residue: |648.76 - 648.86| is just over 0.1 => TOO_FAR
    assert abs(648.76 - 648.86) == 0.10000000000002274 and abs(648.76 - 648.86) > 0.1


def test_wal_sim_trades_known_answers():
    # Appendix B.4: TP long @648.74 -> 649.01 lvl 648.76 R=+0.82 ; @649.21 -> 649.45 lvl 649.20 R=+0.67
    for entry, lvl, r_txt, usd_txt in [(648.74, 648.76, "+0.82", "+0.27"), (649.21, 649.20, "+0.67", "+0.24")]:
        stop, target = lvl - 0.35, lvl + 0.25
        risk = abs(entry - stop) or 1e-9
        r = round(abs(target - entry) / risk, 2)
        assert f"{r:+.2f}" == r_txt
        assert f"{round(target, 2) - round(entry, 2):+.2f}" == usd_txt


def test_duplicate_price_first_wins():
    lv = H.sort_levels(levels_by_name("fix30"))
    k, dist = H.nearest_level_first(lv, 646.6)
    assert lv[k].key == ("black", "solid", 1) and dist == 0.0


def test_shipped_online_policy_weights_probe():
    # Appendix B.6 probe: x=[1,.02,0,0,1,1,100/390] -> go_long 0.37472116060423755
    w = [-0.3408250760675784, -0.009429527396658328, 0.0, -0.1698232270674892, -0.1710018490000892, 0.0, 0.0]
    x = [1.0, 0.02, 0.0, 0.0, 1.0, 1.0, 100 / 390.0]
    assert H.sigmoid_policy(sum(a * b for a, b in zip(w, x))) == 0.37472116060423755
    assert H.sigmoid_policy(-51) == 0.0 and H.sigmoid_policy(51) == 1.0


@pytest.mark.parametrize("case", JITTER_CASES, ids=[c["name"] for c in JITTER_CASES])
def test_jitter_mc_matches_reference_text(case):
    """with_gates False and True (M:3414-3426: touch limit first, then the engine gate on one engine whose latches and
    touch counts carry over): same candidates => same Mersenne-Twister draws => the reference's text, character for
    character."""
    lv = H.sort_levels(levels_by_name(case["levels"]))
    wg, P, lr, _ = jitter_case_setup(case)
    s = H.jitter_monte_carlo(case["bars"], lv, P, n=case["n"], touch_limit=case["touch_limit"],
                             trials=case["trials"], seed=case["seed"], with_gates=wg, lr=lr)
    assert format_mc_message(s) == case["mc_message"]


@pytest.mark.parametrize("case", JITTER_CASES, ids=[c["name"] for c in JITTER_CASES])
def test_simulate_last_bars_matches_reference_rows(case):
    lv = H.sort_levels(levels_by_name(case["levels"]))
    wg, P, lr, _ = jitter_case_setup(case)
    skips = []
    trades, s = H.simulate_last_bars(case["bars"], lv, P, n=case["n"], touch_limit=case["touch_limit"], with_gates=wg,
                                     lr=lr, skips=skips)
    rows = case["sim_rows"]
    # the GATE_SKIP breadcrumbs (M:3594-3597): code, the gate's own message and the proximity, bar by bar
    gate_rows = [(r[1], r[2]) for r in rows if r[1].startswith("GATE_SKIP")]
    mine = [("GATE_SKIP:" + ("CONF_LOW" if r in ("ML_CONF_LOW", "COMBINED_LOW") else r), f"{msg} prox={d:.03f}")
            for _, r, msg, d in skips]
    assert mine == gate_rows and (len(gate_rows) > 0) == wg
    summary = [r for r in rows if r[1] == "SUMMARY"][0][2]
    expect = (f"N={case['n']} | trades={s['trades']} | wins={s['wins']} loss={s['loss']} open={s['open']} | "
              f"total_R={s['total_R']} avg_R(closed)={s['avg_R']} | "
              f"max_win_R={s['max_win_R']} max_loss_R={s['max_loss_R']} | "
              f"max_profit=${s['max_profit_usd']} max_loss=${s['max_loss_usd']} | "
              f"prox={P.contact_prox} stop={P.stop_padding} tp={P.tp_padding}")
    assert summary == expect
    trows = [r[2] for r in rows if r[1] == "TRADE"]
    assert len(trows) == len(trades)
    for txt, t in zip(trows, trades):
        usd = (t["exit"] - t["entry"]) * (1.0 if t["side"] == 0 else -1.0)
        side = "long" if t["side"] == 0 else "short"
        mine = (f"{t['outcome'].upper():5s} | {side:5s} @ {t['entry']:.2f} → {t['exit']:.2f} "
                f"| lvl {t['lvl']:.2f} | R={t['R']:+.2f} | ${usd:+.2f} "
                f"| prox={t['prox']:.03f} touch#{t['touch']}")
        assert txt == mine
