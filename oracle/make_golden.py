"""TEST INFRASTRUCTURE ONLY — generates tests/golden/* by running the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python -m oracle.make_golden
Outputs (committed):
  tests/golden/engine_cases.json + engine_cases.npz   per-step reason codes and trades from
      the reference `MonolithicEngine.evaluate_entry` driven with the live-loop order
      (`oracle/ref_harness.drive_reference_path`), for seeded fp32 close paths.
  tests/golden/jitter_mc.json     `QMMXApp.simulate_monte_carlo` / `simulate_last_bars`
      outputs (messagebox text, SIM audit rows) for seeded OHLC bars.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import headless as H  # noqa: E402
from oracle import ref_harness as rh  # noqa: E402
from qmmx_monolithic_monte_carlo_b200 import fixtures as F  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

LR_A = dict(coef=[0.3, -0.8, -0.45, 0.6], intercept=0.4)
LR_B = dict(coef=[-0.2, 1.1, 0.15, -0.5], intercept=0.1)
POL_A = {"go_long": [0.9, -2.0, 0.0, 0.1, 0.2, 0.3, -0.8],
         "go_short": [0.5, 1.0, 0.0, 0.2, -0.1, 0.1, 0.4],
         "skip": [-0.2, 0.5, 0.0, 0.0, 0.1, 0.2, 0.3]}
POL_B = {"go_long": [0.45, 0.5, 0.0, 0.0, 0.1, -0.1, 0.2],
         "go_short": [1.2, -3.0, 0.0, 0.1, 0.0, 0.2, -0.5],
         "skip": [0.1, 0.9, 0.0, 0.1, 0.0, -0.3, 0.0]}

# name, levels, params (headless.Params kwargs), lr per-day list, policy per-day list, gen kwargs
CASES = [
    dict(name="c1_blue_solid", levels="blue4", params=dict(contact_prox=0.1, cooldown_s=120),
         paths=32, steps=390, seed=11),
    dict(name="fixture30_default_cooldown", levels="fix30", params=dict(contact_prox=0.1, cooldown_s=8, enable_veto=True),
         paths=32, steps=390, seed=12),
    dict(name="fixture30_cents", levels="fix30", params=dict(contact_prox=0.1, cooldown_s=120),
         paths=32, steps=390, seed=13, round_cents=True),
    dict(name="fixture30_lr_and", levels="fix30", params=dict(contact_prox=0.1, cooldown_s=120),
         lr=[LR_A], paths=32, steps=390, seed=14),
    dict(name="fixture30_lr_blend", levels="fix30",
         params=dict(contact_prox=0.1, cooldown_s=300, use_blend=True, w_rules=0.6, w_ml=0.5),
         lr=[LR_A], paths=32, steps=390, seed=15, round_cents=True),
    dict(name="fixture30_policy", levels="fix30",
         params=dict(contact_prox=0.1, cooldown_s=60, disable_policy_gate=False),
         lr=[LR_A], policy=[POL_A], paths=32, steps=390, seed=16),
    dict(name="fixture30_multiday_swap", levels="fix30",
         params=dict(contact_prox=0.1, cooldown_s=900, disable_policy_gate=False),
         lr=[LR_A, LR_B, LR_A], policy=[POL_A, POL_B, POL_A], paths=16, steps=3 * 390, seed=17),
    dict(name="fixture30_tight_prox_wide_pads", levels="fix30",
         params=dict(contact_prox=0.05, stop_padding=0.85, tp_padding=0.5, cooldown_s=0, q_min_prob=0.55),
         paths=32, steps=390, seed=18, sigma=0.30),
    dict(name="fixture30_flat_ticks", levels="fix30", params=dict(contact_prox=0.1, cooldown_s=120),
         paths=32, steps=390, seed=19, round_cents=True, sigma=0.02),
]

SETTING_KEYS = dict(contact_prox="CONTACT_PROX", stop_padding="STOP_PADDING", tp_padding="TP_PADDING",
                    q_min_prob="Q_MIN_PROB", cooldown_s="Q_SIGNAL_COOLDOWN", w_rules="W_RULES", w_ml="W_ML",
                    veto_vol_strong="VETO_VOL_STRONG", veto_prox="VETO_PROX")


def settings_for(params: dict) -> dict:
    s = {"ENABLE_VETO": "1" if params.get("enable_veto") else "0"}
    for k, v in params.items():
        if k in SETTING_KEYS:
            s[SETTING_KEYS[k]] = repr(v) if isinstance(v, float) else str(v)
    if params.get("use_blend"):
        s["USE_BLEND"] = "1"
    if params.get("disable_ml_gate"):
        s["DISABLE_ML_GATE"] = "1"
    return s


def levels_by_name(name):
    return F.levels_30() if name == "fix30" else F.levels_blue_solid_4()


def gen_engine_cases():
    meta, arrays = [], {}
    for case in CASES:
        levels = levels_by_name(case["levels"])
        lv_sorted = H.sort_levels(levels)
        key_to_idx = {lv.key: i for i, lv in enumerate(lv_sorted)}
        closes = F.host_gbm_closes(case["paths"], case["steps"], seed=case["seed"],
                                   round_cents=case.get("round_cents", False),
                                   sigma=case.get("sigma", F.GBM_SIGMA_ANNUAL))
        lr = case.get("lr")
        pol = case.get("policy")
        models = [rh.make_sklearn_model(m["coef"], m["intercept"]) for m in lr] if lr else None
        codes = np.zeros((case["steps"], case["paths"]), dtype=np.uint8)
        trades = []
        for j in range(case["paths"]):
            r = rh.drive_reference_path(closes[:, j], levels, settings_for(case["params"]),
                                        lr_model=models, policy_w=pol,
                                        disable_policy_gate=case["params"].get("disable_policy_gate", True))
            codes[:, j] = [H.E[c] for c in r["codes"]]
            trades.append([dict(entry_bar=t["entry_bar"], exit_bar=t["exit_bar"],
                                side=0 if t["side"] == "long" else 1, level=key_to_idx[t["level"]],
                                reason=H.X[t["reason"]], entry=t["entry"], exit=t["exit"], pnl=t["pnl"],
                                mtm=t.get("mtm", 0.0)) for t in r["trades"]])
        arrays[case["name"] + "__close"] = closes
        arrays[case["name"] + "__codes"] = codes
        meta.append(dict(name=case["name"], levels=case["levels"], params=case["params"], lr=lr,
                         policy=pol, paths=case["paths"], steps=case["steps"], trades=trades))
        print(case["name"], "trades", sum(len(t) for t in trades),
              "codes", np.bincount(codes.ravel(), minlength=H.N_ENTRY_CODES).tolist())
    np.savez_compressed(os.path.join(GOLD, "engine_cases.npz"), **arrays)
    with open(os.path.join(GOLD, "engine_cases.json"), "w") as f:
        json.dump(dict(source="reference MonolithicEngine.evaluate_entry via oracle/ref_harness.py",
                       entry_codes=H.ENTRY_CODES, exit_codes=H.EXIT_CODES, cases=meta), f)


def synth_ohlc(n, seed, sigma=0.15, sub=4):
    """OHLC bars from `sub` GBM sub-steps per bar, rounded to cents like Polygon bars."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    vol = sigma * np.sqrt(F.GBM_DT / sub)
    z = rng.standard_normal(n * sub)
    px = F.GBM_S0 * np.exp(np.cumsum(vol * z))
    px = np.round(px, 2).reshape(n, sub)
    o = np.concatenate([[F.GBM_S0], px[:-1, -1]])
    h = np.maximum(px.max(axis=1), o)
    l = np.minimum(px.min(axis=1), o)
    c = px[:, -1]
    return dict(o=o.tolist(), h=h.tolist(), l=l.tolist(), c=c.tolist())


def gen_jitter_cases():
    """with_gates=False cases (the detector alone) and with_gates=True cases (the reference's own default, M:3353, and
    what its GUI buttons pass, M:2655-2663): every run builds a fresh engine.  `gates` holds the engine settings a gated
    case was run under (strings as written with settings_set) and the logistic model, if any."""
    out = []
    for name, n, seed, trials, tl, levels_name, prox, gates in [
        ("mc_n390_fix30", 390, 21, 500, 1, "fix30", 0.1, None),
        ("mc_n120_fix30", 120, 22, 500, 1, "fix30", 0.1, None),
        ("mc_n390_blue4_tl2", 390, 23, 300, 2, "blue4", 0.1, None),
        ("mc_n600_fix30_prox05", 600, 24, 200, 1, "fix30", 0.05, None),
        ("gated_n390_fix30", 390, 21, 500, 1, "fix30", 0.1, dict()),
        ("gated_n600_fix30_tl3", 600, 25, 300, 3, "fix30", 0.1, dict()),
        ("gated_n1200_fix30_tl6_overtouch", 1200, 26, 100, 6, "fix30", 0.1, dict(sigma=0.05)),
        ("gated_n390_blue4_qmin075", 390, 23, 300, 2, "blue4", 0.1, dict(q_min_prob=0.75)),
        ("gated_n600_fix30_lr_and", 600, 27, 200, 2, "fix30", 0.1, dict(lr=LR_A)),
        ("gated_n600_fix30_lr_blend", 600, 28, 200, 2, "fix30", 0.1, dict(lr=LR_B, use_blend=True, w_rules=0.6, w_ml=0.5)),
        ("gated_n390_fix30_flat_ticks", 390, 29, 100, 4, "fix30", 0.1, dict(sigma=0.02)),
    ]:
        g = gates or {}
        bars = synth_ohlc(n, seed, sigma=g.get("sigma", 0.15))
        levels = levels_by_name(levels_name)
        settings = {"CONTACT_PROX": repr(prox), "ENABLE_VETO": "0"}
        if "q_min_prob" in g:
            settings["Q_MIN_PROB"] = repr(g["q_min_prob"])
        if g.get("use_blend"):
            settings.update(USE_BLEND="1", W_RULES=repr(g["w_rules"]), W_ML=repr(g["w_ml"]))
        model = rh.make_sklearn_model(g["lr"]["coef"], g["lr"]["intercept"]) if "lr" in g else None
        wg = gates is not None
        dict_bars = [dict(t=60000 * i, o=bars["o"][i], h=bars["h"][i], l=bars["l"][i], c=bars["c"][i])
                     for i in range(n)]
        mc = rh.run_reference_monte_carlo(dict_bars, levels, settings, lr_model=model, n=n, touch_limit=tl, trials=trials,
                                          with_gates=wg, seed=seed)
        sim_msg, sim_rows = rh.run_reference_last_bars(dict_bars, levels, settings, lr_model=model, n=n, touch_limit=tl,
                                                       with_gates=wg)
        out.append(dict(name=name, n=n, seed=seed, trials=trials, touch_limit=tl, levels=levels_name,
                        contact_prox=prox, with_gates=wg, gates={k: v for k, v in g.items() if k != "sigma"},
                        bars=bars, mc_message=mc[2] if mc else None,
                        sim_message=sim_msg[2] if sim_msg else None,
                        sim_rows=[list(r) for r in sim_rows]))
        print(name, (mc[2] if mc else None), "| GATE_SKIP rows:", sum(1 for r in sim_rows if r[1].startswith("GATE_SKIP")))
    with open(os.path.join(GOLD, "jitter_mc.json"), "w") as f:
        json.dump(dict(source="reference QMMXApp.simulate_monte_carlo / simulate_last_bars, with_gates False and True",
                       cases=out), f)


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    if "--jitter-only" not in sys.argv:
        gen_engine_cases()
    gen_jitter_cases()
