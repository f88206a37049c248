"""GPU parity tests: libqmmx_mc.so (through its C ABI) against the oracle on the same inputs.

Bar: bit-exact entry/exit bar indices, sides, level ids, reason codes, trade counts and every integer
counter; per-trade PnL equal as fp32; fp64 sums within 1e-9 relative (the north-star asks 1e-6).
"""
import numpy as np
import pytest

from oracle import c_oracle as CO
from oracle import headless as H
from qmmx_monolithic_monte_carlo_b200 import abi
from qmmx_monolithic_monte_carlo_b200 import engine as E
from qmmx_monolithic_monte_carlo_b200 import fixtures as F
from tests.golden_util import load_engine_cases, load_jitter_cases, levels_by_name
from tests.test_c_oracle import c_inputs

pytestmark = pytest.mark.gpu

ENGINE_CASES = load_engine_cases()


@pytest.fixture(scope="module")
def eng():
    e = E.MCEngine(0)
    yield e
    e.close()


def assert_agg_equal(a, b, *, sums_rel=1e-9):
    """a: abi.Agg from the GPU, b: abi.Agg from the oracle."""
    for f in ("n_paths", "n_steps", "n_trades", "n_wins", "n_open", "n_trade_overflow"):
        assert getattr(a, f) == getattr(b, f), f
    assert list(a.entry_hist) == list(b.entry_hist)
    assert list(a.exit_hist) == list(b.exit_hist)
    assert list(a.pnl_hist) == list(b.pnl_hist)
    for f in ("sum_pnl", "sum_pnl2", "sum_path_pnl2", "sum_open_mtm"):
        x, y = getattr(a, f), getattr(b, f)
        assert x == pytest.approx(y, rel=sums_rel, abs=1e-9), f
    if b.n_paths:
        assert a.max_path_pnl == b.max_path_pnl and a.min_path_pnl == b.min_path_pnl


def assert_trades_equal(tr, cnt, otr, ocnt):
    assert np.array_equal(cnt, ocnt)
    for f in ("entry_bar", "exit_bar", "side", "reason", "level_id", "entry", "exit", "pnl"):
        m = np.arange(tr.shape[1])[None, :] < np.minimum(cnt, tr.shape[1])[:, None]
        assert np.array_equal(tr[f][m], otr[f][m]), f


def configure(eng, lv_price, lv_solid, params, lr=None, pol=None):
    eng.set_levels_raw(lv_price, lv_solid)
    eng.set_params(params)
    eng.set_lr(lr)
    eng.set_policy(pol)


@pytest.mark.parametrize("case", ENGINE_CASES, ids=[c["name"] for c in ENGINE_CASES])
def test_gpu_matches_reference_golden(eng, case):
    lv, params, lr_a, pol_a = c_inputs(case)
    configure(eng, [l.price for l in lv], [l.solid for l in lv], params, lr_a, pol_a)
    close = case["close"]
    eng.upload_bars(close)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=close.shape)
    assert np.array_equal(codes, case["codes"])
    for j in range(case["paths"]):
        gt = case["trades"][j]
        assert cnt[j] == len(gt)
        for g, t in zip(gt, tr[j]):
            assert (g["entry_bar"], g["exit_bar"], g["side"], g["level"], g["reason"]) == \
                   (t["entry_bar"], t["exit_bar"], t["side"], t["level_id"], t["reason"])
            assert np.float32(g["pnl"]) == t["pnl"] and np.float32(g["entry"]) == t["entry"]
    a = res[0].agg
    assert list(a.entry_hist) == np.bincount(case["codes"].ravel(), minlength=20).tolist()


PARAM_SETS = [
    dict(contact_prox=0.1, cooldown_s=120),
    dict(contact_prox=0.1, cooldown_s=8),
    dict(contact_prox=0.05, cooldown_s=300, stop_padding=0.6, tp_padding=0.5),
    dict(contact_prox=0.2, cooldown_s=0, q_min_prob=0.3, overtouch_limit=7),
    dict(contact_prox=0.1, cooldown_s=61, q_min_prob=0.9),
]


@pytest.mark.parametrize("pi", range(len(PARAM_SETS)))
@pytest.mark.parametrize("levels", ["fix30", "blue4"])
def test_gpu_matches_oracle_random_paths(eng, pi, levels):
    rows, price, solid = E.sort_levels(levels_by_name(levels))
    params = abi.make_params(**PARAM_SETS[pi])
    close = F.host_gbm_closes(5000 + 37 * pi, 390, seed=100 + pi, round_cents=(pi % 2 == 1))
    configure(eng, price, solid, params)
    eng.upload_bars(close)
    res, tr, cnt, codes = eng.run_scan(trades=48, codes=True, paths_steps=close.shape)
    o = CO.run_paths(close, price, solid, params, want_codes=True, max_trades=48, n_threads=4)
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])
    # aggregate-only (production) kernel gives the same aggregates as the parity kernel
    res2 = eng.run_scan()
    o["agg"].n_trade_overflow = 0            # only the parity-mode record buffer can overflow
    assert_agg_equal(res2[0].agg, o["agg"])


def _lr_tables(rng, days):
    lr = np.column_stack([rng.normal(0, 0.5, days), rng.normal(0, 0.5, days), rng.normal(-0.3, 0.4, days),
                          rng.normal(0, 0.8, days), rng.normal(0.6, 0.5, days)])
    pol = rng.normal(0, 0.5, (days, 3, 7))
    pol[:, 0, 0] += 1.0; pol[:, 1, 0] += 1.0; pol[:, 2, 0] -= 1.0
    return lr, pol


@pytest.mark.parametrize("mode", ["and", "blend", "policy", "and+policy", "ml_off"])
def test_gpu_lr_gates_multi_day(eng, mode):
    rng = np.random.default_rng(11)
    days = 3
    lr, pol = _lr_tables(rng, days)
    rows, price, solid = E.sort_levels(F.levels_30())
    kw = dict(contact_prox=0.1, cooldown_s=120)
    if mode == "blend":
        kw.update(use_blend=True, w_rules=0.6, w_ml=0.4)
    if "policy" in mode:
        kw.update(disable_policy_gate=False)
    if mode == "ml_off":
        kw.update(disable_ml_gate=True)
    params = abi.make_params(**kw)
    close = F.host_gbm_closes(3000, 390 * days, seed=7, sigma=0.10)
    use_lr = None if mode == "policy" else lr
    use_pol = pol if "policy" in mode else None
    configure(eng, price, solid, params, use_lr, use_pol)
    eng.upload_bars(close)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=close.shape)
    o = CO.run_paths(close, price, solid, params, lr=use_lr, policy=use_pol, want_codes=True, max_trades=64,
                     n_threads=4)
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])
    hist = res[0].entry_hist
    if mode in ("and", "and+policy"):
        assert hist["ML_CONF_LOW"] > 0
    if mode == "blend":
        assert hist["COMBINED_LOW"] > 0
    if "policy" in mode:
        assert hist["ONLINE_POLICY"] > 0
    assert hist["OK"] > 0


def test_gpu_hl_exit_mode(eng):
    g = abi.make_gbm(sub_steps=4)
    c, h, l = CO.gen_gbm(77, 0, 4000, 390, g, ohlc=True)
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=True)
    configure(eng, price, solid, params)
    eng.upload_bars(c, h, l)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=c.shape)
    o = CO.run_paths(c, price, solid, params, high=h, low=l, want_codes=True, max_trades=64, n_threads=4)
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])


def test_gpu_edge_cases(eng):
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120)
    # no levels: every step NOLEVELS
    configure(eng, [], [], params)
    close = F.host_gbm_closes(100, 50, seed=3)
    eng.upload_bars(close)
    r = eng.run_scan()[0]
    assert r.entry_hist["NOLEVELS"] == 100 * 50 and r.n_trades == 0
    # ragged path count (not a multiple of 32), 1 step, flat path, trade-record overflow
    configure(eng, price, solid, params)
    for paths, steps in [(1, 1), (33, 7), (257, 391), (31, 2)]:
        close = F.host_gbm_closes(paths, steps, seed=paths)
        eng.upload_bars(close)
        res, tr, cnt, codes = eng.run_scan(trades=2, codes=True, paths_steps=close.shape)
        o = CO.run_paths(close, price, solid, params, want_codes=True, max_trades=2)
        assert np.array_equal(codes, o["codes"])
        assert_trades_equal(tr, cnt, o["trades"], o["counts"])
        assert_agg_equal(res[0].agg, o["agg"])
    flat = np.full((40, 64), 648.76, dtype=np.float32)
    flat[20:, :32] = 648.80
    eng.upload_bars(flat)
    res, tr, cnt, codes = eng.run_scan(trades=4, codes=True, paths_steps=flat.shape)
    o = CO.run_paths(flat, price, solid, params, want_codes=True, max_trades=4)
    assert np.array_equal(codes, o["codes"]) and (codes == abi.ENTRY_CODES.index("DIR_UNKNOWN")).sum() >= 64 * 20
    # empty input
    eng.upload_bars(np.zeros((0, 0), dtype=np.float32))
    r = eng.run_scan()[0]
    assert r.n_paths == 0 and r.n_steps == 0


def test_gpu_tiny_prices_flag(eng):
    # prices around 0.01: the +-1e-9 guard band matters; QMMX_F_TINY_PRICES evaluates it in float64
    rng = np.random.default_rng(5)
    base = np.float32(0.01)
    steps, paths = 200, 512
    inc = rng.integers(-2, 3, size=(steps, paths)).astype(np.float32) * np.float32(9.3e-10)
    close = (base + np.cumsum(inc, axis=0)).astype(np.float32)
    price = np.array([0.01, 0.0100001]); solid = np.array([1, 0], dtype=np.uint8)
    params = abi.make_params(contact_prox=1e-6, stop_padding=3e-8, tp_padding=3e-8, cooldown_s=60)
    params.flags |= abi.F_TINY_PRICES
    configure(eng, price, solid, params)
    eng.upload_bars(close)
    res, tr, cnt, codes = eng.run_scan(trades=32, codes=True, paths_steps=close.shape)
    o = CO.run_paths(close, price, solid, params, want_codes=True, max_trades=32)
    assert np.array_equal(codes, o["codes"])
    assert_agg_equal(res[0].agg, o["agg"])


def test_gpu_param_grid_sweep(eng):
    rows, price, solid = E.sort_levels(F.levels_30())
    grid = [abi.make_params(contact_prox=0.1, stop_padding=sp, tp_padding=tp, cooldown_s=cd)
            for sp in (0.10, 0.35, 0.85) for tp in (0.15, 0.5) for cd in (0, 60, 900)]
    grid.append(abi.make_params(contact_prox=0.05, q_min_prob=0.5, cooldown_s=120))   # second table
    close = F.host_gbm_closes(4096, 390, seed=21)
    configure(eng, price, solid, grid)
    eng.upload_bars(close)
    res = eng.run_scan()
    assert len(res) == len(grid)
    for p, r in zip(grid, res):
        o = CO.run_paths(close, price, solid, p, n_threads=4)
        assert_agg_equal(r.agg, o["agg"])


def test_gpu_host_pipeline_equals_resident(eng):
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120)
    close = F.host_gbm_closes(70001, 390, seed=9)
    configure(eng, price, solid, params)
    eng.upload_bars(close)
    a = eng.run_scan()[0]
    eng.pin(close)
    try:
        b = eng.run_scan_host(close, chunk_paths=16384)[0]
    finally:
        eng.unpin(close)
    o = CO.run_paths(close, price, solid, params, n_threads=8)
    assert_agg_equal(a.agg, o["agg"])
    assert_agg_equal(b.agg, o["agg"])


def test_gpu_philox_gbm_generator(eng):
    g = abi.make_gbm()
    paths, steps = 4096, 390
    eng.gen_gbm(20250829, 1000, paths, steps, g)
    c, _, _ = eng.download_bars(paths, steps)
    oc, _, _ = CO.gen_gbm(20250829, 1000, paths, steps, g)
    # same Philox stream and operation order; MUFU vs libm transcendentals differ in the last bits
    assert np.allclose(c, oc, rtol=2e-5, atol=0)
    lr = np.diff(np.log(np.vstack([np.full((1, paths), g.s0, dtype=np.float64), c.astype(np.float64)])), axis=0)
    sd = 0.15 * np.sqrt(1 / (252 * 390))
    assert abs(lr.std() / sd - 1) < 0.01 and abs(lr.mean()) < 5 * sd / np.sqrt(lr.size)
    # sharding: path ids are global
    eng.gen_gbm(20250829, 1000 + 96, 64, steps, g)
    c2, _, _ = eng.download_bars(64, steps)
    assert np.array_equal(c2, c[:, 96:160])
    # OHLC recipe
    g4 = abi.make_gbm(sub_steps=4, round_cents=True)
    eng.gen_gbm(5, 0, 1000, 100, g4)
    c4, h4, l4 = eng.download_bars(1000, 100, hl=True)
    o4 = CO.gen_gbm(5, 0, 1000, 100, g4, ohlc=True)
    assert np.all(h4 >= c4) and np.all(l4 <= c4)
    # MUFU vs libm last-bit differences can flip a cent rounding: most bars equal, none off by more than a cent
    assert np.mean(c4 == o4[0]) > 0.85, np.mean(c4 == o4[0])
    assert np.allclose(c4, o4[0], atol=0.011)


def test_gpu_fused_equals_generate_then_scan(eng):
    rows, price, solid = E.sort_levels(F.levels_30())
    rng = np.random.default_rng(2)
    lr, _ = _lr_tables(rng, 1)
    params = abi.make_params(contact_prox=0.1, cooldown_s=120)
    configure(eng, price, solid, params, lr)
    g = abi.make_gbm()
    paths, steps = 100_003, 390
    fused = eng.run_fused_gbm(42, 7, paths, steps, g)[0]
    eng.gen_gbm(42, 7, paths, steps, g)
    scan = eng.run_scan()[0]
    assert_agg_equal(fused.agg, scan.agg, sums_rel=1e-12)
    # and the device-generated bars scanned by the oracle give the same aggregates again
    c, _, _ = eng.download_bars(paths, steps)
    o = CO.run_paths(c, price, solid, params, lr=lr, n_threads=8)
    assert_agg_equal(scan.agg, o["agg"])
    # statistical tier: oracle-generated (libm) paths vs device-generated paths, same seed
    oc, _, _ = CO.gen_gbm(42, 7, 20000, steps, g)
    oo = CO.run_paths(oc, price, solid, params, lr=lr, n_threads=8)["agg"]
    eng.gen_gbm(42, 7, 20000, steps, g)
    dd = eng.run_scan()[0].agg
    assert abs(dd.n_trades - oo.n_trades) <= 5 * np.sqrt(oo.n_trades) * 0.05 + 20
    # OHLC fused
    p2 = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=True)
    configure(eng, price, solid, p2)
    g4 = abi.make_gbm(sub_steps=4)
    f4 = eng.run_fused_gbm(3, 0, 20011, 390, g4)[0]
    eng.gen_gbm(3, 0, 20011, 390, g4)
    s4 = eng.run_scan()[0]
    assert_agg_equal(f4.agg, s4.agg, sums_rel=1e-12)


def test_gpu_bootstrap_generator(eng):
    g = abi.make_gbm(sub_steps=4)
    wc, wh, wl = CO.gen_gbm(5, 0, 1, 40, g, ohlc=True)
    wo = np.concatenate([[g.s0], wc[:-1, 0]]).astype(np.float32)
    window = np.stack([wo, wh[:, 0], wl[:, 0], wc[:, 0]], axis=1)
    paths, steps = 3000, 5 * 390
    eng.gen_bootstrap(9, 11, paths, steps, window, 5)
    c, h, l = eng.download_bars(paths, steps, hl=True)
    oc, oh, ol = CO.gen_bootstrap(9, 11, paths, steps, window, 5)
    assert np.array_equal(c, oc) and np.array_equal(h, oh) and np.array_equal(l, ol)   # integer draws + IEEE products
    # config 5: multi-day paths, per-day LR weights, exits on high/low
    rng = np.random.default_rng(4)
    lr, pol = _lr_tables(rng, 5)
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=True)
    configure(eng, price, solid, params, lr)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=c.shape)
    o = CO.run_paths(c, price, solid, params, high=h, low=l, lr=lr, want_codes=True, max_trades=64, n_threads=8)
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])


def test_gpu_lr_score_matches_fp32_reference(eng):
    import torch
    rng = np.random.default_rng(0)
    for dim, heads in [(7, 3), (4, 1), (32, 8)]:
        x = rng.normal(0, 1, (10007, dim)).astype(np.float32)
        w = rng.normal(0, 1, (heads, dim)).astype(np.float32)
        b = rng.normal(0, 1, heads).astype(np.float32)
        out = eng.lr_score(x, w, b)
        ref = torch.sigmoid(torch.from_numpy(x) @ torch.from_numpy(w).T + torch.from_numpy(b)).numpy()
        assert np.allclose(out, ref, atol=2e-6, rtol=0)     # fp32 FMA accumulation, tolerance 2e-6 absolute
        out2 = eng.lr_score(x, w, None)
        ref2 = torch.sigmoid(torch.from_numpy(x) @ torch.from_numpy(w).T).numpy()
        assert np.allclose(out2, ref2, atol=2e-6, rtol=0)


@pytest.mark.parametrize("case", load_jitter_cases()[:3], ids=lambda c: c["name"])
def test_gpu_jitter_mc_replay_is_exact(eng, case):
    lv = H.sort_levels(levels_by_name(case["levels"]))
    P = H.Params(contact_prox=case["contact_prox"])
    bars = case["bars"]
    n = case["n"]
    c, h, l = list(bars["c"])[-n:], list(bars["h"])[-n:], list(bars["l"])[-n:]
    cands = H.find_candidates(c, lv, P, touch_limit=case["touch_limit"])
    if not cands:
        pytest.skip("no candidates")
    trials = 2000
    rng = np.random.default_rng(1)
    normals = rng.standard_normal((trials, len(cands), 4))
    uniforms = rng.random((trials, len(cands)))
    sig = (0.02, 0.01, 0.005, 0.003)
    ref = H.jitter_monte_carlo(bars, lv, P, n=n, touch_limit=case["touch_limit"], trials=trials,
                               level_jitter_std=sig[0], entry_slip_std=sig[1], stop_slip_std=sig[2],
                               target_slip_std=sig[3], draws=(normals, uniforms))
    out = eng.run_jitter(h, l, cands, P.stop_padding, P.tp_padding, sig, trials, replay=(normals, uniforms))
    assert np.array_equal(out["totals"], np.array(ref["totals"]))
    assert np.array_equal(out["drawdowns"], np.array(ref["drawdowns"]))
    s = H.mc_summary(out["totals"].tolist(), out["drawdowns"].tolist(), out["wins"].tolist(),
                     out["losses"].tolist(), out["opens"].tolist())
    for k in ("mean", "median", "stdev", "best", "worst", "var05", "cvar05", "dd_mean", "dd_worst", "wins",
              "losses", "opens"):
        assert s[k] == ref[k], k
    # statistical tier: device Philox draws vs the reference's Mersenne-Twister draws
    big = 200_000
    dev = eng.run_jitter(h, l, cands, P.stop_padding, P.tp_padding, (0.02, 0.01, 0.0, 0.0), big, seed=99)
    refs = H.jitter_monte_carlo(bars, lv, P, n=n, touch_limit=case["touch_limit"], trials=4000, seed=5)
    se = refs["stdev"] / np.sqrt(4000) + dev["totals"].std() / np.sqrt(big)
    assert abs(dev["totals"].mean() - refs["mean"]) < 5 * se + 1e-9     # 5-sigma bound on the mean total R


def test_gpu_monte_carlo_dropin_message(eng):
    """`montecarlo.monte_carlo` (reference call shape) with replayed draws gives the oracle's summary text."""
    from qmmx_monolithic_monte_carlo_b200 import montecarlo as MC
    from tests.golden_util import format_mc_message
    case = load_jitter_cases()[0]
    lv = H.sort_levels(levels_by_name(case["levels"]))
    P = H.Params(contact_prox=case["contact_prox"])
    n, trials = case["n"], 500
    cands = H.find_candidates(list(case["bars"]["c"])[-n:], lv, P, touch_limit=case["touch_limit"])
    rng = np.random.default_rng(3)
    normals = rng.standard_normal((trials, len(cands), 4)); uniforms = rng.random((trials, len(cands)))
    ref = H.jitter_monte_carlo(case["bars"], lv, P, n=n, touch_limit=case["touch_limit"], trials=trials,
                               draws=(normals, uniforms))
    got = MC.monte_carlo(eng, case["bars"], levels_by_name(case["levels"]), contact_prox=case["contact_prox"], n=n,
                         touch_limit=case["touch_limit"], trials=trials, with_gates=False, replay=(normals, uniforms))
    assert got.message() == format_mc_message(ref)
    assert got.candidates == ref["candidates"]


@pytest.mark.parametrize("case", [c for c in load_jitter_cases() if c.get("with_gates")], ids=lambda c: c["name"])
def test_gpu_monte_carlo_dropin_with_gates(eng, case):
    """`montecarlo.monte_carlo(..., with_gates=True)` -- the reference's default and what its GUI passes (M:2662): the
    gated candidate pass finds the reference's candidate count (read from its own MC text) and, with the oracle's draws
    replayed, the device walk gives the oracle's summary text; the gate removed candidates the detector alone keeps."""
    from qmmx_monolithic_monte_carlo_b200 import montecarlo as MC
    from tests.golden_util import format_mc_message, jitter_case_setup
    lv = H.sort_levels(levels_by_name(case["levels"]))
    wg, P, lr, gs = jitter_case_setup(case)
    n, trials = case["n"], 400
    closes = list(case["bars"]["c"])[-n:]
    cands = H.find_candidates(closes, lv, P, touch_limit=case["touch_limit"], with_gates=True, lr=lr)
    ungated = H.find_candidates(closes, lv, P, touch_limit=case["touch_limit"])
    assert [c["i"] for c in cands] != [c["i"] for c in ungated]
    assert case["mc_message"].startswith(f"Candidates: {len(cands)} |")
    rng = np.random.default_rng(4)
    normals = rng.standard_normal((trials, len(cands), 4)); uniforms = rng.random((trials, len(cands)))
    ref = H.jitter_monte_carlo(case["bars"], lv, P, n=n, touch_limit=case["touch_limit"], trials=trials,
                               draws=(normals, uniforms), with_gates=True, lr=lr)
    got = MC.monte_carlo(eng, case["bars"], levels_by_name(case["levels"]), contact_prox=case["contact_prox"], n=n,
                         touch_limit=case["touch_limit"], trials=trials, replay=(normals, uniforms), **gs)   # with_gates defaults to True
    assert got.message() == format_mc_message(ref) and got.candidates == len(cands)


def test_gpu_full_size_properties(eng):
    """BASELINE config sizes through size-independent properties: counters partition the steps, the histogram
    counts every path, shards add up to the whole (linearity), and the run is idempotent."""
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120)
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    configure(eng, price, solid, params, lr)
    g = abi.make_gbm()
    P, T = 1_000_000, 390
    whole = eng.run_fused_gbm(11, 0, P, T, g)[0]
    a = whole.agg
    assert a.n_paths == P and a.n_steps == P * T == sum(a.entry_hist)
    assert sum(a.pnl_hist) == P and a.n_trades == a.exit_hist[0] + a.exit_hist[1] and a.n_open == a.exit_hist[3]
    assert a.entry_hist[0] == a.n_trades + a.n_open and 0 < a.n_wins < a.n_trades
    parts = [eng.run_fused_gbm(11, lo, hi - lo, T, g)[0].agg for lo, hi in [(0, 300_001), (300_001, 1_000_000)]]
    for f in ("n_paths", "n_steps", "n_trades", "n_wins", "n_open"):
        assert getattr(a, f) == sum(getattr(p, f) for p in parts), f
    assert list(a.entry_hist) == [x + y for x, y in zip(parts[0].entry_hist, parts[1].entry_hist)]
    assert list(a.pnl_hist) == [x + y for x, y in zip(parts[0].pnl_hist, parts[1].pnl_hist)]
    assert a.sum_pnl == pytest.approx(parts[0].sum_pnl + parts[1].sum_pnl, rel=1e-9)
    assert a.max_path_pnl == max(p.max_path_pnl for p in parts) and a.min_path_pnl == min(p.min_path_pnl for p in parts)
    again = eng.run_fused_gbm(11, 0, P, T, g)[0].agg
    assert list(again.entry_hist) == list(a.entry_hist) and again.n_wins == a.n_wins
    # resident-bar scan of the same generated paths agrees with the fused kernel
    eng.gen_gbm(11, 0, P, T, g)
    scan = eng.run_scan()[0].agg
    assert list(scan.entry_hist) == list(a.entry_hist) and list(scan.pnl_hist) == list(a.pnl_hist)
    # and the oracle itself at the full C2 size: all 1M paths of the resident bars through the C port on every host
    # thread (~3 s), every counter, both histograms, the PnL histogram and the float64 sums
    import os
    bars = eng.download_bars(P, T)[0]
    o = CO.run_paths(bars, price, solid, params, lr=lr, n_threads=os.cpu_count() or 4)["agg"]
    assert_agg_equal(scan, o)
    assert_agg_equal(a, o)


@pytest.mark.parametrize("hl", [False, True], ids=["close", "hl"])
def test_gpu_lean_kernels_equal_first_generation(eng, hl, monkeypatch):
    """The production (lean) scan and fused kernels against the first-generation lockstep kernels (QMMX_SCAN_V1=1,
    read by qmmx_create) on the full C2 size: two independent implementations of the same engine, every integer
    aggregate identical and the per-path trade records identical on a 64k-path slice."""
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=hl)
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    g = abi.make_gbm(sub_steps=4 if hl else 1)
    P, T = 1_000_000, 390
    monkeypatch.setenv("QMMX_SCAN_V1", "1")
    v1 = E.MCEngine(0)
    monkeypatch.delenv("QMMX_SCAN_V1")
    try:
        out = []
        for e in (eng, v1):
            configure(e, price, solid, params, lr)
            fused = e.run_fused_gbm(5, 0, P, T, g)[0].agg
            e.gen_gbm(5, 0, P, T, g)
            scan = e.run_scan()[0].agg
            e.gen_gbm(5, 0, 65_536, T, g)
            res, tr, cnt, _ = e.run_scan(trades=48, paths_steps=(T, 65_536))
            out.append((fused, scan, res[0].agg, tr, cnt))
        (f0, s0, a0, tr0, c0), (f1, s1, a1, tr1, c1) = out
        assert_agg_equal(f0, f1)
        assert_agg_equal(s0, s1)
        assert_agg_equal(f0, s0)
        assert_agg_equal(a0, a1)
        assert_trades_equal(tr0, c0, tr1, c1)
    finally:
        v1.close()


@pytest.mark.parametrize("variant", ["lean", "lean_hl", "first_generation(tiny)"])
def test_gpu_volume_soft_veto(eng, variant):
    """per-bar volume feeds `_calc_volume_slope` (M:1796-1811) and the soft veto
    (M:1678-1705, M:1773-1794), which sits between the confidence gate and the LR gate.  Codes, trades and
    aggregates against the oracle (itself pinned against the reference with volume in recent_bars)."""
    from tests.test_c_oracle import _volume_for, VETO_LONG, VETO_SHORT
    rng = np.random.default_rng(12)
    hl = variant == "lean_hl"
    tiny = variant.startswith("first")
    if hl:
        close, high, low = CO.gen_gbm(41, 0, 4096, 390, abi.make_gbm(sub_steps=4), ohlc=True)
    else:
        close, high, low = F.host_gbm_closes(4096, 390, seed=33), None, None
    vol = _volume_for(close, rng)
    rows, price, solid = E.sort_levels(F.levels_30())
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    params = abi.make_params(contact_prox=0.1, cooldown_s=60, enable_veto=True, veto_prox=0.08, exit_on_hl=hl,
                             tiny_prices=tiny)
    configure(eng, price, solid, params, lr)
    eng.upload_bars(close, high, low)
    # veto enabled but no volume: inert, exactly the run without the flag
    inert = eng.run_scan()[0].agg
    assert inert.entry_hist[VETO_LONG] == 0 and inert.entry_hist[VETO_SHORT] == 0
    eng.upload_volume(vol)
    assert np.array_equal(eng.download_volume(*close.shape[::-1]), vol)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=close.shape)
    o = CO.run_paths(close, price, solid, params, high=high, low=low, volume=vol, lr=lr, want_codes=True,
                     max_trades=64, n_threads=8)
    assert o["agg"].entry_hist[VETO_LONG] > 100 and o["agg"].entry_hist[VETO_SHORT] > 100
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])
    assert_agg_equal(eng.run_scan()[0].agg, o["agg"])          # aggregate-only kernel variant
    eng.upload_volume(None)
    assert_agg_equal(eng.run_scan()[0].agg, inert)


def test_gpu_bootstrap_volume_plane_and_veto(eng):
    g = abi.make_gbm(sub_steps=4)
    wc, wh, wl = CO.gen_gbm(5, 0, 1, 40, g, ohlc=True)
    wo = np.concatenate([[g.s0], wc[:-1, 0]]).astype(np.float32)
    window = np.stack([wo, wh[:, 0], wl[:, 0], wc[:, 0]], axis=1)
    wv = np.random.default_rng(2).uniform(1e4, 3e5, 40).astype(np.float32)
    paths, steps = 2000, 2 * 390
    eng.gen_bootstrap(9, 11, paths, steps, window, 5, window_volume=wv)
    c, h, l = eng.download_bars(paths, steps, hl=True)
    v = eng.download_volume(paths, steps)
    oc, oh, ol, ov = CO.gen_bootstrap(9, 11, paths, steps, window, 5, window_volume=wv)
    assert np.array_equal(c, oc) and np.array_equal(h, oh) and np.array_equal(l, ol) and np.array_equal(v, ov)
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=True, enable_veto=True)
    configure(eng, price, solid, params, None)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=c.shape)
    o = CO.run_paths(c, price, solid, params, high=h, low=l, volume=v, want_codes=True, max_trades=64, n_threads=8)
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])
    eng.gen_bootstrap(9, 11, paths, steps, window, 5)           # without volume: the plane is dropped
    with pytest.raises(E.QmmxError):
        eng.download_volume(paths, steps)


@pytest.mark.parametrize("hl", [True, False], ids=["hl", "close"])
def test_gpu_fused_bootstrap_equals_generate_then_scan(eng, hl):
    """qmmx_run_fused_bootstrap draws the bars of qmmx_gen_bootstrap (checked against the oracle generator above)
    and walks them in registers: every aggregate equals generate-then-scan, multi-day LR tables included."""
    g = abi.make_gbm(sub_steps=4)
    wc, wh, wl = CO.gen_gbm(5, 0, 1, 40, g, ohlc=True)
    wo = np.concatenate([[g.s0], wc[:-1, 0]]).astype(np.float32)
    window = np.stack([wo, wh[:, 0], wl[:, 0], wc[:, 0]], axis=1)
    rng = np.random.default_rng(4)
    lr, pol = _lr_tables(rng, 3)
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=hl)
    configure(eng, price, solid, params, lr)
    paths, steps = 50_001, 3 * 390
    fused = eng.run_fused_bootstrap(9, 7, paths, steps, window, 5)[0].agg
    eng.gen_bootstrap(9, 7, paths, steps, window, 5)
    scan = eng.run_scan()[0].agg
    assert scan.n_trades > 0
    assert_agg_equal(fused, scan, sums_rel=1e-12)
    # sharding: two shards of the fused run add up to the whole
    a = eng.run_fused_bootstrap(9, 7, 20_000, steps, window, 5)[0].agg
    b = eng.run_fused_bootstrap(9, 7 + 20_000, paths - 20_000, steps, window, 5)[0].agg
    assert list(fused.entry_hist) == [x + y for x, y in zip(a.entry_hist, b.entry_hist)]
    assert list(fused.pnl_hist) == [x + y for x, y in zip(a.pnl_hist, b.pnl_hist)]


@pytest.mark.parametrize("variant", ["lean", "first_generation(tiny)"])
def test_gpu_regime_guard_gate(eng, variant):
    """the AccumulationBreakoutGuard regime plane (regime_kernel pre-pass over close + volume) and the
    904 gate (M:1652-1666) between the confidence gate and the soft veto; against the oracle, which is pinned
    against the reference driven through ingest_tick."""
    from tests.test_c_oracle import breakout_paths
    closes, vol = breakout_paths(1536, 8)
    rows, price, solid = E.sort_levels(F.levels_30())
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    params = abi.make_params(contact_prox=0.1, cooldown_s=60, enable_veto=True, tiny_prices=variant.startswith("first"))
    configure(eng, price, solid, params, lr)
    eng.upload_bars(closes)
    eng.upload_volume(vol)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=closes.shape)
    o = CO.run_paths(closes, price, solid, params, volume=vol, lr=lr, want_codes=True, max_trades=64, n_threads=8)
    acc = abi.ENTRY_CODES.index("ACC_BREAKOUT_GATE")
    assert o["agg"].entry_hist[acc] > 100
    assert set(np.unique(CO.regime_series(closes[:, :64], vol[:, :64]))) == {0, 1, 2, 3}
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])
    assert_agg_equal(eng.run_scan()[0].agg, o["agg"])


@pytest.mark.gpu
@pytest.mark.parametrize("case", load_jitter_cases(), ids=lambda c: c["name"])
def test_gpu_last_bars_dropin_matches_reference_rows(eng, case):
    """Row a15: `montecarlo.last_bars` (device walk) reproduces the SIM/TRADE rows and the "Sim complete" text the
    reference's `simulate_last_bars` wrote for the golden bars (tests/golden/jitter_mc.json)."""
    from qmmx_monolithic_monte_carlo_b200 import montecarlo as MC
    from tests.golden_util import jitter_case_setup
    wg, P, lr, gs = jitter_case_setup(case)
    skips = []
    trades, s = MC.last_bars(eng, case["bars"], levels_by_name(case["levels"]), contact_prox=case["contact_prox"],
                             n=case["n"], touch_limit=case["touch_limit"], with_gates=wg, gate_skips=skips, **gs)
    assert s.message() == case["sim_message"]
    assert [list(r) for r in skips] == [r for r in case["sim_rows"] if r[1].startswith("GATE_SKIP")]   # M:3594-3597
    trows = [r[2] for r in case["sim_rows"] if r[1] == "TRADE"]
    assert len(trows) == len(trades)
    for txt, t in zip(trows, trades):
        usd = (t["exit"] - t["entry"]) * (1.0 if t["side"] == "long" else -1.0)
        assert txt == (f"{t['outcome'].upper():5s} | {t['side']:5s} @ {t['entry']:.2f} → {t['exit']:.2f} "
                       f"| lvl {t['lvl']:.2f} | R={t['R']:+.2f} | ${usd:+.2f} | prox={t['prox']:.03f} touch#{t['touch#']}")


@pytest.mark.gpu
def test_gpu_last_bars_walk_equals_oracle_on_long_series(eng):
    """12000 bars (the ABI maximum), ~hundreds of candidates, ties where one bar reaches stop and target."""
    from qmmx_monolithic_monte_carlo_b200 import montecarlo as MC
    rng = np.random.default_rng(15)
    n = 12000
    c = 648.9 + np.cumsum(rng.normal(0, 0.12, n))
    spread = np.abs(rng.normal(0, 0.25, n))
    bars = dict(o=c.tolist(), c=c.tolist(), h=(c + spread).tolist(), l=(c - spread).tolist())
    lv = H.sort_levels(F.levels_30())
    P = H.Params(contact_prox=0.1, stop_padding=0.2, tp_padding=0.15)
    ref_trades, ref_s = H.simulate_last_bars(bars, lv, P, n=n, touch_limit=12)
    trades, s = MC.last_bars(eng, bars, F.levels_30(), contact_prox=0.1, stop_padding=0.2, tp_padding=0.15, n=n,
                             touch_limit=12)
    assert len(trades) == len(ref_trades) >= 100
    assert [(t["i"], t["j"], t["outcome"], t["R"], t["exit"]) for t in trades] == \
           [(t["i"], t["j"], t["outcome"], t["R"], t["exit"]) for t in ref_trades]
    assert {k: getattr(s, k) for k in ref_s} == ref_s
    assert {t["outcome"] for t in trades} >= {"tp", "stop"}


@pytest.mark.gpu
def test_gpu_fp32_probe_is_plausible(eng):
    """The FFMA-only probe behind bench.py's roofline_fp32: a B200 sustains tens of TFLOP/s, never above nominal."""
    v = eng.measure_fp32_peak()
    assert 20.0 < v < 90.0, v


@pytest.mark.gpu
@pytest.mark.parametrize("hl", [False, True], ids=["close", "hl"])
def test_gpu_sweep_kernel_equals_per_combo_kernels(eng, hl, monkeypatch):
    """Parameter sweeps run 32 combos to a warp over one path (lean_sweep_kernel).  77 combos = two full warps of lanes
    + a 13-combo tail that stays on the per-combo kernel, a second family (other CONTACT_PROX) of 24 lanes, LR gate
    on (its mask depends on STOP_PADDING, so it differs per lane).  Against the per-combo kernels (QMMX_NO_SWEEP=1)
    on every aggregate, and against the C oracle on a few combos."""
    rows, price, solid = E.sort_levels(F.levels_30())
    grid = [abi.make_params(contact_prox=0.1, stop_padding=sp, tp_padding=tp, cooldown_s=cd, exit_on_hl=hl)
            for sp in (0.10, 0.2, 0.35, 0.5, 0.85, 0.0, 1.5) for tp in (0.15, 0.25, 0.5) for cd in (0, 60, 300, 7200)][:77]
    grid += [abi.make_params(contact_prox=0.05, q_min_prob=0.5, stop_padding=0.1 + 0.03 * i, cooldown_s=120, exit_on_hl=hl)
             for i in range(24)]
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    g = abi.make_gbm(sub_steps=4 if hl else 1)
    P, T = 3001, 395                       # ragged: not a multiple of the warp count, last chunk of 11 steps
    monkeypatch.setenv("QMMX_NO_SWEEP", "1")
    plain = E.MCEngine(0)
    monkeypatch.delenv("QMMX_NO_SWEEP")
    try:
        out = []
        for e in (eng, plain):
            configure(e, price, solid, grid, lr)
            e.gen_gbm(11, 0, P, T, g)
            n0 = e.launch_count()
            out.append(e.run_scan())
            launches = e.launch_count() - n0
            if e is eng:
                assert launches < len(grid) // 2          # 2 + 1 sweep launches + the 13-combo tail (+ the reduction)
        for a, b in zip(*out):
            assert_agg_equal(a.agg, b.agg)
        c, h, l = eng.download_bars(P, T, hl=hl)
        # the host pipeline (chunked H2D + scan per chunk) sweeps each chunk and adds up to the same aggregates
        for a, b in zip(eng.run_scan_host(c, h if hl else None, l if hl else None, chunk_paths=1024), out[0]):
            assert_agg_equal(a.agg, b.agg, sums_rel=1e-12)
        for ci in (0, 31, 32, 76, 77, 100):
            o = CO.run_paths(c, price, solid, grid[ci], lr=lr, high=h if hl else None, low=l if hl else None, n_threads=4)
            assert_agg_equal(out[0][ci].agg, o["agg"])
        assert len({r.n_trades for r in out[0]}) > 20        # the combos really differ
        for P2, T2 in ((1, 1), (5, 33), (40, 2)):            # degenerate shapes: one step, a chunk boundary, idle warps
            got = []
            for e in (eng, plain):
                e.gen_gbm(12, 0, P2, T2, g)
                got.append(e.run_scan())
            for a, b in zip(*got):
                assert_agg_equal(a.agg, b.agg)
        if not hl:   # constant prices (no direction ever) and a first price change beyond the first 32-step chunk
            close = F.host_gbm_closes(64, 100, seed=5)
            close[:, :8] = np.float32(648.76)
            close[:45, 8:16] = close[45, 8:16]
            got = []
            for e in (eng, plain):
                e.upload_bars(close)
                got.append(e.run_scan())
            for a, b in zip(*got):
                assert_agg_equal(a.agg, b.agg)
            assert got[0][0].entry_hist["DIR_UNKNOWN"] >= 8 * 100 + 8 * 45
    finally:
        plain.close()


@pytest.mark.gpu
def test_gpu_sweep_kernel_full_c4_size(eng, monkeypatch):
    """BASELINE config 4 at full size (256 combos x 1M resident paths x 390): the sweep kernel's aggregates equal the
    per-combo kernels' for every combo, and the counters of each combo partition its path-steps."""
    rows, price, solid = E.sort_levels(F.levels_30())
    grid = [abi.make_params(contact_prox=0.10, stop_padding=0.10 + 0.05 * i, tp_padding=tp, cooldown_s=cd)
            for i in range(16) for tp in (0.15, 0.25, 0.35, 0.50) for cd in (0, 60, 300, 900)]
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    P, T = 1_000_000, 390
    monkeypatch.setenv("QMMX_NO_SWEEP", "1")
    plain = E.MCEngine(0)
    monkeypatch.delenv("QMMX_NO_SWEEP")
    try:
        out = []
        for e in (eng, plain):
            configure(e, price, solid, grid, lr)
            e.gen_gbm(F.HOST_SEED, 0, P, T, abi.make_gbm())
            out.append(e.run_scan())
        for a, b in zip(*out):
            assert_agg_equal(a.agg, b.agg, sums_rel=1e-12)
            assert a.n_paths == P and a.n_steps == P * T == sum(a.entry_hist.values())
            assert a.n_trades == a.exit_hist["STOP"] + a.exit_hist["TARGET"] and a.entry_hist["OK"] == a.n_trades + a.n_open
    finally:
        plain.close()


@pytest.mark.gpu
@pytest.mark.parametrize("hl", [False, True], ids=["close", "hl"])
def test_gpu_online_policy_retrain(eng, hl):
    """SURVEY 8f-4 on the device: every path's OnlinePolicy go heads learn from its own closed trades (policy_learn,
    two SGD steps per trade as the reference's two labelled rows give).  Per-tick codes, trade records and all integer
    aggregates equal the C oracle's; the learnt float64 weights agree to 1e-12 relative (CUDA exp vs libm exp differ by
    at most one ulp per update).  The learning must matter: ONLINE_POLICY skips appear and a frozen policy decides
    differently."""
    rows, price, solid = E.sort_levels(F.levels_30())
    pol = np.array([[[0.55, -0.3, 0.0, 0.0, 0.0, 0.05, -0.1], [0.5, -0.2, 0.0, 0.0, 0.0, 0.05, 0.05],
                     [-0.6, 0.2, 0.0, 0.0, 0.1, -0.1, 0.1]]])
    kw = dict(contact_prox=0.1, cooldown_s=60, disable_policy_gate=False, exit_on_hl=hl)
    P, T = 2053, 3 * 390
    g = abi.make_gbm(sub_steps=4 if hl else 1)
    params = abi.make_params(online_retrain=True, **kw)
    configure(eng, price, solid, params, pol=pol)
    eng.gen_gbm(77, 0, P, T, g)
    c, h, l = eng.download_bars(P, T, hl=hl)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=(T, P))
    w = eng.download_policy_weights(P)
    o = CO.run_paths(c, price, solid, params, policy=pol, high=h, low=l, want_codes=True, max_trades=64, n_threads=8)
    assert np.array_equal(codes, o["codes"])
    assert_agg_equal(res[0].agg, o["agg"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert np.allclose(w, o["policy_weights"], rtol=1e-12, atol=0)
    assert res[0].entry_hist["ONLINE_POLICY"] > 0 and np.abs(w - pol[0, :2].reshape(1, 14)).max() > 0.1
    # aggregate-only run (no parity outputs) gives the same aggregates and weights
    agg2 = eng.run_scan()[0].agg
    assert list(agg2.entry_hist) == list(o["agg"].entry_hist) and list(agg2.exit_hist) == list(o["agg"].exit_hist)
    assert agg2.n_wins == o["agg"].n_wins and list(agg2.pnl_hist) == list(o["agg"].pnl_hist)
    assert np.array_equal(eng.download_policy_weights(P), w)
    frozen = abi.make_params(online_retrain=False, **kw)
    configure(eng, price, solid, frozen, pol=pol)
    assert eng.run_scan()[0].n_trades != res[0].n_trades
    # what the flag cannot do fails loudly
    configure(eng, price, solid, [params, params], pol=pol)
    with pytest.raises(E.QmmxError):
        eng.run_scan()
    # the fused generate+scan kernels carry the per-path heads too (SURVEY 8f-4 over Philox GBM and bootstrap paths): the
    # same paths scanned from HBM give the same aggregates and bit-identical learnt weights
    configure(eng, price, solid, params, pol=pol)
    eng.gen_gbm(77, 0, P, T, g)
    scan = eng.run_scan()[0].agg
    w_scan = eng.download_policy_weights(P)
    fused = eng.run_fused_gbm(77, 0, P, T, g)[0].agg
    assert_agg_equal(fused, scan)
    assert np.array_equal(eng.download_policy_weights(P), w_scan) and fused.entry_hist[abi.ENTRY_CODES.index("ONLINE_POLICY")] > 0
    window = F.synthetic_window(40)
    if hl:
        eng.gen_bootstrap(5, 1000, P, T, window, 5)
        scan = eng.run_scan()[0].agg
        w_scan = eng.download_policy_weights(P)
        fused = eng.run_fused_bootstrap(5, 1000, P, T, window, 5)[0].agg
        assert_agg_equal(fused, scan)
        assert np.array_equal(eng.download_policy_weights(P), w_scan)


def _levels_48(seed=9):
    """48 levels (the 64-level kernel variants: 64-bit latch, 104 KB of tables), clustered so that windows overlap."""
    rng = np.random.default_rng(seed)
    price = np.round(np.concatenate([rng.uniform(641.0, 657.0, 40), 648.9 + rng.normal(0, 0.08, 8)]), 2)
    return [dict(color=("blue", "orange", "black", "teal")[i % 4], type="solid" if i % 3 else "dashed", index=i // 4,
                 price=float(p)) for i, p in enumerate(price)]


@pytest.mark.gpu
@pytest.mark.parametrize("hl", [False, True], ids=["close", "hl"])
def test_gpu_wide_level_sets(eng, hl, monkeypatch):
    """More than 32 levels selects the WIDE kernel variants everywhere: scan (parity outputs), fused GBM, sweep,
    online retrain -- each against the C oracle or against its generate-then-scan twin."""
    rows, price, solid = E.sort_levels(_levels_48())
    assert len(price) == 48
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    g = abi.make_gbm(sub_steps=4 if hl else 1)
    P, T = 4099, 390
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=hl)
    configure(eng, price, solid, params, lr)
    eng.gen_gbm(31, 5, P, T, g)
    c, h, l = eng.download_bars(P, T, hl=hl)
    res, tr, cnt, codes = eng.run_scan(trades=48, codes=True, paths_steps=(T, P))
    o = CO.run_paths(c, price, solid, params, lr=lr, high=h, low=l, want_codes=True, max_trades=48, n_threads=8)
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])
    assert_agg_equal(eng.run_fused_gbm(31, 5, P, T, g)[0].agg, eng.run_scan()[0].agg, sums_rel=1e-12)
    # sweep (lanes = combos) on 64-level tables
    grid = [abi.make_params(contact_prox=0.1, stop_padding=0.1 + 0.05 * i, tp_padding=tp, cooldown_s=cd, exit_on_hl=hl)
            for i in range(6) for tp in (0.15, 0.4) for cd in (0, 300)]
    monkeypatch.setenv("QMMX_NO_SWEEP", "1")
    plain = E.MCEngine(0)
    monkeypatch.delenv("QMMX_NO_SWEEP")
    try:
        out = []
        for e in (eng, plain):
            configure(e, price, solid, grid, lr)
            e.gen_gbm(31, 5, P, T, g)
            out.append(e.run_scan())
        for a, b in zip(*out):
            assert_agg_equal(a.agg, b.agg, sums_rel=1e-12)
        assert_agg_equal(out[0][7].agg, CO.run_paths(c, price, solid, grid[7], lr=lr, high=h, low=l, n_threads=8)["agg"])
    finally:
        plain.close()
    # online retrain on 64-level tables, three days with per-day LR tables (MULTIDAY + LEARN variant)
    pol = np.array([[[0.55, -0.3, 0.0, 0.0, 0.0, 0.05, -0.1], [0.5, -0.2, 0.0, 0.0, 0.0, 0.05, 0.05],
                     [-0.6, 0.2, 0.0, 0.0, 0.1, -0.1, 0.1]]])
    lr3 = np.array([[0.2, -0.5, -0.45, 0.35, 1.0], [0.1, -0.4, -0.3, 0.2, 0.9], [0.3, -0.6, -0.5, 0.4, 1.1]])
    pr = abi.make_params(contact_prox=0.1, cooldown_s=60, disable_policy_gate=False, online_retrain=True, exit_on_hl=hl)
    configure(eng, price, solid, pr, lr3, pol)
    P2, T2 = 1031, 3 * 390
    eng.gen_gbm(32, 0, P2, T2, g)
    c, h, l = eng.download_bars(P2, T2, hl=hl)
    res, tr, cnt, codes = eng.run_scan(trades=64, codes=True, paths_steps=(T2, P2))
    o = CO.run_paths(c, price, solid, pr, lr=lr3, policy=pol, high=h, low=l, want_codes=True, max_trades=64, n_threads=8)
    assert np.array_equal(codes, o["codes"])
    assert_trades_equal(tr, cnt, o["trades"], o["counts"])
    assert_agg_equal(res[0].agg, o["agg"])
    assert np.allclose(eng.download_policy_weights(P2), o["policy_weights"], rtol=1e-12, atol=0)


@pytest.mark.gpu
def test_gpu_wide_level_sets_float64_gates_and_bootstrap(eng):
    """The remaining 64-level variants: OnlinePolicy gate (frozen weights), blend gate, soft veto + regime gate on a volume
    plane (the float64 SLOW kernels), and the fused block bootstrap, all against the oracle / the generate-then-scan twin."""
    rows, price, solid = E.sort_levels(_levels_48())
    rng = np.random.default_rng(12)
    lr, pol = _lr_tables(rng, 2)
    P, T = 2051, 2 * 390
    for kw in (dict(disable_policy_gate=False), dict(use_blend=True), dict(enable_veto=True, veto_prox=0.08, exit_on_hl=True)):
        hl = bool(kw.get("exit_on_hl"))
        params = abi.make_params(contact_prox=0.1, cooldown_s=60, **kw)
        configure(eng, price, solid, params, lr, pol)
        eng.gen_gbm(41, 0, P, T, abi.make_gbm(sub_steps=4 if hl else 1))
        c, h, l = eng.download_bars(P, T, hl=hl)
        vol = None
        if kw.get("enable_veto"):
            vol = (1e5 * np.exp(0.5 * np.cumsum(rng.normal(0, 0.35, size=c.shape), axis=0))).astype(np.float32)
            eng.upload_volume(vol)
        res, tr, cnt, codes = eng.run_scan(trades=48, codes=True, paths_steps=(T, P))
        o = CO.run_paths(c, price, solid, params, lr=lr, policy=pol, high=h, low=l, volume=vol, want_codes=True,
                         max_trades=48, n_threads=8)
        assert np.array_equal(codes, o["codes"]), kw
        assert_trades_equal(tr, cnt, o["trades"], o["counts"])
        assert_agg_equal(res[0].agg, o["agg"])
    g = abi.make_gbm(sub_steps=4)
    wc, wh, wl = CO.gen_gbm(5, 0, 1, 40, g, ohlc=True)
    wo = np.concatenate([[g.s0], wc[:-1, 0]]).astype(np.float32)
    window = np.stack([wo, wh[:, 0], wl[:, 0], wc[:, 0]], axis=1)
    configure(eng, price, solid, abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=True), lr)
    fused = eng.run_fused_bootstrap(9, 7, 20_011, T, window, 5)[0].agg
    eng.gen_bootstrap(9, 7, 20_011, T, window, 5)
    assert_agg_equal(fused, eng.run_scan()[0].agg, sums_rel=1e-12)
    assert fused.n_trades > 0


@pytest.mark.gpu
def test_gpu_wide_level_sets_first_generation_kernels(monkeypatch):
    """The first-generation kernels (the TINY_PRICES path, QMMX_SCAN_V1=1) on 48 levels, scan and fused, against the oracle."""
    rows, price, solid = E.sort_levels(_levels_48())
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    monkeypatch.setenv("QMMX_SCAN_V1", "1")
    v1 = E.MCEngine(0)
    monkeypatch.delenv("QMMX_SCAN_V1")
    try:
        for hl in (False, True):
            g = abi.make_gbm(sub_steps=4 if hl else 1)
            params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=hl)
            configure(v1, price, solid, params, lr)
            P, T = 3001, 390
            v1.gen_gbm(51, 0, P, T, g)
            c, h, l = v1.download_bars(P, T, hl=hl)
            res, tr, cnt, codes = v1.run_scan(trades=48, codes=True, paths_steps=(T, P))
            o = CO.run_paths(c, price, solid, params, lr=lr, high=h, low=l, want_codes=True, max_trades=48, n_threads=8)
            assert np.array_equal(codes, o["codes"])
            assert_trades_equal(tr, cnt, o["trades"], o["counts"])
            assert_agg_equal(res[0].agg, o["agg"])
            assert_agg_equal(v1.run_fused_gbm(51, 0, P, T, g)[0].agg, o["agg"], sums_rel=1e-9)
    finally:
        v1.close()


def _sigmoid_threshold(target):
    """smallest float64 z with 1 / (1 + exp(-z)) >= target, as the library derives it (M:20-26 with libm exp)"""
    import math
    sig = lambda x: 1.0 / (1.0 + math.exp(-x))
    lo, hi = -50.0, 50.0
    while True:
        mid = lo + (hi - lo) / 2
        if mid <= lo or mid >= hi:
            break
        if sig(mid) >= target:
            hi = mid
        else:
            lo = mid
    while sig(math.nextafter(hi, -math.inf)) >= target:
        hi = math.nextafter(hi, -math.inf)
    return hi


@pytest.mark.gpu
@pytest.mark.parametrize("bump", [-1, 0, 1], ids=["below", "at", "above"])
def test_gpu_policy_gate_at_the_threshold(eng, bump):
    """The OnlinePolicy gate (M:3083-3085) decides on a plain float64 sum when that sum is clear of the threshold by more
    than its rounding band and on CPython's compensated sum inside the band.  Heads whose only non-zero weight is the bias
    put EVERY evaluation inside the band: bias = the go threshold (0.60) or one ulp either side, skip bias = the skip
    threshold (0.55) or one ulp either side; the oracle evaluates the sigmoids themselves."""
    import math
    z_go, z_sk = _sigmoid_threshold(0.60), _sigmoid_threshold(0.55)
    step = lambda z, n: z if n == 0 else math.nextafter(z, math.inf if n > 0 else -math.inf)
    rows, price, solid = E.sort_levels(F.levels_30())
    close = F.host_gbm_closes(3000, 390, seed=91)
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, disable_policy_gate=False)
    outcomes = []
    for go_n, sk_n in ((bump, -1), (1, bump)):
        pol = np.zeros((1, 3, 7))
        pol[0, 0, 0] = pol[0, 1, 0] = step(z_go, go_n)
        pol[0, 2, 0] = step(z_sk, sk_n)
        configure(eng, price, solid, params, None, pol)
        eng.upload_bars(close)
        a = eng.run_scan()[0]
        o = CO.run_paths(close, price, solid, params, policy=pol, n_threads=4)
        assert_agg_equal(a.agg, o["agg"])
        outcomes.append((a.n_trades > 0, a.entry_hist["ONLINE_POLICY"] > 0))
    # go head: passes at and above its threshold (>= 0.60); skip head: passes strictly below its threshold (< 0.55)
    assert outcomes[0] == ((bump >= 0), (bump < 0)) and outcomes[1] == ((bump < 0), (bump >= 0))


@pytest.mark.gpu
@pytest.mark.parametrize("hl", [False, True], ids=["close", "hl"])
def test_gpu_class_partition_at_capacity(eng, hl):
    """64 levels ~0.2 apart with CONTACT_PROX 0.2: every window overlaps its neighbours', the class partition of the price
    line (touched level + levels whose window holds the price) needs ~180 classes, the single-boundary records that are
    left (254 - classes) run out and the remaining boundary cells resolve through the zone search.  Parity outputs,
    aggregates and the fused kernel against the C oracle; 32 of the same levels on the tile-event kernel."""
    rng = np.random.default_rng(8)
    prices = 643.0 + np.round(np.cumsum(rng.uniform(0, 0.4, 64)), 2)
    levels = [dict(color=("blue", "orange", "black", "teal")[i % 4], type="solid" if i % 3 else "dashed", index=i // 4,
                   price=float(p)) for i, p in enumerate(prices)]
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    g = abi.make_gbm(sub_steps=4 if hl else 1)
    for lv, overtouch in ((levels, 4), (levels[16:48], 6)):
        rows, price, solid = E.sort_levels(lv)
        ncls, nruns, nrec = E.debug_class_partition(price, solid, 0.2, 0.6)
        assert ncls > 2.5 * len(price) and ncls + nrec <= 254
        if len(price) == 64:
            assert ncls + nrec == 254                      # capacity reached: some boundary cells use the zone search
        P, T = 3001, 390
        params = abi.make_params(contact_prox=0.2, cooldown_s=60, exit_on_hl=hl, overtouch_limit=overtouch)
        configure(eng, price, solid, params, lr)
        eng.gen_gbm(77, 3, P, T, g)
        c, h, l = eng.download_bars(P, T, hl=hl)
        res, tr, cnt, codes = eng.run_scan(trades=200, codes=True, paths_steps=(T, P))
        o = CO.run_paths(c, price, solid, params, lr=lr, high=h, low=l, want_codes=True, max_trades=200, n_threads=8)
        assert np.array_equal(codes, o["codes"])
        assert_trades_equal(tr, cnt, o["trades"], o["counts"])
        assert_agg_equal(res[0].agg, o["agg"])
        assert_agg_equal(eng.run_scan()[0].agg, o["agg"])                                  # aggregate-only kernels (tile for 32 levels)
        assert_agg_equal(eng.run_fused_gbm(77, 3, P, T, g)[0].agg, o["agg"], sums_rel=1e-9)
        assert o["agg"].n_trades > 100 and o["agg"].n_trade_overflow == 0


# ---- tile-event kernels (scan_tile.cuh): the production close-only scan and fused GBM kernels -----------------------
TILE_CASES = [
    # (paths, steps, params, days, mode, cents, sigma)
    (4999, 390, dict(contact_prox=0.1, cooldown_s=120), 1, "and", False, 0.15),
    (1, 1, dict(contact_prox=0.1, cooldown_s=120), 1, "and", False, 0.15),
    (33, 7, dict(contact_prox=0.1, cooldown_s=8), 1, "none", False, 0.15),
    (257, 391, dict(contact_prox=0.1, cooldown_s=0), 1, "and", True, 0.15),
    (2050, 9, dict(contact_prox=0.2, cooldown_s=300, overtouch_limit=2), 1, "none", True, 0.15),
    (3001, 64, dict(contact_prox=0.05, cooldown_s=61, stop_padding=0.05, tp_padding=0.05), 1, "and", True, 0.15),
    (3000, 1170, dict(contact_prox=0.1, cooldown_s=120), 3, "and", False, 0.10),
    (3000, 1170, dict(contact_prox=0.1, cooldown_s=120, disable_policy_gate=False), 3, "and+policy", False, 0.10),
    (2000, 780, dict(contact_prox=0.1, cooldown_s=120, use_blend=True, w_rules=0.6, w_ml=0.4), 2, "blend", False, 0.10),
    (1500, 400, dict(contact_prox=0.1, cooldown_s=1800, stop_padding=2.5, tp_padding=2.0), 1, "none", False, 0.15),
    (1000, 390, dict(contact_prox=0.1, cooldown_s=120), 1, "and", False, 0.004),      # long flat stretches in cents
]


@pytest.mark.parametrize("ci", range(len(TILE_CASES)))
def test_gpu_tile_kernel_equals_oracle_and_lean(eng, ci, monkeypatch):
    """The tile-event kernel (per-lane cursors in a 32-row ring, masks + rounds) against the C oracle and against the
    lockstep lean kernel (QMMX_NO_TILE=1) on ragged path/step counts, flat and cent-rounded closes (runs of equal
    prices: direction carried across chunks), zero and day-long cooldowns, trades that outlast many chunks, per-day
    logistic masks, the blend and OnlinePolicy gates."""
    paths, steps, kw, days, mode, cents, sigma = TILE_CASES[ci]
    rng = np.random.default_rng(100 + ci)
    lr, pol = _lr_tables(rng, days)
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(**kw)
    close = F.host_gbm_closes(paths, steps, seed=500 + ci, sigma=sigma, round_cents=cents)
    if ci == 10:
        close[5:40, ::3] = close[5, ::3]              # 35 equal closes in a row, inside a contact window for some
        close[100:140, 1::3] = np.float32(648.76)
    use_lr = lr if mode in ("and", "and+policy", "blend") else None
    use_pol = pol if "policy" in mode else None
    o = CO.run_paths(close, price, solid, params, lr=use_lr, policy=use_pol, n_threads=4)
    monkeypatch.setenv("QMMX_NO_TILE", "1")
    lean = E.MCEngine(0)
    monkeypatch.delenv("QMMX_NO_TILE")
    try:
        for e in (eng, lean):
            configure(e, price, solid, params, use_lr, use_pol)
            e.upload_bars(close)
            assert_agg_equal(e.run_scan()[0].agg, o["agg"])
    finally:
        lean.close()


@pytest.mark.parametrize("cents", [False, True], ids=["raw", "cents"])
def test_gpu_tile_fused_gbm_equals_generate_then_scan(eng, cents):
    """Fused Philox GBM + tile scan == gen_gbm_kernel bars scanned by the tile kernel == the same bars through the
    C oracle, for a ragged path count and a step count that is not a multiple of the chunk."""
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120)
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    g = abi.make_gbm(round_cents=cents)
    configure(eng, price, solid, params, lr)
    for P, T in ((70_001, 390), (4097, 397), (999, 5)):
        fused = eng.run_fused_gbm(11, 1234, P, T, g)[0].agg
        eng.gen_gbm(11, 1234, P, T, g)
        scan = eng.run_scan()[0].agg
        close = eng.download_bars(P, T)[0]
        o = CO.run_paths(close, price, solid, params, lr=lr, n_threads=4)
        assert_agg_equal(scan, o["agg"])
        assert_agg_equal(fused, o["agg"])


def test_gpu_tile_kernel_volume_and_regime_gates(eng):
    """SLOW tile variant with the per-bar volume plane: soft veto (M:1678-1705) and the breakout regime gate (M:1652-1666)
    read global planes by (step, path) from inside the rounds."""
    rng = np.random.default_rng(21)
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, enable_veto=True)
    P, T = 3000, 390
    close = F.host_gbm_closes(P, T, seed=77, sigma=0.05)
    vol = rng.lognormal(10.0, 0.9, size=(T, P)).astype(np.float32)
    configure(eng, price, solid, params)
    eng.upload_bars(close)
    eng.upload_volume(vol)
    a = eng.run_scan()[0]
    o = CO.run_paths(close, price, solid, params, volume=vol, n_threads=4)
    assert_agg_equal(a.agg, o["agg"])
    assert a.entry_hist["VETO_CONTRA_VOL_LONG"] + a.entry_hist["VETO_CONTRA_VOL_SHORT"] > 0
    eng.upload_volume(None)


def test_gpu_aggregate_reads_need_a_matching_run(eng):
    """Round-1 advisor finding: fetch / agg_device / merge size their access by the CURRENT combo count, so they must
    refuse to read after the configuration changed (more combos than the buffer of the last run holds) or when only a
    generator / scorer has run; a new run makes them valid again."""
    rows, price, solid = E.sort_levels(F.levels_30())
    fresh = E.MCEngine(0)
    try:
        configure(fresh, price, solid, abi.make_params(contact_prox=0.1))
        fresh.gen_gbm(1, 0, 1000, 50, abi.make_gbm())                   # a generator is not a scan: nothing to fetch
        with pytest.raises(E.QmmxError):
            fresh.fetch_agg()
        assert fresh.run_scan()[0].n_paths == 1000
        fresh.set_params([abi.make_params(contact_prox=0.1, cooldown_s=c) for c in (8, 60, 120, 300)])
        for call in (fresh.fetch_agg, fresh.agg_device, lambda: fresh.merge_flat(0, 1), fresh.allreduce):
            with pytest.raises(E.QmmxError):
                call()
        res = fresh.run_scan()
        assert len(res) == 4 and all(r.n_paths == 1000 for r in res)
        assert len(fresh.fetch_agg()) == 4
    finally:
        fresh.close()
