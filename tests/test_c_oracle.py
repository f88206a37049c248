"""oracle/qmmx_oracle.c (C restatement) against oracle/headless.py and the reference golden vectors."""
import random

import numpy as np
import pytest

from oracle import c_oracle as CO
from oracle import headless as H
from qmmx_monolithic_monte_carlo_b200 import abi
from qmmx_monolithic_monte_carlo_b200 import engine as E
from qmmx_monolithic_monte_carlo_b200 import fixtures as F
from tests.golden_util import load_engine_cases, oracle_inputs

ENGINE_CASES = load_engine_cases()


def c_inputs(case):
    lv, P, lr, pol = oracle_inputs(case)
    params = abi.make_params(**case["params"])
    lr_a = np.array([list(m.coef) + [m.intercept] for m in lr]) if lr else None
    pol_a = np.array([[w["go_long"], w["go_short"], w["skip"]] for w in pol]) if pol else None
    return lv, params, lr_a, pol_a


@pytest.mark.parametrize("case", ENGINE_CASES, ids=[c["name"] for c in ENGINE_CASES])
def test_c_oracle_matches_reference_golden(case):
    lv, params, lr_a, pol_a = c_inputs(case)
    r = CO.run_paths(case["close"], [l.price for l in lv], [l.solid for l in lv], params,
                     lr=lr_a, policy=pol_a, want_codes=True, max_trades=64, n_threads=2)
    assert np.array_equal(r["codes"], case["codes"])
    for j in range(case["paths"]):
        gt = case["trades"][j]
        assert r["counts"][j] == len(gt)
        for g, t in zip(gt, r["trades"][j]):
            assert (g["entry_bar"], g["exit_bar"], g["side"], g["level"], g["reason"]) == \
                   (t["entry_bar"], t["exit_bar"], t["side"], t["level_id"], t["reason"])
            assert np.float32(g["pnl"]) == t["pnl"] and np.float32(g["entry"]) == t["entry"]
    a = r["agg"]
    assert a.n_paths == case["paths"] and a.n_steps == case["paths"] * case["steps"]
    assert list(a.entry_hist) == np.bincount(case["codes"].ravel(), minlength=20).tolist()
    closed = [t for tr in case["trades"] for t in tr if t["reason"] != 3]
    assert a.n_trades == len(closed) and a.n_wins == sum(1 for t in closed if t["pnl"] > 0)
    assert a.sum_pnl == pytest.approx(sum(t["pnl"] for t in closed), rel=1e-12, abs=1e-12)


def test_c_oracle_hl_exit_mode_matches_headless():
    rng = np.random.default_rng(3)
    g = abi.make_gbm(sub_steps=4)
    c, h, l = CO.gen_gbm(77, 0, 64, 390, g, ohlc=True)
    assert np.all(h >= c) and np.all(l <= c) and np.all(h >= l)
    lv = H.sort_levels(F.levels_30())
    P = H.Params(contact_prox=0.1, cooldown_s=120, exit_on_hl=True)
    params = abi.make_params(contact_prox=0.1, cooldown_s=120, exit_on_hl=True)
    r = CO.run_paths(c, [x.price for x in lv], [x.solid for x in lv], params, high=h, low=l,
                     want_codes=True, max_trades=64)
    for j in range(0, 64, 7):
        o = H.run_path(c[:, j], lv, P, high=h[:, j], low=l[:, j])
        assert o.codes == r["codes"][:, j].tolist()
        assert len(o.trades) == r["counts"][j]
        for a, b in zip(o.trades, r["trades"][j]):
            assert (a.entry_bar, a.exit_bar, a.side, a.level, a.reason) == \
                   (b["entry_bar"], b["exit_bar"], b["side"], b["level_id"], b["reason"])
            assert np.float32(a.pnl) == b["pnl"]


def test_dot_py312_is_builtin_sum():
    rnd = random.Random(4)
    for _ in range(2000):
        w = [rnd.gauss(0, 1) * 10 ** rnd.randint(-8, 8) for _ in range(7)]
        x = [rnd.gauss(0, 1) for _ in range(7)]
        expect = sum(a * b for a, b in zip(w, x))   # CPython >= 3.12: Neumaier
        assert H.dot_py312(w, x) == expect
        wa, xa = np.array(w), np.array(x)
        assert CO.lib().oq_dot_py312(wa.ctypes.data, xa.ctypes.data, 7) == expect


def test_confidence_c_equals_python():
    rnd = random.Random(5)
    for _ in range(5000):
        d, s, tc, prox = rnd.random() * 0.12, rnd.random() < 0.5, rnd.randint(1, 5), rnd.choice([0.05, 0.1, 0.2])
        assert CO.lib().oq_confidence(d, int(s), tc, prox) == H.compute_confidence(d, s, tc, prox)


# Random123 known-answer vectors for philox4x32-10 (kat_vectors in the Random123 distribution)
PHILOX_KAT = [
    ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
     [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
]


@pytest.mark.parametrize("ctr,key,out", PHILOX_KAT)
def test_philox_known_answers(ctr, key, out):
    assert CO.philox(ctr, key).tolist() == out


def test_gbm_generator_statistics_and_sharding():
    g = abi.make_gbm()
    c, _, _ = CO.gen_gbm(20250829, 0, 4096, 390, g)
    lr = np.diff(np.log(np.vstack([np.full((1, 4096), g.s0, dtype=np.float64), c.astype(np.float64)])), axis=0)
    sd = 0.15 * np.sqrt(1 / (252 * 390))
    assert abs(lr.std() / sd - 1) < 0.01 and abs(lr.mean()) < 4 * sd / np.sqrt(lr.size)
    # path ids are global: a shard starting at offset 1000 reproduces columns 1000.. of the full run
    c2, _, _ = CO.gen_gbm(20250829, 1000, 96, 390, g)
    assert np.array_equal(c2, c[:, 1000:1096])


def test_bootstrap_generator_uses_window_returns():
    g = abi.make_gbm(sub_steps=4)
    wc, wh, wl = CO.gen_gbm(5, 0, 1, 40, g, ohlc=True)
    wo = np.concatenate([[g.s0], wc[:-1, 0]]).astype(np.float32)
    window = np.stack([wo, wh[:, 0], wl[:, 0], wc[:, 0]], axis=1)
    c, h, l = CO.gen_bootstrap(9, 0, 32, 2 * 390, window, 5)
    rc = (window[1:, 3] / window[:-1, 3]).astype(np.float32)
    prev = np.vstack([np.full((1, 32), window[-1, 3], dtype=np.float32), c[:-1]])
    ratio = c / prev
    # every step's close ratio is one of the window's 39 ratios (up to fp32 rounding of the product)
    d = np.abs(ratio[..., None] - rc[None, None, :]).min(axis=-1)
    assert d.max() < 2e-7
    assert np.all(h >= l)


VETO_LONG, VETO_SHORT = abi.ENTRY_CODES.index("VETO_CONTRA_VOL_LONG"), abi.ENTRY_CODES.index("VETO_CONTRA_VOL_SHORT")


def _volume_for(close, rng):
    """Per-bar volume with swings big enough that the 6-bar slope crosses +-0.25 regularly."""
    steps, paths = close.shape
    base = rng.uniform(5e4, 2e5, size=(1, paths))
    v = base * np.exp(np.cumsum(rng.normal(0, 0.35, size=(steps, paths)), axis=0) * 0.5)
    v[rng.random(size=v.shape) < 0.01] = 0.0          # a few empty bars (`v or 0.0`)
    return v.astype(np.float32)


def test_c_oracle_volume_veto_matches_headless():
    """Row a7 (M:1678-1705, M:1773-1811): with per-bar volume the soft veto fires; C restatement == headless.py."""
    rng = np.random.default_rng(11)
    close = F.host_gbm_closes(48, 390, seed=31)
    vol = _volume_for(close, rng)
    lv = H.sort_levels(F.levels_30())
    kw = dict(contact_prox=0.1, cooldown_s=60, enable_veto=True, veto_prox=0.08)
    P, params = H.Params(**kw), abi.make_params(**kw)
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    r = CO.run_paths(close, [x.price for x in lv], [x.solid for x in lv], params, volume=vol, lr=lr,
                     want_codes=True, max_trades=64, n_threads=2)
    n_veto = 0
    for j in range(48):
        o = H.run_path(close[:, j], lv, P, volume=[float(x) for x in vol[:, j]],
                       lr=[H.LRModel(coef=[0.2, -0.5, -0.45, 0.35], intercept=1.0)])
        assert o.codes == r["codes"][:, j].tolist(), j
        assert len(o.trades) == r["counts"][j]
        n_veto += sum(1 for c in o.codes if c in (VETO_LONG, VETO_SHORT))
    assert n_veto > 20          # the gate is exercised, both codes
    a = r["agg"]
    assert a.entry_hist[VETO_LONG] > 0 and a.entry_hist[VETO_SHORT] > 0
    # slope helper alone, every tick of one path
    for i in range(390):
        assert CO.lib().oq_volume_slope(vol.ctypes.data, 48, 3, i) == H.volume_slope([float(x) for x in vol[max(0, i - 239):i + 1, 3]])
    # without volume the veto is inert (the reference on its own data)
    r0 = CO.run_paths(close, [x.price for x in lv], [x.solid for x in lv], params, lr=lr, want_codes=True)
    assert r0["agg"].entry_hist[VETO_LONG] == 0 and r0["agg"].entry_hist[VETO_SHORT] == 0


def test_bootstrap_generator_carries_window_volume():
    g = abi.make_gbm(sub_steps=4)
    wc, wh, wl = CO.gen_gbm(5, 0, 1, 40, g, ohlc=True)
    wo = np.concatenate([[g.s0], wc[:-1, 0]]).astype(np.float32)
    window = np.stack([wo, wh[:, 0], wl[:, 0], wc[:, 0]], axis=1)
    wv = np.arange(1000, 1040, dtype=np.float32)
    c, h, l, v = CO.gen_bootstrap(9, 0, 16, 390, window, 5, window_volume=wv)
    c2, _, _ = CO.gen_bootstrap(9, 0, 16, 390, window, 5)
    assert np.array_equal(c, c2)
    assert set(np.unique(v)) <= set(wv[1:].tolist())
    idx = (v - 1000).astype(int)                       # the window bar each step was drawn from
    prev = np.vstack([np.full((1, 16), window[-1, 3], dtype=np.float32), c[:-1]])
    assert np.array_equal(c, prev * (window[idx, 3] / window[idx - 1, 3]).astype(np.float32))


def breakout_paths(n, seed):
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
    vol[np.random.default_rng(seed + 1).random(size=vol.shape) < 0.01] = 0.0
    return closes, vol


def test_c_oracle_regime_guard_matches_headless():
    """AccumulationBreakoutGuard regime per tick and the 904 gate, C restatement == headless.py."""
    closes, vol = breakout_paths(40, 5)
    reg = CO.regime_series(closes, vol)
    assert set(np.unique(reg)) == {0, 1, 2, 3}
    for j in range(40):
        assert reg[:, j].tolist() == H.regime_series(closes[:, j], vol[:, j]), j
    lv = H.sort_levels(F.levels_30())
    kw = dict(contact_prox=0.1, cooldown_s=60, enable_veto=True)
    r = CO.run_paths(closes, [x.price for x in lv], [x.solid for x in lv], abi.make_params(**kw), volume=vol,
                     want_codes=True, max_trades=64, n_threads=2)
    for j in range(40):
        o = H.run_path(closes[:, j], lv, H.Params(**kw), volume=[float(x) for x in vol[:, j]])
        assert o.codes == r["codes"][:, j].tolist(), j
    assert r["agg"].entry_hist[abi.ENTRY_CODES.index("ACC_BREAKOUT_GATE")] >= 5


def test_c_oracle_online_policy_retrain_matches_headless():
    """the C restatement of the OnlinePolicy learning path against oracle/headless.py (itself pinned against
    the reference in test_oracle_vs_reference.py): per-tick codes and every learnt float64 weight identical."""
    closes = F.host_gbm_closes(6, 3 * 390, seed=404)
    pol = [{"go_long": [0.55, -0.3, 0.0, 0.0, 0.0, 0.05, -0.1], "go_short": [0.5, -0.2, 0.0, 0.0, 0.0, 0.05, 0.05],
            "skip": [-0.6, 0.2, 0.0, 0.0, 0.1, -0.1, 0.1]}]
    pa = np.array([[pol[0]["go_long"], pol[0]["go_short"], pol[0]["skip"]]])
    rows, price, solid = E.sort_levels(F.levels_30())
    kw = dict(contact_prox=0.1, cooldown_s=60, disable_policy_gate=False, online_retrain=True)
    o = CO.run_paths(closes, price, solid, abi.make_params(**kw), policy=pa, want_codes=True, max_trades=64)
    lv = H.sort_levels(F.levels_30())
    n_pol = 0
    for j in range(closes.shape[1]):
        h = H.run_path(closes[:, j], lv, H.Params(**kw), policy=pol)
        assert list(o["codes"][:, j]) == h.codes
        assert (h.policy_weights["go_long"] + h.policy_weights["go_short"]) == list(o["policy_weights"][j])
        n_pol += h.codes.count(H.ENTRY_CODES.index("ONLINE_POLICY"))
    assert n_pol > 0


@pytest.mark.parametrize("lr_on,cents,cool", [(True, False, 120), (False, True, 8), (True, True, 0)])
def test_numpy_batch_oracle_equals_c_oracle(lr_on, cents, cool):
    """oracle/numpy_batch.py (the NumPy batch headless reference of SURVEY 8d, timed by bench.py as a CPU baseline)
    against the C oracle: every integer counter and both reason histograms equal, PnL sums to 1e-12."""
    from oracle import numpy_batch as NB
    lv = H.sort_levels(F.levels_30())
    kw = dict(contact_prox=0.1, cooldown_s=cool)
    close = F.host_gbm_closes(3000, 390, seed=41 + cool, round_cents=cents)
    lr = H.LRModel([0.2, -0.5, -0.45, 0.35], 1.0) if lr_on else None
    lr_a = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]]) if lr_on else None
    o = CO.run_paths(close, [x.price for x in lv], [x.solid for x in lv], abi.make_params(**kw), lr=lr_a, n_threads=4)["agg"]
    b = NB.run_batch(close, lv, H.Params(**kw), lr)
    assert list(o.entry_hist) == b["entry_hist"].tolist() and list(o.exit_hist) == b["exit_hist"].tolist()
    assert (o.n_trades, o.n_wins, o.n_open) == (b["n_trades"], b["n_wins"], b["n_open"])
    assert b["sum_pnl"] == pytest.approx(o.sum_pnl, rel=1e-12, abs=1e-9)
    assert b["sum_pnl2"] == pytest.approx(o.sum_pnl2, rel=1e-12, abs=1e-9)
    assert o.n_trades > 1000 and (not lr_on or o.entry_hist[abi.ENTRY_CODES.index("ML_CONF_LOW")] > 0)


def test_numpy_batch_oracle_matches_reference_golden():
    from oracle import numpy_batch as NB
    for case in ENGINE_CASES:
        lv, P, lr, pol = oracle_inputs(case)
        if P.use_blend or not P.disable_policy_gate or (lr and len(lr) > 1):
            continue
        P.enable_veto = False             # inert without volume (the golden run had it on with volume-less bars)
        b = NB.run_batch(case["close"], lv, P, lr[0] if lr else None)
        assert b["entry_hist"].tolist() == np.bincount(case["codes"].ravel(), minlength=20).tolist(), case["name"]
        closed = [t for tr in case["trades"] for t in tr if t["reason"] != 3]
        assert b["n_trades"] == len(closed)


def test_generators_follow_the_host_numpy_laws():
    """The C restatement of the Philox generators (bit-equal to the device's, tests/test_gpu_parity.py) against
    independent numpy-Philox draws of the same laws, through the engine: the statistical tier of
    tests/test_gpu_statistical.py at CPU-suite size (two-sample z tests per reason code, chi-square on the PnL histogram)."""
    import tests.test_gpu_statistical as S
    lv = H.sort_levels(F.levels_30())
    price, solid = [x.price for x in lv], [x.solid for x in lv]
    # GBM: oracle Philox4x32-10 + Box-Muller (fp32) vs numpy float64 GBM
    params = abi.make_params(contact_prox=0.10, cooldown_s=120)
    c1 = F.host_gbm_closes(60_000, 390, seed=11)
    c2, _, _ = CO.gen_gbm(12, 0, 90_000, 390, abi.make_gbm())
    a1 = CO.run_paths(c1, price, solid, params, lr=S.LR, n_threads=8)["agg"]
    a2 = CO.run_paths(c2, price, solid, params, lr=S.LR, n_threads=8)["agg"]
    var = S.per_path_variances(CO.run_paths(np.ascontiguousarray(c1[:, :30_000]), price, solid, params, lr=S.LR, want_codes=True,
                                            max_trades=96, n_threads=8), 390, 30_000)
    S.compare(a2, 90_000, a1, 60_000, var)
    # block bootstrap: oracle generator vs numpy
    w = F.synthetic_window(40)
    p2 = abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True)
    c, h, l = S.numpy_block_bootstrap(w, 50_000, 390, 5, seed=5)
    oc, oh, ol = CO.gen_bootstrap(77, 0, 70_000, 390, w, 5)
    b1 = CO.run_paths(c, price, solid, p2, high=h, low=l, lr=S.LR, n_threads=8)["agg"]
    b2 = CO.run_paths(oc, price, solid, p2, high=oh, low=ol, lr=S.LR, n_threads=8)["agg"]
    var = S.per_path_variances(CO.run_paths(np.ascontiguousarray(c[:, :30_000]), price, solid, p2, lr=S.LR, want_codes=True, max_trades=128,
                                            high=np.ascontiguousarray(h[:, :30_000]), low=np.ascontiguousarray(l[:, :30_000]),
                                            n_threads=8), 390, 30_000)
    S.compare(b2, 70_000, b1, 50_000, var)
