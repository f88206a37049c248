"""`qmmx_run_last_bars_static` (include/qmmx_mc.h): the last-bars walk with its scratch in the library's file-scope
`static __device__` array instead of a cudaMalloc'ed arena.  No counterpart in the reference.
Collected last on purpose (the file name sorts after every other test file).

The check is bit-exact (integer outcomes and bar indices) against the existing entry point `qmmx_run_last_bars`
through the same C ABI, on two calls in a row that share the array: the second call is smaller and different, so
anything read from what the first call left behind would show.
"""

import numpy as np
import pytest

from qmmx_monolithic_monte_carlo_b200 import abi
from qmmx_monolithic_monte_carlo_b200 import engine as E
from qmmx_monolithic_monte_carlo_b200 import fixtures as F

pytestmark = pytest.mark.gpu

MAX_BARS, MAX_CAND = 12000, 16384          # QMMX_STATIC_MAX_BARS / QMMX_STATIC_MAX_CAND


def make_case(n_bars, n_cand, seed):
    """a synthetic high/low series and candidates whose stops and targets sit around the price at their bar"""
    rng = np.random.default_rng(seed)
    w = F.synthetic_window(n_bars, seed=seed)
    high, low, close = w[:, 1].astype(np.float64), w[:, 2].astype(np.float64), w[:, 3].astype(np.float64)
    cands = []
    for i in rng.integers(0, n_bars, size=n_cand):
        side = int(rng.integers(0, 2))
        a, b = rng.uniform(0.02, 1.5, size=2)
        stop, target = (close[i] - a, close[i] + b) if side == 0 else (close[i] + a, close[i] - b)
        cands.append(dict(i=int(i), side=side, stop=float(stop), target=float(target)))
    return high, low, cands


def test_gpu_last_bars_static_equals_the_allocating_entry_point_twice_in_a_row():
    with E.MCEngine(0) as eng:
        seen = set()
        for n_bars, n_cand, seed in ((MAX_BARS, MAX_CAND, 41), (777, 1001, 42), (5000, 1, 43)):
            high, low, cands = make_case(n_bars, n_cand, seed)
            oc0, xb0 = eng.run_last_bars(high, low, cands)             # existing entry point (cudaMalloc'ed arena)
            n0 = eng.launch_count()
            oc1, xb1 = eng.run_last_bars_static(high, low, cands)      # static scratch, first pass
            oc2, xb2 = eng.run_last_bars_static(high, low, cands)      # and again on the used array
            assert eng.launch_count() == n0 + 2                        # the kernel ran, twice
            for oc, xb in ((oc1, xb1), (oc2, xb2)):
                assert np.array_equal(oc, oc0) and np.array_equal(xb, xb0)
            seen |= set(oc0.tolist())
        assert seen == {0, 1, 2}                                       # tp, stop and open all occur


def test_gpu_last_bars_static_rejects_sizes_beyond_the_static_scratch():
    with E.MCEngine(0) as eng:
        high, low, cands = make_case(64, 4, 7)
        big = cands * (MAX_CAND // 4) + cands[:1]                      # MAX_CAND + 1 candidates
        with pytest.raises(E.QmmxError) as ei:
            eng.run_last_bars_static(high, low, big)
        assert ei.value.status == abi.E_UNSUPPORTED and "QMMX_STATIC_MAX_CAND 16384" in str(ei.value)
        long_h = np.full(MAX_BARS + 1, 101.0); long_l = np.full(MAX_BARS + 1, 99.0)
        with pytest.raises(E.QmmxError) as ei:
            eng.run_last_bars_static(long_h, long_l, cands)
        assert ei.value.status == abi.E_UNSUPPORTED and "QMMX_STATIC_MAX_BARS 12000" in str(ei.value)
        oc, xb = eng.run_last_bars_static(high, low, [])               # nothing to do is not an error
        assert len(oc) == 0 and len(xb) == 0
        oc, xb = eng.run_last_bars_static(high, low, cands)            # and the entry point still works afterwards
        assert np.array_equal(oc, eng.run_last_bars(high, low, cands)[0])


def test_gpu_last_bars_static_two_contexts_share_the_array_in_turn():
    """two ctxs on one device resolve the same array; calls in turn (not overlapping) do not disturb each other"""
    high, low, cands = make_case(2000, 3000, 5)
    with E.MCEngine(0) as a, E.MCEngine(0) as b:
        ref = a.run_last_bars(high, low, cands)
        ra = a.run_last_bars_static(high, low, cands)
        rb = b.run_last_bars_static(high[:1500], low[:1500], [c for c in cands if c["i"] < 1500])
        ra2 = a.run_last_bars_static(high, low, cands)
        assert np.array_equal(ra[0], ref[0]) and np.array_equal(ra[1], ref[1])
        assert np.array_equal(ra2[0], ref[0]) and np.array_equal(ra2[1], ref[1])
        rb_ref = b.run_last_bars(high[:1500], low[:1500], [c for c in cands if c["i"] < 1500])
        assert np.array_equal(rb[0], rb_ref[0]) and np.array_equal(rb[1], rb_ref[1])
