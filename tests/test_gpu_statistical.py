"""Statistical tier of the parity contract (north star: "with on-device Philox paths, the PnL and exit-reason
distributions match the reference's statistics within stated confidence bounds").

Device side: the fused generate+scan kernels on their own Philox4x32-10 streams.  Host side: paths drawn with
`numpy.random.Generator(Philox)` (float64 GBM / block bootstrap, cast to fp32) and run through the C oracle.  The two
samples are independent draws of the same law, so every aggregate is compared with a two-sample test:

* per-path means of every entry reason code (20 bins), every exit code (4 bins), trades and wins per path, PnL per path:
  z = (m_dev - m_host) / sqrt(s^2 (1/n_dev + 1/n_host)), s^2 = the per-path variance measured on a 100k-path host
  subsample (steps of one path are correlated, paths are not, so the path is the sampling unit); the family is held at
  ALPHA = 1e-4 with a Bonferroni split over its bins;
* the 256-bin histogram of per-path PnL: two-sample chi-square on the bins with at least 25 paths, p > ALPHA.
A wrong constant anywhere in the generators or the engine (drift, volatility, a gate threshold, a padding) moves these
by tens of sigmas at this sample size; a correct implementation fails a family once in 10^4 runs.
"""
import os

import numpy as np
import pytest
from scipy import stats

from oracle import c_oracle as CO
from qmmx_monolithic_monte_carlo_b200 import abi
from qmmx_monolithic_monte_carlo_b200 import engine as E
from qmmx_monolithic_monte_carlo_b200 import fixtures as F

pytestmark = pytest.mark.gpu
ALPHA = 1e-4
LR = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
THREADS = os.cpu_count() or 4


@pytest.fixture(scope="module")
def eng():
    e = E.MCEngine(0)
    yield e
    e.close()


def per_path_variances(o, steps, paths):
    """variances of the per-path counts from an oracle run with codes and trades: entry codes, exit codes, trades, wins, pnl"""
    codes, tr, cnt = o["codes"], o["trades"], o["counts"]
    v_entry = np.array([np.var((codes == c).sum(axis=0)) for c in range(abi.N_ENTRY_CODES)])
    m = np.arange(tr.shape[1])[None, :] < np.minimum(cnt, tr.shape[1])[:, None]
    v_exit = np.array([np.var(((tr["reason"] == r) & m).sum(axis=1)) for r in range(abi.N_EXIT_CODES)])
    closed = m & (tr["exit_bar"] >= 0)
    v_trades = np.var(closed.sum(axis=1))
    v_wins = np.var((closed & (tr["pnl"] > 0)).sum(axis=1))
    v_pnl = np.var(np.where(closed, tr["pnl"].astype(np.float64), 0.0).sum(axis=1))
    assert cnt.max() <= tr.shape[1]
    return v_entry, v_exit, v_trades, v_wins, v_pnl


def z_family(name, dev_tot, n_dev, host_tot, n_host, var, alpha=ALPHA):
    dev_tot, host_tot, var = (np.atleast_1d(np.asarray(x, dtype=np.float64)) for x in (dev_tot, host_tot, var))
    # (nearly) constant bins are compared directly, to 1e-3: DIR_UNKNOWN is 1 step per path plus a ~1e-4 chance per step
    # that two consecutive fp32 closes coincide, and THAT depends on the generator's arithmetic (the device recurses in
    # fp32, where exp(x) rounds to exactly 1 for |x| < 6e-8; the host rounds a float64 path once), not on the engine
    live = var > 1e-3
    assert np.allclose(dev_tot[~live] / n_dev, host_tot[~live] / n_host, rtol=1e-3, atol=1e-3), name
    z = (dev_tot[live] / n_dev - host_tot[live] / n_host) / np.sqrt(var[live] * (1.0 / n_dev + 1.0 / n_host))
    zc = stats.norm.isf(alpha / (2 * max(1, live.sum())))
    assert np.all(np.abs(z) < zc), (name, z.round(2).tolist(), round(float(zc), 2))
    return z


def chi2_two_sample(h1, h2, alpha=ALPHA, min_count=25):
    h1, h2 = np.asarray(h1, dtype=np.float64), np.asarray(h2, dtype=np.float64)
    keep = (h1 + h2) >= min_count
    a = np.append(h1[keep], h1[~keep].sum()); b = np.append(h2[keep], h2[~keep].sum())
    a, b = a[(a + b) > 0], b[(a + b) > 0]
    k1, k2 = np.sqrt(b.sum() / a.sum()), np.sqrt(a.sum() / b.sum())
    stat = float(np.sum((k1 * a - k2 * b) ** 2 / (a + b)))
    p = float(stats.chi2.sf(stat, len(a) - 1))
    assert p > alpha, (stat, len(a) - 1, p)
    return stat, p


def compare(dev, n_dev, host, n_host, var):
    v_entry, v_exit, v_trades, v_wins, v_pnl = var
    z_family("entry reason codes", list(dev.entry_hist), n_dev, list(host.entry_hist), n_host, v_entry)
    z_family("exit reason codes", list(dev.exit_hist), n_dev, list(host.exit_hist), n_host, v_exit)
    z_family("trades per path", dev.n_trades, n_dev, host.n_trades, n_host, v_trades)
    z_family("wins per path (hit rate numerator)", dev.n_wins, n_dev, host.n_wins, n_host, v_wins)
    z_family("mean PnL per path", dev.sum_pnl, n_dev, host.sum_pnl, n_host, v_pnl)
    # second moment of the per-path PnL: its sampling variance is E[x^4] - E[x^2]^2, bounded here through the histogram
    centers = dev.pnl_hist_lo + (np.arange(abi.PNL_HIST_BINS) + 0.5) * (dev.pnl_hist_hi - dev.pnl_hist_lo) / abi.PNL_HIST_BINS
    hh = np.array(list(host.pnl_hist), dtype=np.float64)
    m2 = float((hh * centers ** 2).sum() / hh.sum()); m4 = float((hh * centers ** 4).sum() / hh.sum())
    z_family("variance of PnL per path", dev.sum_path_pnl2, n_dev, host.sum_path_pnl2, n_host, max(m4 - m2 * m2, 1e-12))
    chi2_two_sample(list(dev.pnl_hist), list(host.pnl_hist))


def test_gpu_philox_gbm_statistics_match_host_numpy_philox(eng):
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.10, cooldown_s=120)
    n_host, n_dev, T = 1_000_000, 4_000_000, 390
    close = F.host_gbm_closes(n_host, T, seed=777_001)                       # numpy Philox, float64 GBM -> fp32
    host = CO.run_paths(close, price, solid, params, lr=LR, n_threads=THREADS)["agg"]
    sub = 100_000
    var = per_path_variances(CO.run_paths(np.ascontiguousarray(close[:, :sub]), price, solid, params, lr=LR, want_codes=True,
                                          max_trades=96, n_threads=THREADS), T, sub)
    del close
    eng.set_levels(F.levels_30()); eng.set_params(params); eng.set_lr(LR); eng.set_policy(None)
    dev = eng.run_fused_gbm(424_242, 0, n_dev, T, abi.make_gbm())[0].agg       # device Philox4x32-10 + Box-Muller, fp32
    assert dev.n_paths == n_dev and host.n_paths == n_host
    compare(dev, n_dev, host, n_host, var)


def numpy_block_bootstrap(window, paths, steps, block_len, seed):
    """Circular block bootstrap with numpy's Philox: the law of gen_bootstrap_kernel (start uniform over the window's
    returns, fp32 products in sequence), independent draws."""
    w = np.asarray(window, dtype=np.float32)
    pc = w[:-1, 3]
    rh, rl, rc = w[1:, 1] / pc, w[1:, 2] / pc, w[1:, 3] / pc
    nret = len(rc)
    rng = np.random.Generator(np.random.Philox(key=seed))
    nblk = (steps + block_len - 1) // block_len
    start = rng.integers(0, nret, size=(nblk, paths))
    idx = ((start[:, None, :] + np.arange(block_len)[None, :, None]) % nret).reshape(nblk * block_len, paths)[:steps]
    close = np.multiply.accumulate(np.concatenate([np.full((1, paths), w[-1, 3], dtype=np.float32), rc[idx]]), axis=0, dtype=np.float32)
    prev = close[:-1]
    return close[1:], (prev * rh[idx]).astype(np.float32), (prev * rl[idx]).astype(np.float32)


def test_gpu_fused_bootstrap_statistics_match_host_numpy_philox(eng):
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True)
    window = F.synthetic_window(40)
    n_host, n_dev, T = 400_000, 2_000_000, 390
    c, h, l = numpy_block_bootstrap(window, n_host, T, 5, seed=31_337)
    host = CO.run_paths(c, price, solid, params, high=h, low=l, lr=LR, n_threads=THREADS)["agg"]
    sub = 100_000
    var = per_path_variances(CO.run_paths(np.ascontiguousarray(c[:, :sub]), price, solid, params, lr=LR, want_codes=True, max_trades=128,
                                          high=np.ascontiguousarray(h[:, :sub]), low=np.ascontiguousarray(l[:, :sub]),
                                          n_threads=THREADS), T, sub)
    del c, h, l
    eng.set_levels(F.levels_30()); eng.set_params(params); eng.set_lr(LR); eng.set_policy(None)
    dev = eng.run_fused_bootstrap(99, 0, n_dev, T, window, 5)[0].agg
    compare(dev, n_dev, host, n_host, var)
