"""Level fixtures and synthetic-input recipes shared by tests and bench.

The 30-level Blue/Orange/Black/Teal fixture is the one recovered from the reference's
`qmmx.db-wal`; `BLUE_SOLID_4` is config 1's "one Blue/Solid set".
"""
from __future__ import annotations

import numpy as np

_FIXTURE = {
    ("black", "dashed"): [649.20, 647.41, 644.50, 641.90, 640.29],
    ("black", "solid"): [652.11, 646.60, 643.20, 637.69],
    ("blue", "dashed"): [651.45, 649.60, 646.60, 643.91, 642.25],
    ("blue", "solid"): [654.44, 648.76, 645.24, 639.56],
    ("orange", "dashed"): [652.83, 650.95, 647.90, 645.16, 643.46],
    ("orange", "solid"): [655.88, 650.09, 646.51, 640.72],
    ("teal", "solid"): [500.00, 600.00, 700.00],
}


def levels_30():
    """The recovered 30-level fixture as the reference's list-of-dicts (M:140-144)."""
    out = []
    for (color, typ), prices in _FIXTURE.items():
        for i, p in enumerate(prices):
            out.append({"color": color, "type": typ, "index": i, "price": p})
    return out


def levels_blue_solid_4():
    return [{"color": "blue", "type": "solid", "index": i, "price": p}
            for i, p in enumerate(_FIXTURE[("blue", "solid")])]


GBM_S0 = 648.90
GBM_SIGMA_ANNUAL = 0.15
GBM_MU_ANNUAL = 0.0
GBM_DT = 1.0 / (252.0 * 390.0)
HOST_SEED = 20250829


def host_gbm_closes(paths: int, steps: int = 390, *, seed: int = HOST_SEED, s0: float = GBM_S0,
                    sigma: float = GBM_SIGMA_ANNUAL, mu: float = GBM_MU_ANNUAL,
                    round_cents: bool = False, chunk: int = 1 << 16) -> np.ndarray:
    """Host GBM closes, fp32 `[step][path]`: numpy Philox(key=seed)."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    drift = (mu - 0.5 * sigma * sigma) * GBM_DT
    vol = sigma * np.sqrt(GBM_DT)
    out = np.empty((steps, paths), dtype=np.float32)
    for lo in range(0, paths, chunk):
        hi = min(paths, lo + chunk)
        z = rng.standard_normal((steps, hi - lo))
        px = s0 * np.exp(np.cumsum(drift + vol * z, axis=0))
        if round_cents:
            px = np.round(px, 2)
        out[:, lo:hi] = px.astype(np.float32)
    return out


def synthetic_window(n_bars: int = 40, *, seed: int = 5, s0: float = GBM_S0, sigma: float = GBM_SIGMA_ANNUAL) -> np.ndarray:
    """A seeded OHLC candle window `[n_bars][4]` (o, h, l, c), fp32: 4 GBM sub-steps per bar."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    vol = sigma * np.sqrt(GBM_DT / 4.0)
    sub = s0 * np.exp(np.cumsum(-0.5 * vol * vol + vol * rng.standard_normal(4 * n_bars))).reshape(n_bars, 4)
    o = np.concatenate([[s0], sub[:-1, 3]])
    h = np.maximum(o, sub.max(axis=1)); l = np.minimum(o, sub.min(axis=1))
    return np.stack([o, h, l, sub[:, 3]], axis=1).astype(np.float32)
