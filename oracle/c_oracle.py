"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/qmmx_oracle.c (the C restatement).

Importable only from tests/, `__graft_entry__.smoke()` and bench.py's CPU legs.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from qmmx_monolithic_monte_carlo_b200 import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libqmmx_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "qmmx_oracle.c")
    hdr = os.path.join(_HERE, "..", "include", "qmmx_mc.h")
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(f) and os.path.getmtime(f) > os.path.getmtime(_SO) for f in (src, hdr))
    if force or stale:
        subprocess.check_call(["make", "-s", "-C", _HERE] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.oq_run_paths.restype = C.c_int
        L.oq_run_paths.argtypes = [
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
            C.c_void_p, C.c_void_p, C.c_int, C.POINTER(abi.Params),
            C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_double,
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(abi.Agg), C.c_int]
        L.oq_run_paths_v.restype = C.c_int
        L.oq_run_paths_v.argtypes = [
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
            C.c_void_p, C.c_void_p, C.c_int, C.POINTER(abi.Params),
            C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_double,
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(abi.Agg), C.c_int]
        L.oq_gen_gbm.restype = C.c_int
        L.oq_gen_gbm.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int32, C.POINTER(abi.Gbm),
                                 C.c_void_p, C.c_void_p, C.c_void_p]
        L.oq_gen_bootstrap.restype = C.c_int
        L.oq_gen_bootstrap.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int32, C.c_void_p,
                                       C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oq_gen_bootstrap_v.restype = C.c_int
        L.oq_gen_bootstrap_v.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                         C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oq_regime_series.restype = None
        L.oq_regime_series.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
        L.oq_volume_slope.restype = C.c_double
        L.oq_volume_slope.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int]
        L.oq_philox4x32_10.restype = None
        L.oq_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.oq_dot_py312.restype = C.c_double
        L.oq_dot_py312.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.oq_confidence.restype = C.c_double
        L.oq_confidence.argtypes = [C.c_double, C.c_int, C.c_int, C.c_double]
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def run_paths(close, levels_price, levels_solid, params: "abi.Params", *, high=None, low=None, volume=None,
              lr=None, policy=None, hist_range=(-8.0, 8.0), want_codes=False, max_trades=0,
              n_threads=1):
    """Oracle batch run. close/high/low: fp32 [steps][paths]. Returns dict(agg, codes, trades, counts)."""
    close = np.ascontiguousarray(close, dtype=np.float32)
    steps, paths = close.shape
    high = None if high is None else np.ascontiguousarray(high, dtype=np.float32)
    low = None if low is None else np.ascontiguousarray(low, dtype=np.float32)
    volume = None if volume is None else np.ascontiguousarray(volume, dtype=np.float32)
    assert volume is None or volume.shape == close.shape
    lp = np.ascontiguousarray(levels_price, dtype=np.float64)
    ls = np.ascontiguousarray(levels_solid, dtype=np.uint8)
    lr_a = None if lr is None else np.ascontiguousarray(lr, dtype=np.float64).reshape(-1, 5)
    pol_a = None if policy is None else np.ascontiguousarray(policy, dtype=np.float64).reshape(-1, 21)
    codes = np.empty((steps, paths), dtype=np.uint8) if want_codes else None
    trades = np.zeros((paths, max_trades), dtype=abi.TRADE_DTYPE) if max_trades > 0 else None
    counts = np.zeros(paths, dtype=np.int32) if max_trades > 0 else None
    agg = abi.Agg()
    polw = None
    if params.flags & abi.F_ONLINE_RETRAIN:        # learnt go_long / go_short heads per path
        polw = np.zeros((paths, 14), dtype=np.float64)
        lib().oq_set_policy_weights_out.argtypes = [C.c_void_p]
        lib().oq_set_policy_weights_out(_ptr(polw))
    rc = lib().oq_run_paths_v(_ptr(close), _ptr(high), _ptr(low), _ptr(volume), paths, steps, _ptr(lp), _ptr(ls), len(lp),
                            C.byref(params), _ptr(lr_a), 0 if lr_a is None else len(lr_a),
                            _ptr(pol_a), 0 if pol_a is None else len(pol_a),
                            float(hist_range[0]), float(hist_range[1]),
                            _ptr(codes), _ptr(trades), _ptr(counts), max_trades, C.byref(agg), n_threads)
    if polw is not None:
        lib().oq_set_policy_weights_out(None)
    if rc != 0:
        raise RuntimeError(f"oq_run_paths failed: {rc}")
    return dict(agg=agg, codes=codes, trades=trades, counts=counts, policy_weights=polw)


def regime_series(close, volume):
    """uint8 [steps][paths]: AccumulationBreakoutGuard regime seen by evaluate_entry at every tick."""
    close = np.ascontiguousarray(close, dtype=np.float32)
    volume = np.ascontiguousarray(volume, dtype=np.float32)
    steps, paths = close.shape
    out = np.empty((paths, steps), dtype=np.uint8)
    for j in range(paths):
        lib().oq_regime_series(_ptr(close), _ptr(volume), paths, j, steps, out[j].ctypes.data_as(C.c_void_p))
    return np.ascontiguousarray(out.T)


def gen_gbm(seed, path_offset, paths, steps, gbm: "abi.Gbm", ohlc=False):
    close = np.empty((steps, paths), dtype=np.float32)
    high = np.empty((steps, paths), dtype=np.float32) if ohlc else None
    low = np.empty((steps, paths), dtype=np.float32) if ohlc else None
    rc = lib().oq_gen_gbm(seed, path_offset, paths, steps, C.byref(gbm), _ptr(close), _ptr(high), _ptr(low))
    if rc != 0:
        raise RuntimeError(f"oq_gen_gbm failed: {rc}")
    return close, high, low


def gen_bootstrap(seed, path_offset, paths, steps, window_ohlc, block_len, window_volume=None):
    """Returns (close, high, low), plus volume when window_volume [window_len] is given."""
    w = np.ascontiguousarray(window_ohlc, dtype=np.float32).reshape(-1, 4)
    wv = None if window_volume is None else np.ascontiguousarray(window_volume, dtype=np.float32)
    assert wv is None or len(wv) == len(w)
    close = np.empty((steps, paths), dtype=np.float32)
    high = np.empty((steps, paths), dtype=np.float32)
    low = np.empty((steps, paths), dtype=np.float32)
    vol = None if wv is None else np.empty((steps, paths), dtype=np.float32)
    rc = lib().oq_gen_bootstrap_v(seed, path_offset, paths, steps, _ptr(w), _ptr(wv), len(w), block_len,
                                  _ptr(close), _ptr(high), _ptr(low), _ptr(vol))
    if rc != 0:
        raise RuntimeError(f"oq_gen_bootstrap failed: {rc}")
    return (close, high, low) if wv is None else (close, high, low, vol)


def philox(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    lib().oq_philox4x32_10(_ptr(c), _ptr(k), _ptr(out))
    return out
