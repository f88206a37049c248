"""Python host side of libqmmx_mc.so: a thin ctypes layer over the C ABI (include/qmmx_mc.h).

The reference has no FFI; its interface for this path is object state on `MonolithicEngine`
(`levels_cache`, `CONTACT_PROX`, `STOP_PADDING`, ... qmmx_monolithic.py:1368-1386) plus the call shapes
`evaluate_entry` (M:1492) / `simulate_monte_carlo` (M:3353).  `MCEngine` keeps that shape: levels are the
reference's list-of-dicts, settings use the reference's setting names, reason codes keep their names.

There is no CPU fallback: if libqmmx_mc.so is missing or no CUDA device is present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import abi

_LIB = None
_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libqmmx_mc.so")


class QmmxError(RuntimeError):
    def __init__(self, status: int, detail: str = ""):
        self.status = status
        super().__init__(f"libqmmx_mc status {status}: {detail}")


def load_library(path: Optional[str] = None):
    """dlopen libqmmx_mc.so and declare the prototypes of include/qmmx_mc.h.  Raises if it is absent."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    p = path or os.environ.get("QMMX_LIB") or _LIB_PATH      # QMMX_LIB: a tuning / debug build beside the product library
    if not os.path.exists(p):
        raise QmmxError(abi.E_STATE, f"{p} not found: build it with `python -m qmmx_monolithic_monte_carlo_b200.build` "
                                     "(there is no CPU fallback)")
    L = C.CDLL(p)
    vp, i32, i64, u64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
    protos = {
        "qmmx_abi_version": (C.c_int, []),
        "qmmx_strerror": (C.c_char_p, [C.c_int]),
        "qmmx_last_error": (C.c_char_p, [vp]),
        "qmmx_create": (C.c_int, [C.c_int, C.POINTER(vp)]),
        "qmmx_destroy": (C.c_int, [vp]),
        "qmmx_set_stream": (C.c_int, [vp, vp]),
        "qmmx_set_levels": (C.c_int, [vp, vp, vp, C.c_int]),
        "qmmx_set_params": (C.c_int, [vp, vp, C.c_int]),
        "qmmx_set_lr": (C.c_int, [vp, vp, C.c_int]),
        "qmmx_set_policy": (C.c_int, [vp, vp, C.c_int]),
        "qmmx_set_pnl_hist_range": (C.c_int, [vp, dbl, dbl]),
        "qmmx_upload_bars": (C.c_int, [vp, vp, vp, vp, i64, i32]),
        "qmmx_bind_device_bars": (C.c_int, [vp, vp, vp, vp, i64, i32, i64]),
        "qmmx_device_bars": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(i64)]),
        "qmmx_pin_host": (C.c_int, [vp, vp, C.c_size_t]),
        "qmmx_unpin_host": (C.c_int, [vp, vp]),
        "qmmx_download_bars": (C.c_int, [vp, vp, vp, vp]),
        "qmmx_upload_volume": (C.c_int, [vp, vp, i64, i32]),
        "qmmx_download_volume": (C.c_int, [vp, vp]),
        "qmmx_set_bootstrap_volume": (C.c_int, [vp, vp, i32]),
        "qmmx_gen_gbm": (C.c_int, [vp, u64, u64, i64, i32, C.POINTER(abi.Gbm)]),
        "qmmx_gen_bootstrap": (C.c_int, [vp, u64, u64, i64, i32, vp, i32, i32]),
        "qmmx_run_scan": (C.c_int, [vp, vp, vp, vp, i32, vp]),
        "qmmx_run_scan_host": (C.c_int, [vp, vp, vp, vp, i64, i32, i64, vp]),
        "qmmx_run_fused_gbm": (C.c_int, [vp, u64, u64, i64, i32, C.POINTER(abi.Gbm), vp]),
        "qmmx_run_scan_async": (C.c_int, [vp]),
        "qmmx_run_fused_gbm_async": (C.c_int, [vp, u64, u64, i64, i32, C.POINTER(abi.Gbm)]),
        "qmmx_run_fused_bootstrap": (C.c_int, [vp, u64, u64, i64, i32, vp, i32, i32, vp]),
        "qmmx_run_fused_bootstrap_async": (C.c_int, [vp, u64, u64, i64, i32, vp, i32, i32]),
        "qmmx_fetch_agg": (C.c_int, [vp, vp]),
        "qmmx_agg_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]),
        "qmmx_agg_unpack": (C.c_int, [vp, vp, C.c_int, vp]),
        "qmmx_merge_flat": (C.c_int, [vp, vp, C.c_int]),
        "qmmx_nccl_unique_id": (C.c_int, [vp, vp]),
        "qmmx_nccl_init": (C.c_int, [vp, vp, C.c_int, C.c_int]),
        "qmmx_allreduce": (C.c_int, [vp, vp]),
        "qmmx_get_stream": (C.c_int, [vp, C.POINTER(vp)]),
        "qmmx_lr_score": (C.c_int, [vp, vp, i64, i32, vp, vp, i32, vp]),
        "qmmx_run_jitter": (C.c_int, [vp, vp, vp, i32, vp, vp, vp, vp, i32, dbl, dbl, vp, i64, u64, u64,
                                      vp, vp, vp, vp, vp, vp, vp]),
        "qmmx_run_last_bars": (C.c_int, [vp, vp, vp, i32, vp, vp, vp, vp, i32, vp, vp]),
        "qmmx_run_last_bars_static": (C.c_int, [vp, vp, vp, i32, vp, vp, vp, vp, i32, vp, vp]),
        "qmmx_debug_classify": (C.c_int, [vp, vp, C.c_int, dbl, dbl, vp, i64, vp, vp, vp, vp, vp, vp]),
        "qmmx_debug_class_partition": (C.c_int, [vp, vp, C.c_int, dbl, dbl, vp, vp, vp]),
        "qmmx_last_timing": (C.c_int, [vp, C.POINTER(C.c_float)]),
        "qmmx_launch_count": (C.c_int, [vp, C.POINTER(u64)]),
        "qmmx_measure_fp32_peak": (C.c_int, [vp, C.POINTER(C.c_double)]),
        "qmmx_measure_issue_peak": (C.c_int, [vp, vp, vp]),
        "qmmx_download_policy_weights": (C.c_int, [vp, vp, i64]),
    }
    for name, (res, args) in protos.items():
        fn = getattr(L, name)          # AttributeError here = header/library mismatch
        fn.restype, fn.argtypes = res, args
    if L.qmmx_abi_version() != abi.ABI_VERSION:
        raise QmmxError(abi.E_STATE, "ABI version mismatch between abi.py and libqmmx_mc.so")
    L._protos = protos
    if path is None:
        _LIB = L
    return L


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32_2d(a, name):
    a = np.asarray(a)
    if a.dtype != np.float32 or not a.flags.c_contiguous:
        a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim != 2:
        raise ValueError(f"{name} must be [steps][paths]")
    return a


def sort_levels(levels: Sequence[dict]):
    """The reference's `load_levels` order: ORDER BY color, level_type, level_index (M:140-144).
    Returns (rows, price float64[L], solid uint8[L])."""
    rows = sorted(levels, key=lambda lv: (str(lv["color"]), str(lv["type"]), int(lv["index"])))
    price = np.array([float(lv["price"]) for lv in rows], dtype=np.float64)
    solid = np.array([1 if str(lv["type"]) == "solid" else 0 for lv in rows], dtype=np.uint8)
    return rows, price, solid


@dataclass
class MCResult:
    """Aggregates of one parameter combo (qmmx_agg) with the reference's reason-code names."""
    agg: "abi.Agg"

    @property
    def n_paths(self): return int(self.agg.n_paths)
    @property
    def n_steps(self): return int(self.agg.n_steps)
    @property
    def n_trades(self): return int(self.agg.n_trades)
    @property
    def n_wins(self): return int(self.agg.n_wins)
    @property
    def n_open(self): return int(self.agg.n_open)
    @property
    def entry_hist(self): return {n: int(self.agg.entry_hist[i]) for i, n in enumerate(abi.ENTRY_CODES)}
    @property
    def exit_hist(self): return {n: int(self.agg.exit_hist[i]) for i, n in enumerate(abi.EXIT_CODES)}
    @property
    def pnl_hist(self): return np.array(list(self.agg.pnl_hist), dtype=np.uint64)
    @property
    def sum_pnl(self): return float(self.agg.sum_pnl)
    @property
    def mean_path_pnl(self): return self.sum_pnl / max(1, self.n_paths)
    @property
    def std_path_pnl(self):
        n = max(1, self.n_paths)
        m = self.sum_pnl / n
        return float(max(0.0, self.agg.sum_path_pnl2 / n - m * m)) ** 0.5
    @property
    def hit_rate(self): return self.n_wins / max(1, self.n_trades)

    def as_dict(self):
        a = self.agg
        return dict(n_paths=self.n_paths, n_steps=self.n_steps, n_trades=self.n_trades, n_wins=self.n_wins,
                    n_open=self.n_open, n_trade_overflow=int(a.n_trade_overflow), entry_hist=self.entry_hist,
                    exit_hist=self.exit_hist, sum_pnl=a.sum_pnl, sum_pnl2=a.sum_pnl2,
                    sum_path_pnl2=a.sum_path_pnl2, sum_open_mtm=a.sum_open_mtm, max_path_pnl=a.max_path_pnl,
                    min_path_pnl=a.min_path_pnl)


class MCEngine:
    """One GPU context.  Mirrors the state the reference keeps on `MonolithicEngine`."""

    def __init__(self, device: int = 0, lib_path: Optional[str] = None):
        self._L = load_library(lib_path)
        h = C.c_void_p()
        rc = self._L.qmmx_create(int(device), C.byref(h))
        if rc != 0:
            raise QmmxError(rc, f"qmmx_create(device={device}) failed: {self._L.qmmx_strerror(rc).decode()} "
                                "(a CUDA sm_100 device is required; there is no CPU fallback)")
        self._h = h
        self.device = int(device)
        self.n_combos = 0
        self.levels_cache = []
        self._pinned = []

    # ---- lifetime ----
    def close(self):
        if getattr(self, "_h", None):
            self._L.qmmx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _ck(self, rc):
        if rc != 0:
            raise QmmxError(rc, self._L.qmmx_last_error(self._h).decode() or self._L.qmmx_strerror(rc).decode())

    # ---- configuration (reference: engine.levels_cache / settings / engine.model / app.policy) ----
    def set_levels(self, levels: Sequence[dict]):
        rows, price, solid = sort_levels(levels)
        self.levels_cache = rows
        self._ck(self._L.qmmx_set_levels(self._h, _ptr(price), _ptr(solid), len(rows)))

    def set_levels_raw(self, price, solid):
        price = np.ascontiguousarray(price, dtype=np.float64)
        solid = np.ascontiguousarray(solid, dtype=np.uint8)
        self._ck(self._L.qmmx_set_levels(self._h, _ptr(price), _ptr(solid), len(price)))

    def set_params(self, params):
        """One `abi.Params` or a sequence of them (a parameter-grid sweep)."""
        plist = [params] if isinstance(params, abi.Params) else list(params)
        arr = (abi.Params * len(plist))(*plist)
        self._ck(self._L.qmmx_set_params(self._h, C.cast(arr, C.c_void_p), len(plist)))
        self.n_combos = len(plist)

    def set_lr(self, coef_intercept=None):
        """sklearn-form LR per day: rows of [coef0..coef3, intercept] (M:1454-1466).  None = no model."""
        if coef_intercept is None:
            self._ck(self._L.qmmx_set_lr(self._h, None, 0))
            return
        a = np.ascontiguousarray(coef_intercept, dtype=np.float64).reshape(-1, abi.LR_DIM)
        self._ck(self._L.qmmx_set_lr(self._h, _ptr(a), len(a)))

    def set_policy(self, weights=None):
        """OnlinePolicy entry heads per day: [n_days][3 (go_long, go_short, skip)][7] (M:296-331)."""
        if weights is None:
            self._ck(self._L.qmmx_set_policy(self._h, None, 0))
            return
        a = np.ascontiguousarray(weights, dtype=np.float64).reshape(-1, abi.POLICY_HEADS, abi.POLICY_DIM)
        self._ck(self._L.qmmx_set_policy(self._h, _ptr(a), len(a)))

    def set_pnl_hist_range(self, lo, hi):
        self._ck(self._L.qmmx_set_pnl_hist_range(self._h, float(lo), float(hi)))

    def set_stream(self, cuda_stream_ptr: Optional[int]):
        self._ck(self._L.qmmx_set_stream(self._h, C.c_void_p(cuda_stream_ptr or 0)))

    # ---- bars ----
    def upload_bars(self, close, high=None, low=None):
        close = _f32_2d(close, "close")
        high = None if high is None else _f32_2d(high, "high")
        low = None if low is None else _f32_2d(low, "low")
        steps, paths = close.shape
        self._ck(self._L.qmmx_upload_bars(self._h, _ptr(close), _ptr(high), _ptr(low), paths, steps))

    def upload_volume(self, volume):
        """Per-bar volume of the resident bars ([steps][paths]); None drops it.  Read by the soft veto only
        (ENABLE_VETO: `_calc_volume_slope` M:1796-1811, `_soft_veto` M:1773-1794)."""
        if volume is None:
            self._ck(self._L.qmmx_upload_volume(self._h, None, 0, 0))
            return
        volume = _f32_2d(volume, "volume")
        steps, paths = volume.shape
        self._ck(self._L.qmmx_upload_volume(self._h, _ptr(volume), paths, steps))

    def download_volume(self, paths, steps):
        v = np.empty((steps, paths), dtype=np.float32)
        self._ck(self._L.qmmx_download_volume(self._h, _ptr(v)))
        return v

    def bind_device_bars(self, close_ptr: int, high_ptr: Optional[int], low_ptr: Optional[int], paths: int,
                         steps: int, pitch: Optional[int] = None):
        self._ck(self._L.qmmx_bind_device_bars(self._h, C.c_void_p(close_ptr), C.c_void_p(high_ptr or 0),
                                               C.c_void_p(low_ptr or 0), paths, steps, pitch or paths))

    def download_bars(self, paths, steps, hl=False):
        c = np.empty((steps, paths), dtype=np.float32)
        h = np.empty((steps, paths), dtype=np.float32) if hl else None
        l = np.empty((steps, paths), dtype=np.float32) if hl else None
        self._ck(self._L.qmmx_download_bars(self._h, _ptr(c), _ptr(h), _ptr(l)))
        return c, h, l

    def pin(self, arr: np.ndarray):
        self._ck(self._L.qmmx_pin_host(self._h, _ptr(arr), arr.nbytes))
        self._pinned.append(arr)

    def unpin(self, arr: np.ndarray):
        self._ck(self._L.qmmx_unpin_host(self._h, _ptr(arr)))
        self._pinned = [a for a in self._pinned if a is not arr]

    def gen_gbm(self, seed, path_offset, paths, steps, gbm: "abi.Gbm"):
        self._ck(self._L.qmmx_gen_gbm(self._h, seed, path_offset, paths, steps, C.byref(gbm)))

    def gen_bootstrap(self, seed, path_offset, paths, steps, window_ohlc, block_len, window_volume=None):
        """window_volume [window_len]: also generate the per-bar volume plane (soft veto input)."""
        w = np.ascontiguousarray(window_ohlc, dtype=np.float32).reshape(-1, 4)
        wv = None if window_volume is None else np.ascontiguousarray(window_volume, dtype=np.float32)
        if wv is not None and len(wv) != len(w):
            raise ValueError("window_volume must have one value per window bar")
        self._ck(self._L.qmmx_set_bootstrap_volume(self._h, _ptr(wv), len(w)))
        self._ck(self._L.qmmx_gen_bootstrap(self._h, seed, path_offset, paths, steps, _ptr(w), len(w), block_len))

    # ---- the hot path ----
    def _aggs(self):
        return (abi.Agg * max(1, self.n_combos))()

    def run_scan(self, *, trades: int = 0, codes: bool = False, paths_steps=None):
        """Scan the resident bars.  trades > 0 / codes=True switch on the parity-mode outputs (combo 0)."""
        aggs = self._aggs()
        if trades <= 0 and not codes:
            self._ck(self._L.qmmx_run_scan(self._h, C.cast(aggs, C.c_void_p), None, None, 0, None))
            return [MCResult(a) for a in aggs]
        steps, paths = paths_steps
        tr = np.zeros((paths, max(trades, 1)), dtype=abi.TRADE_DTYPE) if trades > 0 else None
        cnt = np.zeros(paths, dtype=np.int32) if trades > 0 else None
        cd = np.zeros((steps, paths), dtype=np.uint8) if codes else None
        self._ck(self._L.qmmx_run_scan(self._h, C.cast(aggs, C.c_void_p), _ptr(tr), _ptr(cnt), trades, _ptr(cd)))
        return [MCResult(a) for a in aggs], tr, cnt, cd

    def run_scan_async(self):
        self._ck(self._L.qmmx_run_scan_async(self._h))

    def run_scan_host(self, close, high=None, low=None, chunk_paths: int = 0):
        close = _f32_2d(close, "close")
        high = None if high is None else _f32_2d(high, "high")
        low = None if low is None else _f32_2d(low, "low")
        steps, paths = close.shape
        aggs = self._aggs()
        self._ck(self._L.qmmx_run_scan_host(self._h, _ptr(close), _ptr(high), _ptr(low), paths, steps,
                                            chunk_paths, C.cast(aggs, C.c_void_p)))
        return [MCResult(a) for a in aggs]

    def run_fused_gbm(self, seed, path_offset, paths, steps, gbm: "abi.Gbm"):
        aggs = self._aggs()
        self._ck(self._L.qmmx_run_fused_gbm(self._h, seed, path_offset, paths, steps, C.byref(gbm),
                                            C.cast(aggs, C.c_void_p)))
        return [MCResult(a) for a in aggs]

    def run_fused_gbm_async(self, seed, path_offset, paths, steps, gbm: "abi.Gbm"):
        self._ck(self._L.qmmx_run_fused_gbm_async(self._h, seed, path_offset, paths, steps, C.byref(gbm)))

    def run_fused_bootstrap(self, seed, path_offset, paths, steps, window_ohlc, block_len):
        """Block-bootstrap the candle window and scan in one kernel (the bars of gen_bootstrap, never written)."""
        w = np.ascontiguousarray(window_ohlc, dtype=np.float32).reshape(-1, 4)
        aggs = self._aggs()
        self._ck(self._L.qmmx_run_fused_bootstrap(self._h, seed, path_offset, paths, steps, _ptr(w), len(w), block_len,
                                                  C.cast(aggs, C.c_void_p)))
        return [MCResult(a) for a in aggs]

    def run_fused_bootstrap_async(self, seed, path_offset, paths, steps, window_ohlc, block_len):
        w = np.ascontiguousarray(window_ohlc, dtype=np.float32).reshape(-1, 4)
        self._ck(self._L.qmmx_run_fused_bootstrap_async(self._h, seed, path_offset, paths, steps, _ptr(w), len(w), block_len))

    def fetch_agg(self):
        aggs = self._aggs()
        self._ck(self._L.qmmx_fetch_agg(self._h, C.cast(aggs, C.c_void_p)))
        return [MCResult(a) for a in aggs]

    def agg_device(self):
        """(device pointer, bytes) of the flat reduction buffer: n_combos * AGG_SLOTS 8-byte slots."""
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self._L.qmmx_agg_device(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def merge_flat(self, d_parts_ptr: int, n_parts: int):
        """Reduce all-gathered flat buffers (device pointer, [n_parts][n_combos * AGG_SLOTS]) into this ctx's."""
        self._ck(self._L.qmmx_merge_flat(self._h, C.c_void_p(d_parts_ptr), n_parts))

    def nccl_unique_id(self) -> bytes:
        """128 bytes from ncclGetUniqueId (rank 0 calls this and ships them to the other ranks)."""
        buf = C.create_string_buffer(128)
        self._ck(self._L.qmmx_nccl_unique_id(self._h, buf))
        return buf.raw

    def nccl_init(self, unique_id: bytes, world_size: int, rank: int):
        """Join the library's own NCCL communicator (ncclCommInitRank; collective over all ranks)."""
        if len(unique_id) != 128:
            raise ValueError("an NCCL unique id is 128 bytes")
        self._ck(self._L.qmmx_nccl_init(self._h, C.c_char_p(unique_id), world_size, rank))

    def allreduce(self, nccl_comm: Optional[int] = None):
        """`qmmx_allreduce`: one ncclAllGather + the merge kernel on the ctx stream; afterwards fetch_agg() on any rank
        returns the whole job's aggregates.  nccl_comm: a raw ncclComm_t, or None for the communicator of nccl_init."""
        self._ck(self._L.qmmx_allreduce(self._h, C.c_void_p(nccl_comm) if nccl_comm else None))

    def stream(self) -> int:
        """Raw cudaStream_t the ctx launches on."""
        p = C.c_void_p()
        self._ck(self._L.qmmx_get_stream(self._h, C.byref(p)))
        return p.value or 0

    def unpack_flat(self, flat_host: np.ndarray):
        flat_host = np.ascontiguousarray(flat_host).view(np.uint64).reshape(-1, abi.AGG_SLOTS)
        aggs = (abi.Agg * len(flat_host))()
        self._ck(self._L.qmmx_agg_unpack(self._h, _ptr(flat_host), len(flat_host), C.cast(aggs, C.c_void_p)))
        return [MCResult(a) for a in aggs]

    def lr_score(self, x, w, bias=None):
        x = np.ascontiguousarray(x, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
        n, dim = x.shape
        heads = w.shape[0]
        out = np.empty((n, heads), dtype=np.float32)
        self._ck(self._L.qmmx_lr_score(self._h, _ptr(x), n, dim, _ptr(w), _ptr(b), heads, _ptr(out)))
        return out

    def run_jitter(self, high, low, cands, stop_padding, tp_padding, sigmas, trials, seed=0, trial_offset=0,
                   replay=None):
        """Trials x candidates walk-forward of the reference MC (M:3449-3510).
        cands: list of dict(i, side, lvl, entry).  replay = (normals[trials][K][4], uniforms[trials][K])."""
        high = np.ascontiguousarray(high, dtype=np.float64)
        low = np.ascontiguousarray(low, dtype=np.float64)
        cb = np.array([c["i"] for c in cands], dtype=np.int32)
        cs = np.array([c["side"] for c in cands], dtype=np.uint8)
        cl = np.array([c["lvl"] for c in cands], dtype=np.float64)
        ce = np.array([c["entry"] for c in cands], dtype=np.float64)
        sg = np.ascontiguousarray(sigmas, dtype=np.float64)
        rn = ru = None
        if replay is not None:
            rn = np.ascontiguousarray(replay[0], dtype=np.float64)
            ru = np.ascontiguousarray(replay[1], dtype=np.float64)
        tot = np.empty(trials, dtype=np.float64); dd = np.empty(trials, dtype=np.float64)
        w = np.empty(trials, dtype=np.int32); lo = np.empty(trials, dtype=np.int32); op = np.empty(trials, dtype=np.int32)
        self._ck(self._L.qmmx_run_jitter(self._h, _ptr(high), _ptr(low), len(high), _ptr(cb), _ptr(cs), _ptr(cl),
                                         _ptr(ce), len(cands), float(stop_padding), float(tp_padding), _ptr(sg),
                                         trials, seed, trial_offset, _ptr(rn), _ptr(ru), _ptr(tot), _ptr(dd),
                                         _ptr(w), _ptr(lo), _ptr(op)))
        return dict(totals=tot, drawdowns=dd, wins=w, losses=lo, opens=op)

    def run_last_bars(self, high, low, cands):
        """Forward walk of `simulate_last_bars` (M:3619-3639) for cands = list of dict(i, side, stop, target).
        Returns (outcome[K] 0 tp / 1 stop / 2 open, exit_bar[K] or -1)."""
        high = np.ascontiguousarray(high, dtype=np.float64)
        low = np.ascontiguousarray(low, dtype=np.float64)
        cb = np.array([c["i"] for c in cands], dtype=np.int32)
        cs = np.array([c["side"] for c in cands], dtype=np.uint8)
        st = np.array([c["stop"] for c in cands], dtype=np.float64)
        tg = np.array([c["target"] for c in cands], dtype=np.float64)
        oc = np.full(len(cands), 2, dtype=np.int32); xb = np.full(len(cands), -1, dtype=np.int32)
        self._ck(self._L.qmmx_run_last_bars(self._h, _ptr(high), _ptr(low), len(high), _ptr(cb), _ptr(cs), _ptr(st),
                                            _ptr(tg), len(cands), _ptr(oc), _ptr(xb)))
        return oc, xb

    def run_last_bars_static(self, high, low, cands):
        """`run_last_bars` through `qmmx_run_last_bars_static`: no device allocation, scratch in the library's static
        1 MiB device array (len(high) <= 12000, len(cands) <= 16384; no reference counterpart, see include/qmmx_mc.h)."""
        high = np.ascontiguousarray(high, dtype=np.float64)
        low = np.ascontiguousarray(low, dtype=np.float64)
        cb = np.array([c["i"] for c in cands], dtype=np.int32)
        cs = np.array([c["side"] for c in cands], dtype=np.uint8)
        st = np.array([c["stop"] for c in cands], dtype=np.float64)
        tg = np.array([c["target"] for c in cands], dtype=np.float64)
        oc = np.full(len(cands), 2, dtype=np.int32); xb = np.full(len(cands), -1, dtype=np.int32)
        self._ck(self._L.qmmx_run_last_bars_static(self._h, _ptr(high), _ptr(low), len(high), _ptr(cb), _ptr(cs), _ptr(st),
                                                   _ptr(tg), len(cands), _ptr(oc), _ptr(xb)))
        return oc, xb

    def last_timing_ms(self) -> float:
        ms = C.c_float()
        self._ck(self._L.qmmx_last_timing(self._h, C.byref(ms)))
        return float(ms.value)

    def download_policy_weights(self, paths: int) -> np.ndarray:
        """ONLINE_RETRAIN: the go_long[7] + go_short[7] heads each path has learnt in the last run_scan, `[paths][14]` f64."""
        out = np.empty((paths, 14), dtype=np.float64)
        self._ck(self._L.qmmx_download_policy_weights(self._h, _ptr(out), paths))
        return out

    def measure_fp32_peak(self) -> float:
        """FFMA-only TFLOP/s of this device, measured now (the FP32 roofline denominator of SURVEY 8(d))."""
        v = C.c_double()
        self._ck(self._L.qmmx_measure_fp32_peak(self._h, C.byref(v)))
        return float(v.value)

    def measure_issue_peak(self):
        """(ALU-pipe, ALU + FMA) issue-rate probes of this device in 1e9 warp-instructions per second (measured now)."""
        a, m = C.c_double(0.0), C.c_double(0.0)
        self._ck(self._L.qmmx_measure_issue_peak(self._h, C.byref(a), C.byref(m)))
        return a.value, m.value

    def launch_count(self) -> int:
        n = C.c_uint64()
        self._L.qmmx_launch_count(self._h, C.byref(n))
        return int(n.value)


def debug_classify(price, solid, prox, qmin, p):
    """Host-only check of the exact classification tables (no device)."""
    L = load_library()
    price = np.ascontiguousarray(price, dtype=np.float64)
    solid = np.ascontiguousarray(solid, dtype=np.uint8)
    p = np.ascontiguousarray(p, dtype=np.float32)
    n = len(p)
    out = dict(level_table=np.empty(n, np.int32), level_brute=np.empty(n, np.int32),
               conf_mask=np.empty(n, np.uint8), conf_brute=np.empty(n, np.uint8),
               near_table=np.empty(n, np.uint64), near_brute=np.empty(n, np.uint64))
    rc = L.qmmx_debug_classify(_ptr(price), _ptr(solid), len(price), float(prox), float(qmin), _ptr(p), n,
                               *[_ptr(out[k]) for k in ("level_table", "level_brute", "conf_mask", "conf_brute",
                                                        "near_table", "near_brute")])
    if rc != 0:
        raise QmmxError(rc, "qmmx_debug_classify")
    return out


def debug_class_partition(price, solid, prox, qmin):
    """(classes, runs, single-boundary records) of the class partition the lean / tile kernels read; classes = -1 when the
    level set does not fit it (host only)."""
    L = load_library()
    price = np.ascontiguousarray(price, dtype=np.float64)
    solid = np.ascontiguousarray(solid, dtype=np.uint8)
    out = [C.c_int32(0) for _ in range(3)]
    rc = L.qmmx_debug_class_partition(_ptr(price), _ptr(solid), len(price), float(prox), float(qmin), *[C.byref(o) for o in out])
    if rc != 0:
        raise QmmxError(rc, "qmmx_debug_class_partition")
    return tuple(o.value for o in out)
