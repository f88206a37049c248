"""Multi-GPU plumbing: paths shard across ranks, one merge of the aggregate vector at the end (an NCCL all-gather of
2.3 KB per combo + the library's merge kernel; `allreduce_flat_` is the three-message all-reduce form for host tensors).

The reference is single-process; its Monte Carlo trials share nothing but read-only bars
(qmmx_monolithic.py:3491-3498), so ranks never exchange bars.  Each rank owns global path ids
[rank*P/R, (rank+1)*P/R); Philox counters are functions of the global id, so integer aggregates are identical
for any rank count.  The only collective is a SUM of counters and float64 sums and a MAX of {max, -min}.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi


def shard_range(total_paths: int, rank: int, world: int):
    """Global path ids of `rank`: contiguous, sizes differ by at most one."""
    lo = total_paths * rank // world
    hi = total_paths * (rank + 1) // world
    return lo, hi


def agg_to_flat(agg: "abi.Agg") -> np.ndarray:
    """abi.Agg -> the flat 8-byte-slot layout of include/qmmx_mc.h (uint64 view)."""
    u = np.zeros(abi.AGG_NU, dtype=np.uint64)
    u[0:6] = [agg.n_paths, agg.n_steps, agg.n_trades, agg.n_wins, agg.n_open, agg.n_trade_overflow]
    u[6:6 + abi.N_ENTRY_CODES] = list(agg.entry_hist)
    u[6 + abi.N_ENTRY_CODES:6 + abi.N_ENTRY_CODES + abi.N_EXIT_CODES] = list(agg.exit_hist)
    u[6 + abi.N_ENTRY_CODES + abi.N_EXIT_CODES:] = list(agg.pnl_hist)
    d = np.array([agg.sum_pnl, agg.sum_pnl2, agg.sum_path_pnl2, agg.sum_open_mtm], dtype=np.float64)
    x = np.array([agg.max_path_pnl, -agg.min_path_pnl], dtype=np.float64)
    return np.concatenate([u, d.view(np.uint64), x.view(np.uint64)])


def flat_to_agg(flat: np.ndarray, hist_range=(-8.0, 8.0), combo: int = 0) -> "abi.Agg":
    flat = np.ascontiguousarray(flat).view(np.uint64)
    nu, nd = abi.AGG_NU, abi.AGG_ND
    u, d, x = flat[:nu], flat[nu:nu + nd].view(np.float64), flat[nu + nd:nu + nd + abi.AGG_NX].view(np.float64)
    a = abi.Agg()
    a.struct_size = C.sizeof(abi.Agg); a.combo = combo
    a.n_paths, a.n_steps, a.n_trades, a.n_wins, a.n_open, a.n_trade_overflow = [int(v) for v in u[0:6]]
    o = 6
    for i in range(abi.N_ENTRY_CODES):
        a.entry_hist[i] = int(u[o + i])
    o += abi.N_ENTRY_CODES
    for i in range(abi.N_EXIT_CODES):
        a.exit_hist[i] = int(u[o + i])
    o += abi.N_EXIT_CODES
    for i in range(abi.PNL_HIST_BINS):
        a.pnl_hist[i] = int(u[o + i])
    a.sum_pnl, a.sum_pnl2, a.sum_path_pnl2, a.sum_open_mtm = [float(v) for v in d]
    a.max_path_pnl, a.min_path_pnl = float(x[0]), -float(x[1])
    a.pnl_hist_lo, a.pnl_hist_hi = hist_range
    return a


def allreduce_flat_(flat_i64, group=None):
    """In-place all-reduce of a torch int64 tensor [n_combos, AGG_SLOTS] holding the flat aggregate(s):
    SUM over the integer counters, SUM over the float64 sums, MAX over {max_path_pnl, -min_path_pnl}.
    Works for CUDA tensors (NCCL over NVLink) and CPU tensors (gloo).  Three latency-bound messages of
    <= 2.3 KB per combo; issue once after the last kernel."""
    import torch
    import torch.distributed as dist
    t = flat_i64.view(-1, abi.AGG_SLOTS)
    nu, nd = abi.AGG_NU, abi.AGG_ND
    u = t[:, :nu].contiguous()
    d = t[:, nu:nu + nd].contiguous().view(torch.float64)
    x = t[:, nu + nd:].contiguous().view(torch.float64)
    dist.all_reduce(u, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(d, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(x, op=dist.ReduceOp.MAX, group=group)
    t[:, :nu].copy_(u)
    t[:, nu:nu + nd].copy_(d.view(torch.int64))
    t[:, nu + nd:].copy_(x.view(torch.int64))
    return flat_i64


def allgather_merge_(eng, device_index: int, group=None, scratch=None):
    """The N>1 merge through torch.distributed: ONE NCCL all-gather of the 2.3 KB-per-combo vector, then the library's
    own merge kernel (qmmx_merge_flat).  The gather runs on torch's current stream and the scan and merge kernels on
    the ctx stream, so the two are ordered explicitly here: torch's stream waits for the ctx stream before the gather
    (the aggregates are final) and the ctx stream waits for torch's before the merge (the gather has landed).  Returns
    the gather scratch tensor for reuse.  A caller without torch uses `MCEngine.allreduce` (qmmx_allreduce) instead."""
    import torch
    import torch.distributed as dist
    flat = device_flat_tensor(eng, device_index)
    world = dist.get_world_size(group)
    if scratch is None or scratch.shape != (world,) + tuple(flat.shape):
        scratch = torch.empty((world,) + tuple(flat.shape), dtype=torch.int64, device=flat.device)
    cur = torch.cuda.current_stream(flat.device)
    raw = eng.stream()
    ctx_stream = cur if raw == cur.cuda_stream else torch.cuda.ExternalStream(raw, device=flat.device)
    if ctx_stream is not cur:
        cur.wait_stream(ctx_stream)
    dist.all_gather_into_tensor(scratch, flat, group=group)
    if ctx_stream is not cur:
        ctx_stream.wait_stream(cur)
    eng.merge_flat(scratch.data_ptr(), world)
    return scratch


def nccl_init_from_torch(eng, group=None):
    """Give the engine its own NCCL communicator (qmmx_nccl_init) using torch.distributed only to ship rank 0's
    128-byte unique id; afterwards `eng.allreduce()` merges the ranks without touching torch."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [eng.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    eng.nccl_init(box[0], world, rank)


def device_flat_tensor(eng, device_index: int):
    """torch int64 view [n_combos, AGG_SLOTS] of the engine's device reduction buffer (zero copy)."""
    import torch
    ptr, nbytes = eng.agg_device()

    class _Mem:
        __cuda_array_interface__ = {"shape": (nbytes // 8,), "typestr": "<i8", "data": (ptr, False), "version": 3}
    return torch.as_tensor(_Mem(), device=f"cuda:{device_index}").view(-1, abi.AGG_SLOTS)
