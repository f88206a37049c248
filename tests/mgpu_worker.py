"""Worker of tests/test_multi_gpu.py (one rank per GPU under torch.distributed.run): the two multi-GPU merges --
torch's all-gather + qmmx_merge_flat WITHOUT set_stream (the ctx keeps its private stream: ordering must come from
allgather_merge_ itself) and the torch-free qmmx_allreduce on the library's own NCCL communicator -- against one
engine that scans all the paths."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from qmmx_monolithic_monte_carlo_b200 import abi, distributed as D, engine as E, fixtures as F  # noqa: E402


def ints(a):
    return [a.n_paths, a.n_steps, a.n_trades, a.n_wins, a.n_open] + list(a.entry_hist) + list(a.exit_hist) + list(a.pnl_hist)


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    total, T = 400_003, 390
    g = abi.make_gbm()
    grid = [abi.make_params(contact_prox=0.1, cooldown_s=120), abi.make_params(contact_prox=0.1, cooldown_s=8, stop_padding=0.5)]
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    lo, hi = D.shard_range(total, rank, world)
    with E.MCEngine(local) as eng:
        eng.set_levels(F.levels_30()); eng.set_params(grid); eng.set_lr(lr)
        # reference: this rank alone over ALL paths
        whole = [r.agg for r in eng.run_fused_gbm(9, 0, total, T, g)]
        # (a) torch all-gather + merge kernel, ctx on its own private stream (no set_stream)
        for _ in range(3):
            eng.run_fused_gbm_async(9, lo, hi - lo, T, g)
            D.allgather_merge_(eng, local)
        a = [r.agg for r in eng.fetch_agg()]
        # (b) qmmx_allreduce on the library's own communicator
        D.nccl_init_from_torch(eng)
        eng.run_fused_gbm_async(9, lo, hi - lo, T, g)
        eng.allreduce()
        b = [r.agg for r in eng.fetch_agg()]
        for c in range(len(grid)):
            assert ints(a[c]) == ints(whole[c]), ("allgather_merge_", rank, c)
            assert ints(b[c]) == ints(whole[c]), ("qmmx_allreduce", rank, c)
            for x in (a[c], b[c]):
                assert abs(x.sum_pnl - whole[c].sum_pnl) <= 1e-9 * max(1.0, abs(whole[c].sum_pnl))
                assert x.max_path_pnl == whole[c].max_path_pnl and x.min_path_pnl == whole[c].min_path_pnl
    dist.barrier()
    if rank == 0:
        print("MGPU_OK world=%d paths=%d trades=%d" % (world, total, whole[0].n_trades))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
