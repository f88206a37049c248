"""N > 1 on real GPUs (needs >= 2 devices: `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu`; skipped
on the one-GPU box the driver uses).  The host-side merge logic has its own world_size-2 gloo test on CPU
(tests/test_host_logic.py)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_nccl_merges_equal_single_engine():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "tests", "mgpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MGPU_OK" in r.stdout, (r.stdout[-2000:], r.stderr[-4000:])
