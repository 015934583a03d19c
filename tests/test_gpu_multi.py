"""The library's own all-reduce kernels (rg_allreduce_p2p, rg_allreduce_multimem, rg_allreduce_ranges) on real GPUs of
one node, against ncclAllReduce and a rank-ordered torch sum: tests/multi_gpu_check.py under torchrun.  Skipped on a
box with fewer than two GPUs (the round's own runs of it under `gpurun --gpus 2/8` are kept under profiles/)."""
import os
import subprocess
import sys

import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least two GPUs")
def test_own_allreduce_kernels_against_nccl():
    n = min(torch.cuda.device_count(), 8)
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29617", os.path.join(helpers.ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "rg_allreduce_p2p" in r.stdout
