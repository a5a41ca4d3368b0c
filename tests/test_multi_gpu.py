"""SURVEY.md 8(e) on real GPUs: launches tests/mgpu_worker.py under torchrun with one rank per visible GPU
(2..8).  Skipped on a single-GPU box (the host logic is covered on gloo by tests/test_parallel_cpu.py)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_slab_volume_and_icp_round_robin_nccl():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    n = min(n, 8)
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29517", os.path.join(here, "mgpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MGPU OK world=%d" % n in r.stdout
