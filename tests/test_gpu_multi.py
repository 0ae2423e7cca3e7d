"""Multi-GPU checks on hardware: skipped unless the box shows at least two GPUs. The work is in
tests/dp_grad_check.py, launched here under torchrun exactly as bench.py's multi-GPU runs are."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least two GPUs on one box")
def test_two_gpu_gradients_equal_one_gpu_gradients():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29511", os.path.join(ROOT, "tests", "dp_grad_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "DP_CHECK" in r.stdout
