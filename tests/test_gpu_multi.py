"""Runs tests/multi_gpu_check.py under torchrun on every power-of-two GPU count the box has (>= 2):
sharded transform, sharded multiexp and partition scheduling against the single-GPU results."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_multi_gpu_parity():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run: gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu)")
    world = 1 << (n.bit_length() - 1)
    world = min(world, 8)
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert out.stdout.count("multi-GPU checks passed") == world
