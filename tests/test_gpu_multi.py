"""Multi-GPU correctness: launches tests/multi_gpu_worker.py with one rank per GPU."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _run(nproc, timeout=180):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + nproc),
           os.path.join(HERE, "multi_gpu_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    out = p.stdout + p.stderr
    assert p.returncode == 0, out[-4000:]
    assert "MULTI_GPU_RESULT OK" in out, out[-4000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpus_all_styles():
    _run(2)


@pytest.mark.skipif(torch.cuda.device_count() < 8, reason="needs 8 GPUs")
def test_eight_gpus_all_styles():
    _run(8)
