"""Real ranks: `torchrun --nproc-per-node 2 tests/dist_worker.py` -- every multi-GPU transport against the single-GPU fused
kernel, bit for bit (skipped on a box with fewer than two GPUs; the emulated-rank tests of test_gpu_shard.py and the gloo
tests of test_cpu_dist.py cover the host logic there)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_all_transports_equal_the_fused_kernel_on_two_ranks():
    n = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29711", os.path.join(ROOT, "tests", "dist_worker.py")]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    tail = (out.stdout + out.stderr)[-3000:]
    assert out.returncode == 0 and "DIST_WORKER_OK" in out.stdout, tail
