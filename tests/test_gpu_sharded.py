"""GPU, >= 2 devices (-m gpu): one MSA sharded across two GPUs gives the bits of the
single-GPU build (tools/sharded_check.py: SA, ISA, LCP, f, boundaries, nodes, edges)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_build_matches_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tools", "sharded_check.py")]
    pr = subprocess.run(cmd, capture_output=True, timeout=900)
    assert pr.returncode == 0, pr.stdout.decode()[-4000:] + pr.stderr.decode()[-4000:]
    assert b"MISMATCH" not in pr.stdout
