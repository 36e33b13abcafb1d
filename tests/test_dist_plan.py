"""Host-side logic of the multi-GPU build (no GPU): the key-range and column-range plans
exported by the C ABI, alone and across two gloo processes."""
import os
import subprocess
import sys

import numpy as np
import pytest

import founderblockgraphs_b200 as pkg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_key_ranges_partition_and_balance(world):
    rng = np.random.default_rng(world)
    for hist in (rng.integers(0, 50, size=65536).astype(np.uint32),
                 np.zeros(65536, np.uint32),
                 np.concatenate([np.zeros(100, np.uint32), np.array([10 ** 6], np.uint32), rng.integers(0, 5, size=400).astype(np.uint32)])):
        b = pkg.api.plan_key_ranges(hist, world)
        assert b[0] == 0 and b[-1] == len(hist) and np.all(np.diff(b.astype(np.int64)) >= 0)
        total = int(hist.sum())
        cum = np.concatenate([[0], np.cumsum(hist.astype(np.int64))])
        for r in range(1, world):
            target = total // world * r + (total % world) * r // world
            # the cut is the first bin boundary at which the running count reaches the target
            assert cum[b[r]] >= target and (b[r] == 0 or cum[b[r] - 1] < target or target == 0 or b[r] == b[r - 1])
        if total and hist.max() <= 50:
            share = np.diff(cum[b.astype(np.int64)])
            assert share.max() - share.min() <= 2 * 50 + 1


@pytest.mark.parametrize("n,world", [(1, 1), (1, 4), (7, 3), (30000, 8), (5_000_001, 2)])
def test_column_ranges_tile_the_msa(n, world):
    edges = [pkg.api.plan_columns(n, world, r) for r in range(world)]
    assert edges[0][0] == 0 and edges[-1][1] == n
    for (a0, a1), (b0, b1) in zip(edges, edges[1:]):
        assert a1 == b0 and a0 <= a1
    assert max(b - a for a, b in edges) - min(b - a for a, b in edges) <= 1


WORKER = r"""
import os, sys
sys.path.insert(0, %r)
import numpy as np
import torch, torch.distributed as dist
import founderblockgraphs_b200 as pkg
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# every rank holds the replicated text, hence the same histogram: plans must agree without communication
rng = np.random.default_rng(7)
hist = rng.integers(0, 1000, size=65536).astype(np.uint32)
b = pkg.api.plan_key_ranges(hist, world)
mine = torch.tensor([int(b[rank]), int(b[rank + 1]), int(hist[b[rank]:b[rank + 1]].sum())])
allr = [torch.zeros(3, dtype=torch.long) for _ in range(world)]
dist.all_gather(allr, mine)
assert allr[0][0] == 0 and allr[-1][1] == 65536
for a, c in zip(allr, allr[1:]):
    assert a[1] == c[0]
assert sum(int(a[2]) for a in allr) == int(hist.sum())
# column plan of f(x): merge by max reproduces the whole vector
n = 1000
f = np.zeros(n, np.int64)
x0, x1 = pkg.api.plan_columns(n, world, rank)
f[x0:x1] = np.arange(x0, x1) + 1
t = torch.from_numpy(f)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
assert np.array_equal(t.numpy(), np.arange(n) + 1)
dist.destroy_process_group()
"""


def test_plans_agree_across_two_gloo_ranks(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script)]
    pr = subprocess.run(cmd, capture_output=True, timeout=300, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert pr.returncode == 0, pr.stderr.decode()[-3000:]
