"""bench.py's command-line contract on a machine without a GPU: the reference arm prints ONE JSON
line with the keys the driver reads, the product arm refuses to run (no CPU path)."""
import json
import os
import subprocess
import sys

import pytest

from oracle import oracle as O
from util import ROOT

BENCH = os.path.join(ROOT, "bench.py")


@pytest.mark.skipif(not O.have_reference(), reason="oracle/_ref is not built")
def test_reference_arm_prints_one_json_line():
    pr = subprocess.run([sys.executable, BENCH, "--impl", "reference", "--steps", "1", "--warmup", "0", "--sample-cols", "200"],
                        capture_output=True, timeout=300)
    assert pr.returncode == 0, pr.stderr.decode()
    lines = [l for l in pr.stdout.decode().splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "msa_cells_per_sec_end_to_end_efg_build" and d["unit"] == "cells/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_product_arm_has_no_cpu_path():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    pr = subprocess.run([sys.executable, BENCH, "--steps", "1", "--warmup", "0"], capture_output=True, timeout=300)
    assert pr.returncode != 0 and pr.stdout.decode().strip() == ""
    assert "no CUDA device" in pr.stderr.decode()
