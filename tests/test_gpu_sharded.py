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


def test_cli_on_two_gpus_writes_the_same_file(tmp_path):
    """FBG_GPUS=2: the drop-in executable forks a second rank, both build ONE graph; the xGFA bytes are those of the
    1-GPU run (and of the reference, tests/golden/large_cases.json)."""
    import hashlib
    import json
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    sys.path.insert(0, ROOT)
    import founderblockgraphs_b200 as pkg
    from founderblockgraphs_b200 import synth
    from oracle import oracle as O
    with open(os.path.join(ROOT, "tests", "golden", "large_cases.json")) as fh:
        gold = {c["name"]: c for c in json.load(fh)}
    for name, extra in (("cfg4_twin_0.01_N_p", ["--elastic", "--gfa", "--ignore-chars=N", "-p"]), ("cfg3_twin_0.1", ["--elastic", "--gfa"])):
        c = gold[name]
        msa, _ = synth.make_config(c["gen"][1], c["gen"][2], seed=c["gen"][3])
        inp = str(tmp_path / (name + ".fasta"))
        O.write_fasta(inp, [bytes(r) for r in msa])
        digests = []
        for gpus in ("1", "2"):
            out = str(tmp_path / ("%s_%s.xgfa" % (name, gpus)))
            env = dict(os.environ, FBG_GPUS=gpus)
            pr = pkg.api.main(["--input", inp, "--output", out] + extra, capture_output=True, timeout=600, env=env)
            assert pr.returncode == 0, pr.stderr.decode()[-2000:]
            if gpus == "2":
                assert b"Building on 2 GPUs." in pr.stderr
            digests.append(hashlib.sha256(open(out, "rb").read()).hexdigest())
        assert digests[0] == digests[1] == c["xgfa_sha256"]
