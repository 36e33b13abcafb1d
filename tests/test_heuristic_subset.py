"""--heuristic-subset=ROWNUM (fbg.cpp:2158-2178, 3400-3433, 3467-3500, efg_validate 3199-3292).

tests/golden/hs_cases.json holds what the reference binary writes in that mode (tests/golden/make_golden_hs.py).
  CPU: the oracle's restatement (row-chunk f, max, one DP, repair loop with a plain string search) reproduces the
       reference's xGFA bytes, its "blocks to remove" sequence and its optimal segment length.
  GPU: the C-ABI path (fbg_heuristic_f + fbg_segment + fbg_validate_blocks + K4 / K5) and the drop-in executable do.
"""
import base64
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import oracle as O
from util import FIXTURES, GOLDEN

with open(os.path.join(GOLDEN, "hs_cases.json")) as _fh:
    CASES = json.load(_fh)
for _c in CASES:
    _c["xgfa"] = base64.b64decode(_c["xgfa_b64"]) if _c.get("xgfa_b64") else None


def cid(c):
    return c["fixture"].replace(".fasta", "") + ":" + " ".join(c["flags"][2:])


OK_CASES = [c for c in CASES if c["exit"] == 0 and "--threads=2" not in c["flags"]]
ABORT_CASES = [c for c in CASES if c["exit"] != 0]


def load(c):
    rows, _, _ = O.read_fasta(os.path.join(FIXTURES, c["fixture"]), 1, True, False)
    return O.msa_array(rows)


def opts(c):
    return dict(ignore=b"N" if "--ignore-chars=N" in c["flags"] else b"", tricks_off="--disable-elastic-tricks" in c["flags"])


def test_goldens_cover_the_repair_loop():
    assert len(OK_CASES) > 100 and sum(1 for c in OK_CASES if sum(c["removed"]) > 0) > 50
    assert all("-p" in c["flags"] for c in ABORT_CASES)  # the reference asserts in output_efg_external (fbg.cpp:1421)
    # --threads does not change the output
    by = {(c["fixture"], c["hs"]): c for c in CASES if c["flags"][3:] == []}
    for c in CASES:
        if "--threads=2" in c["flags"]:
            assert c["xgfa_sha256"] == by[(c["fixture"], c["hs"])]["xgfa_sha256"]


@pytest.mark.parametrize("c", OK_CASES, ids=cid)
def test_oracle_heuristic_subset_matches_reference(c, tmp_path):
    msa = load(c)
    o = opts(c)
    b, removed, score, nosol = O.heuristic_subset(msa, c["hs"], o["ignore"], o["tricks_off"])
    assert not nosol and removed == c["removed"] and score == c["score"]
    p = str(tmp_path / "o.xgfa")
    O.write_xgfa(p, msa, b)
    assert open(p, "rb").read() == c["xgfa"]


# ------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("c", OK_CASES, ids=cid)
def test_gpu_heuristic_subset_matches_reference(gpu_ctx, c):
    msa = load(c)
    o = opts(c)
    gpu_ctx.set_options(elastic=True, gap_limit=1, ignore_chars=o["ignore"], disable_tricks=o["tricks_off"])
    b, removed, score, f = gpu_ctx.heuristic_subset(msa, c["hs"])
    assert removed == c["removed"] and score == c["score"]
    # f is the maximum over the row chunks of the chunk's own f
    f_ref = np.zeros(msa.shape[1], np.uint64)
    for r0 in range(0, msa.shape[0], c["hs"]):
        f_ref = np.maximum(f_ref, O.compute_f(np.ascontiguousarray(msa[r0: r0 + c["hs"]]), o["ignore"], o["tricks_off"]))
    assert np.array_equal(f, f_ref)
    assert gpu_ctx.xgfa_bytes() == c["xgfa"]
    # there is no index of the whole MSA in this mode, and the library says so
    import founderblockgraphs_b200 as pkg
    with pytest.raises(pkg.api.FbgError):
        gpu_ctx.sa()


@pytest.mark.gpu
@pytest.mark.parametrize("c", [c for c in CASES if c["fixture"] in ("hs_rep_10x150.fasta", "hs_rep_gapsN_12x200.fasta", "hs_rep_40x300.fasta", "msa.fasta")
                               and c["hs"] in (2, 5)], ids=cid)
def test_cli_heuristic_subset_matches_reference(c, tmp_path):
    import re
    import founderblockgraphs_b200 as pkg
    out = str(tmp_path / "o.xgfa")
    pr = pkg.api.main(["--input", os.path.join(FIXTURES, c["fixture"]), "--output", out] + c["flags"], capture_output=True, timeout=300)
    err = pr.stderr.decode()
    if c["exit"] != 0:  # -p: the reference dies of its own assertion while writing (fbg.cpp:1421)
        assert pr.returncode == 134 and not os.path.exists(out)
        return
    assert pr.returncode == 0, err[-2000:]
    assert open(out, "rb").read() == c["xgfa"]
    assert [int(x) for x in re.findall(r"There are (\d+) blocks to remove", err)] == c["removed"]
    assert "(optimal segment length = %d)" % c["score"] in err and "Graph fixed in %diterations" % (len(c["removed"]) - 1) in err
