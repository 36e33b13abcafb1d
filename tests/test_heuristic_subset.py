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
