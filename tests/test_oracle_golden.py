"""CPU: the C restatement (oracle/fbg_oracle.c) against the golden vectors generated from
the reference binary (tests/golden/make_golden.py).  This is what pins the oracle."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from util import FIXTURES, case_id, elastic_cases, flags_to_opts, nonelastic_cases, stat_lines


@pytest.mark.parametrize("case", elastic_cases(), ids=case_id)
def test_elastic_against_reference_golden(case, tmp_path):
    o = flags_to_opts(case["flags"])
    rows, ids, _ = O.read_fasta(os.path.join(FIXTURES, case["fixture"]), elastic=True, output_paths=o["paths"])
    msa = O.msa_array(rows)
    st = stat_lines(case)
    gaps, toign = O.counts(msa, o["ignore"])
    assert gaps == st["gaps"] and toign == st["toignore"]
    f = O.compute_f(msa, o["ignore"], o["tricks_off"])
    if case["exit"] != 0:
        assert o["tricks_off"] and f[0] == msa.shape[1]  # "No valid segmentation found!" (fbg.cpp:1932-1937)
        return
    mml, bt, b = O.dp_elastic(f)
    assert mml[-1] == st["L"]
    if "f" in case:
        n = msa.shape[1]
        assert f.tolist() == case["f"]
        ok = np.array(case["mml"]) <= n
        assert mml[ok].tolist() == np.array(case["mml"])[ok].tolist()
        assert bt[ok].tolist() == np.array(case["backtrack"])[ok].tolist()
        assert b.tolist() == case["boundaries"]
    out = str(tmp_path / "o.xgfa")
    O.write_xgfa(out, msa, b, ids=ids if o["paths"] else None)
    assert open(out, "rb").read() == case["xgfa"]


@pytest.mark.parametrize("case", nonelastic_cases(), ids=case_id)
def test_nonelastic_against_reference_golden(case):
    gl = int(case["flags"][0].split("=")[1])
    rows, _, msgs = O.read_fasta(os.path.join(FIXTURES, case["fixture"]), gap_limit=gl, elastic=False)
    notices = [l for l in case["stderr"] if l.startswith(("NOTICE", "WARNING"))]
    assert msgs == notices
    if not rows:
        assert case["exit"] == 1 and any(l.startswith("Unable") for l in case["stderr"])
        return
    msa = O.msa_array(rows)
    assert msa.shape[0] == case["harness_m"]
    st = stat_lines(case)
    v = O.v_from_f(O.compute_f(msa, b"", True))
    failed, s, prev, b, score = O.dp_nonelastic(v, greedy=(gl != 1))
    assert score == st["Optimal score: "]
    if failed:
        assert case["exit"] == 1 and case["harness_status"] == 1
        return
    assert case["exit"] == -11  # the reference prints its statistics and then crashes
    assert v.tolist() == case["v"] and s.tolist() == case["s"] and prev.tolist() == case["prev"] and b.tolist() == case["boundaries"]
    g = O.graph(msa, b)
    assert len(b) == st["Number of segments: "]
    assert len(g["node_row"]) == st["#nodes="] and g["label_total"] == st["total length of node labels="]
    assert int(g["heights"].max()) == st["#founders="] and len(g["edge_src"]) == st["#edges="]


def test_sa_lcp_restatement_small():
    rng = np.random.default_rng(3)
    for _ in range(20):
        n = int(rng.integers(1, 200))
        T = np.concatenate([rng.integers(65, 68, size=n).astype(np.uint8), np.zeros(1, np.uint8)])
        SA, ISA, LCP = O.sa_lcp(T)
        suf = sorted(range(len(T)), key=lambda i: bytes(T[i:]))
        assert SA.tolist() == suf
        assert all(ISA[SA[k]] == k for k in range(len(T)))
        for k in range(1, len(T)):
            a, b = bytes(T[SA[k - 1]:]), bytes(T[SA[k]:])
            l = 0
            while l < min(len(a), len(b)) and a[l] == b[l]:
                l += 1
            assert LCP[k] == l
