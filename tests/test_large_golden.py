"""Parity at the sizes and on the index paths bench.py times.

tests/golden/large_cases.json holds sha256 digests of what the REFERENCE produced on medium and
benchmark-sized synthetic MSAs (tests/golden/make_golden_large.py: 100 x 20 000 ... the full cfg2
and cfg3 MSAs bench.py builds).  The inputs are regenerated here from the same seeds.

  CPU (-m "not gpu"): the file is complete, the generator still yields the recorded MSAs, and the
      C restatement (oracle/) reproduces the reference's digests on the cases it finishes in seconds.
  GPU (-m gpu): f / DP tables / boundaries / xGFA bytes through the C ABI, in the default index mode
      (sufficient / virtual index where eligible) and with FBG_FULL_INDEX=1, and the xGFA file through
      the drop-in executable -- all compared by digest with the reference's.
"""
import hashlib
import importlib.util
import json
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as O
from util import GOLDEN, ROOT

_spec = importlib.util.spec_from_file_location("make_golden_large", os.path.join(GOLDEN, "make_golden_large.py"))
MG = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(MG)

with open(os.path.join(GOLDEN, "large_cases.json")) as _fh:
    CASES = {c["name"]: c for c in json.load(_fh)}

# cases whose MSA the CPU suite regenerates / the oracle finishes within seconds
CPU_REGEN = ["gapped_100x20000_p", "gapped_200x50000_p", "gapless_300x20000_elastic", "cfg2_full_nonelastic_v0", "cfg3_twin_0.02"]
CPU_ORACLE = ["gapped_100x20000_p"]
FULL = [n for n in CASES if "_full_" in n]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<u8").tobytes()).hexdigest()


def ids_of(m):
    return [b"h%d" % i for i in range(m)]


_msa_cache = {}


def case_msa(name):
    if name not in _msa_cache:
        _msa_cache.clear()  # one large MSA at a time
        _msa_cache[name] = MG.make_case_msa(tuple(CASES[name]["gen"]))
    return _msa_cache[name]


def test_large_golden_file_is_complete():
    assert set(MG.CASES) == set(CASES), "run tests/golden/make_golden_large.py"
    for c in CASES.values():
        assert c["status"] == 0 and len(c["xgfa_sha256"]) == 64 and c["xgfa_bytes"] > 0 and c["n_blocks"] > 0
        keys = ("f", "mml", "backtrack", "boundaries") if c["mode"] == "elastic" else ("v", "s", "prev", "boundaries")
        assert all(len(c[k + "_sha256"]) == 64 for k in keys)


@pytest.mark.parametrize("name", CPU_REGEN)
def test_generator_reproduces_recorded_msa(name):
    msa = case_msa(name)
    assert msa.shape == (CASES[name]["m"], CASES[name]["n"])
    assert hashlib.sha256(msa.tobytes()).hexdigest() == CASES[name]["msa_sha256"]


@pytest.mark.parametrize("name", CPU_ORACLE)
def test_oracle_matches_reference_digests(name, tmp_path):
    c = CASES[name]
    msa = case_msa(name)
    f = O.compute_f(msa, c["ignore"].encode(), False)
    assert digest(f) == c["f_sha256"]
    mml, bt, b = O.dp_elastic(f)
    assert digest(mml) == c["mml_sha256"] and digest(bt) == c["backtrack_sha256"] and digest(b) == c["boundaries_sha256"]
    p = str(tmp_path / "o.xgfa")
    O.write_xgfa(p, msa, b, ids_of(msa.shape[0]) if c["paths"] else None)
    assert hashlib.sha256(open(p, "rb").read()).hexdigest() == c["xgfa_sha256"]


# ------------------------------------------------------------------------------- GPU
def _gpu_check(ctx, name, full_index):
    c = CASES[name]
    msa = case_msa(name)
    m, n = msa.shape
    assert hashlib.sha256(msa.tobytes()).hexdigest() == c["msa_sha256"]
    elastic = c["mode"] == "elastic"
    old = os.environ.pop("FBG_FULL_INDEX", None)
    if full_index:
        os.environ["FBG_FULL_INDEX"] = "1"
    try:
        ctx.set_options(elastic=elastic, gap_limit=1, ignore_chars=c["ignore"].encode())
        ctx.upload(msa)
        f = ctx.compute_f()  # NOT preceded by build_index: the index the later stages choose themselves
        st = ctx.stats()
        if full_index:
            assert st["index_depth"] == 0
        b, score = ctx.segment()
        assert score == c["score"] and len(b) == c["n_blocks"]
        assert digest(b) == c["boundaries_sha256"]
        if elastic:
            assert digest(f) == c["f_sha256"]
            mml, bt = ctx.dp(True)
            assert digest(mml) == c["mml_sha256"] and digest(bt) == c["backtrack_sha256"]
        else:
            s, prev, v = ctx.dp(False)
            assert digest(v) == c["v_sha256"] and digest(s) == c["s_sha256"] and digest(prev) == c["prev_sha256"]
        info = ctx.format_xgfa(ids_of(m) if c["paths"] else None)
        assert info["n_bytes"] == c["xgfa_bytes"]
        h = hashlib.sha256()
        step = 64 << 20
        for off in range(0, info["n_bytes"], step):
            h.update(ctx.read_xgfa(off, min(step, info["n_bytes"] - off)))
        assert h.hexdigest() == c["xgfa_sha256"]
        return st
    finally:
        os.environ.pop("FBG_FULL_INDEX", None)
        if old is not None:
            os.environ["FBG_FULL_INDEX"] = old


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_matches_reference_digests(gpu_ctx, name):
    st = _gpu_check(gpu_ctx, name, full_index=False)
    if name.startswith("cfg2_full") or name.startswith("gapless"):
        # the path bench.py --workload cfg2 times: no SA / ISA / LCP array written at all
        assert st["index_virtual"] == 1 and st["index_depth"] > 0


@pytest.mark.gpu
@pytest.mark.parametrize("name", [n for n in sorted(CASES) if CASES[n]["m"] * CASES[n]["n"] <= 60_000_000])
def test_gpu_full_index_matches_reference_digests(gpu_ctx, name):
    _gpu_check(gpu_ctx, name, full_index=True)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["gapped_100x20000_p", "gapped_200x50000_p", "cfg2_full_nonelastic_v0", "cfg2_full_elastic_v0",
                                  "cfg3_twin_0.1", "cfg4_twin_0.01_N_p", "cfg5_twin_0.01"])
@pytest.mark.parametrize("host_writer", [False, True])
def test_cli_matches_reference_digests(name, host_writer, tmp_path):
    """The drop-in executable, FASTA file in, xGFA file out (device-side and host writer)."""
    import founderblockgraphs_b200 as pkg
    c = CASES[name]
    if host_writer and c["m"] * c["n"] > 30_000_000:
        pytest.skip("host writer checked on the smaller cases")
    msa = case_msa(name)
    inp, out = str(tmp_path / "in.fasta"), str(tmp_path / "out.xgfa")
    O.write_fasta(inp, [bytes(r) for r in msa])
    args = ["--input", inp, "--output", out]
    if c["mode"] == "elastic":
        args += ["--elastic", "--gfa"]
        if c["ignore"]:
            args += ["--ignore-chars=" + c["ignore"]]
        if c["paths"]:
            args += ["-p"]
    else:
        args += ["--gap-limit=1", "--nonelastic-gfa"]
    env = dict(os.environ)
    env.pop("FBG_HOST_WRITER", None)
    if host_writer:
        env["FBG_HOST_WRITER"] = "1"
    pr = pkg.api.main(args, capture_output=True, timeout=900, env=env)
    assert pr.returncode == 0, pr.stderr.decode()[-2000:]
    data = open(out, "rb").read()
    assert len(data) == c["xgfa_bytes"] and hashlib.sha256(data).hexdigest() == c["xgfa_sha256"]
    err = pr.stderr.decode()
    if c["mode"] == "elastic":
        assert "(optimal segment length = %d)" % c["score"] in err
    else:
        assert "Optimal score: %d" % c["score"] in err and "Number of segments: %d" % c["n_blocks"] in err
