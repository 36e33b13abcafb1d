"""CPU, only where the reference binaries exist (this container): seeded random MSAs
through the real reference (oracle/_ref) against the C restatement."""
import os

import numpy as np
import pytest

from founderblockgraphs_b200 import synth
from oracle import oracle as O

pytestmark = pytest.mark.skipif(not O.have_reference(), reason="oracle/_ref binaries not built (no /root/reference)")


@pytest.mark.parametrize("seed", range(6))
def test_elastic_random(seed, tmp_path):
    rng = np.random.default_rng(1000 + seed)
    for it in range(6):
        m, n = int(rng.integers(1, 9)), int(rng.integers(1, 45))
        n_p = float(rng.choice([0, 0.05]))
        msa = synth.random_small_msa(rng, m, n, gap_p=float(rng.choice([0, 0.1, 0.3])), n_p=n_p, alphabet=bytes(rng.choice([b"AC", b"ACGT"])))
        if (msa == 45).all(axis=1).any():
            continue  # all-gap row + -p is undefined behaviour in the reference (Q4)
        ignore = b"N" if n_p > 0 else b""
        tricks_off = bool(rng.random() < 0.3)
        fa = str(tmp_path / ("e%d.fasta" % it))
        O.write_fasta(fa, [bytes(r) for r in msa])
        args = (["--ignore-chars", "N"] if ignore else []) + (["--disable-elastic-tricks"] if tricks_off else []) + ["--output-paths"]
        res, err, ref_x = O.run_harness(fa, "elastic", args, want_output=True)
        f = O.compute_f(msa, ignore, tricks_off)
        if "f" not in res:
            assert tricks_off and f[0] == n
            continue
        assert np.array_equal(f, res["f"])
        mml, bt, b = O.dp_elastic(f)
        assert np.array_equal(b, res["boundaries"])
        out = str(tmp_path / "o.xgfa")
        O.write_xgfa(out, msa, b, ids=[b"h%d" % i for i in range(m)])
        assert open(out, "rb").read() == ref_x


@pytest.mark.parametrize("seed", range(4))
def test_nonelastic_random(seed, tmp_path):
    rng = np.random.default_rng(2000 + seed)
    for it in range(8):
        m, n = int(rng.integers(1, 9)), int(rng.integers(1, 60))
        gapped = rng.random() < 0.5
        msa = synth.random_small_msa(rng, m, n, gap_p=(0.1 if gapped else 0.0), alphabet=bytes(rng.choice([b"AC", b"ACGT"])), div=0.5)
        gl = 0 if gapped else 1
        fa = str(tmp_path / ("n%d.fasta" % it))
        O.write_fasta(fa, [bytes(r) for r in msa])
        res, err, _ = O.run_harness(fa, "nonelastic", ["--gap-limit", str(gl)])
        v = O.v_from_f(O.compute_f(msa, b"", True))
        failed, s, prev, b, score = O.dp_nonelastic(v, greedy=(gl != 1))
        assert score == int(err.split("Optimal score: ")[1].split()[0])
        if failed:
            assert res["status"] == 1
            continue
        for k, mine in (("v", v), ("s", s), ("prev", prev), ("boundaries", b)):
            assert np.array_equal(mine, res[k]), k
