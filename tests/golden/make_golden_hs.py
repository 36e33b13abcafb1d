"""Generates tests/golden/hs_cases.json: what the REFERENCE (oracle/_ref/founderblockgraph, built from
/root/reference by oracle/Makefile) writes in its hidden --heuristic-subset=ROWNUM mode (SURVEY 8(f) f3:
fbg.cpp:2158-2178 row-chunk worker, 3400-3433 driver loop, 3467-3500 + efg_validate 3199-3292 repair loop).

    python tests/golden/make_golden_hs.py

Inputs: some of the fixtures of make_golden.py plus seeded MSAs with repeats ACROSS row chunks (rows of one
chunk carry a copy of a segment other chunks have elsewhere), which is what makes the per-chunk f too small and
the repair loop merge blocks.  All inputs have m <= 100 rows: the reference's transpose worker never clears its
row buffer after the first 100 rows (fbg.cpp:330-337), so beyond that its output is not meaningful.
Per case: exit code, xGFA bytes (base64) and sha256, the "There are K blocks to remove" sequence, the final
"optimal segment length".
"""
import base64
import hashlib
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

FIX = os.path.join(HERE, "fixtures")
ACGT = np.frombuffer(b"ACGT", np.uint8)


def repeat_msa(rng, m, n, reps, mut=0.03, gap_runs=0, n_chars=0):
    base = ACGT[rng.integers(0, 4, size=n)]
    msa = np.tile(base, (m, 1)).copy()
    mm = rng.random((m, n)) < mut
    msa[mm] = ACGT[rng.integers(0, 4, size=int(mm.sum()))]
    for (r0, r1, src, dst, length) in reps:
        msa[r0:r1, dst:dst + length] = msa[0, src:src + length]  # rows r0..r1 carry a copy of row 0's segment elsewhere
    for _ in range(gap_runs):
        i, x, length = int(rng.integers(0, m)), int(rng.integers(0, n)), int(rng.integers(1, 9))
        msa[i, x:x + length] = 0x2D
    for _ in range(n_chars):
        i, x = int(rng.integers(0, m)), int(rng.integers(0, n))
        if msa[i, x] != 0x2D:
            msa[i, x] = 0x4E
    return msa


def fixtures():
    rng = np.random.default_rng(20261105)
    out = {
        "hs_rep_6x80": repeat_msa(rng, 6, 80, [(3, 6, 10, 50, 12)]),
        "hs_rep_8x120": repeat_msa(rng, 8, 120, [(4, 8, 20, 70, 15), (0, 2, 90, 30, 10)]),
        "hs_rep_10x150": repeat_msa(rng, 10, 150, [(5, 10, 5, 100, 20)]),
        "hs_rep_gaps_9x140": repeat_msa(rng, 9, 140, [(4, 9, 15, 90, 18)], gap_runs=14),
        "hs_rep_gapsN_12x200": repeat_msa(rng, 12, 200, [(6, 12, 30, 140, 22), (0, 3, 160, 60, 12)], gap_runs=20, n_chars=25),
        "hs_rep_40x300": repeat_msa(rng, 40, 300, [(20, 40, 40, 200, 25), (0, 10, 250, 100, 15)], mut=0.01, gap_runs=60),
    }
    for name, msa in out.items():
        O.write_fasta(os.path.join(FIX, name + ".fasta"), [bytes(r) for r in msa])
    return sorted(out) + ["msa", "test2", "rand_gaps_12x120", "rand_gapsN_7x70", "rand_gapless_16x200", "rand_m1_1x30"]


def main():
    O.build()
    assert O.have_reference(), "reference binaries missing: make -C oracle ref"
    cases = []
    for fx in fixtures():
        path = os.path.join(FIX, fx + ".fasta")
        for hs in (1, 2, 3, 5):
            for extra in ([], ["-p"], ["--ignore-chars=N"], ["--threads=2"], ["--disable-elastic-tricks"]):
                if extra == ["--threads=2"] and hs != 2:
                    continue
                flags = ["--elastic", "--gfa", "--heuristic-subset=%d" % hs] + extra
                code, xg, err = O.run_reference(path, flags, timeout=300)
                removed = [int(x) for x in re.findall(r"There are (\d+) blocks to remove", err)]
                lens = [int(x) for x in re.findall(r"optimal segment length = (\d+)", err)]
                cases.append(dict(fixture=fx + ".fasta", flags=flags, hs=hs, exit=code, removed=removed, score=lens[-1] if lens else None,
                                  nosolution="No valid segmentation found!" in err,
                                  xgfa_sha256=hashlib.sha256(xg).hexdigest() if xg is not None and code == 0 else None,
                                  xgfa_b64=base64.b64encode(xg).decode() if xg is not None and code == 0 else None))
    with open(os.path.join(HERE, "hs_cases.json"), "w") as fh:
        json.dump(cases, fh, indent=0)
    fixed = sum(1 for c in cases if sum(c["removed"]) > 0)
    print("wrote %d cases, %d of them with blocks merged by the repair loop, exits %s" % (len(cases), fixed, sorted(set(c["exit"] for c in cases))))


if __name__ == "__main__":
    main()
