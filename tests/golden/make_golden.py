"""Generates tests/golden/* by running the REFERENCE (oracle/_ref/founderblockgraph and
oracle/_ref/ref_harness, built from /root/reference by oracle/Makefile) in this container.

    python tests/golden/make_golden.py

Outputs (committed):
  fixtures/*.fasta          copies of the reference's own test inputs (test/*.fasta) plus
                            seeded random MSAs covering gaps, N runs, duplicate rows, m=1, n=1
  cases.json                for every (fixture, flag set): reference exit code, xGFA sha256 and
                            bytes (base64), result-carrying stderr lines, and the harness dumps
                            f / mml / backtrack / boundaries (elastic), v / s / prev / boundaries
                            + statistics (non-elastic)
"""
import base64
import hashlib
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from founderblockgraphs_b200 import synth  # noqa: E402

FIX = os.path.join(HERE, "fixtures")


def stderr_results(err: str):
    keep = ("Input MSA", "MSA contains", "Computing optimal", "Optimal score", "Number of segments", "#nodes=", "total length",
            "#founders=", "#edges=", "No proper", "No valid", "Unable", "WARNING", "NOTICE", "--elastic and", "Output of original", "Invalid number")
    return [l for l in err.splitlines() if l.startswith(keep)]


def main():
    O.build()
    assert O.have_reference(), "reference binaries missing: make -C oracle ref"
    os.makedirs(FIX, exist_ok=True)
    for name in ("msa.fasta", "test.fasta", "test2.fasta", "test3.fasta"):
        shutil.copyfile(os.path.join("/root/reference/test", name), os.path.join(FIX, name))
    rng = np.random.default_rng(20261104)
    rand = {}
    specs = [("rand_gapless_8x60", 8, 60, 0.0, 0.0), ("rand_gaps_6x50", 6, 50, 0.2, 0.0), ("rand_gapsN_7x70", 7, 70, 0.15, 0.06),
             ("rand_dups_5x40", 5, 40, 0.0, 0.0), ("rand_m1_1x30", 1, 30, 0.1, 0.0), ("rand_n1_4x1", 4, 1, 0.0, 0.0),
             ("rand_gaps_12x120", 12, 120, 0.1, 0.03), ("rand_gapless_16x200", 16, 200, 0.0, 0.0)]
    for name, m, n, gp, npb in specs:
        msa = synth.random_small_msa(rng, m, n, gap_p=gp, n_p=npb, alphabet=b"ACGT" if "dups" not in name else b"AC", div=0.3)
        if "dups" in name:
            msa[1] = msa[0]
            msa[3] = msa[2]
        if name == "rand_n1_4x1":
            msa[:] = np.frombuffer(b"ACAG", np.uint8).reshape(4, 1)
        path = os.path.join(FIX, name + ".fasta")
        O.write_fasta(path, [bytes(r) for r in msa], width=(23 if "12x120" in name else 0))
        rand[name] = path

    cases = []
    fixtures = sorted(f for f in os.listdir(FIX) if f.endswith(".fasta"))
    for fx in fixtures:
        path = os.path.join(FIX, fx)
        flagsets = [["--elastic", "--gfa"], ["--elastic", "--gfa", "-p"], ["--elastic", "--gfa", "--disable-elastic-tricks"],
                    ["--elastic", "--gfa", "--ignore-chars=N"], ["--elastic", "--gfa", "--threads=3"],
                    ["--gap-limit=1"], ["--gap-limit=0"], ["--gap-limit=3"], ["--gfa"], ["--elastic"]]
        for flags in flagsets:
            code, xg, err = O.run_reference(path, flags)
            case = dict(fixture=fx, flags=flags, exit=code, stderr=stderr_results(err),
                        xgfa_sha256=hashlib.sha256(xg).hexdigest() if xg is not None else None,
                        xgfa_b64=base64.b64encode(xg).decode() if xg is not None else None)
            elastic = "--elastic" in flags and "--gfa" in flags
            if elastic and "--threads=3" not in flags:
                hargs = []
                if "--disable-elastic-tricks" in flags:
                    hargs.append("--disable-elastic-tricks")
                if "--ignore-chars=N" in flags:
                    hargs += ["--ignore-chars", "N"]
                res, _, _ = O.run_harness(path, "elastic", hargs)
                for k in ("f", "mml", "backtrack", "boundaries"):
                    if k in res:
                        case[k] = [int(v) for v in res[k]]
            if flags[0].startswith("--gap-limit"):
                gl = flags[0].split("=")[1]
                res, herr, _ = O.run_harness(path, "nonelastic", ["--gap-limit", gl])
                for k in ("v", "s", "prev", "boundaries"):
                    if k in res:
                        case[k] = [int(v) for v in res[k]]
                case["harness_status"] = res.get("status")
                case["harness_m"] = res.get("m")
            cases.append(case)
    with open(os.path.join(HERE, "cases.json"), "w") as fh:
        json.dump(cases, fh, indent=0)
    print("wrote", len(cases), "cases over", len(fixtures), "fixtures")


if __name__ == "__main__":
    main()
