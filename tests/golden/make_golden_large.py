"""Generates tests/golden/large_cases.json: sha256 digests of what the REFERENCE (oracle/_ref,
built from /root/reference by oracle/Makefile) produces on medium and benchmark-sized inputs --
the sizes and index paths bench.py times ("pin parity at the sizes that are
benchmarked").  Only digests, sizes and times are committed; the inputs are regenerated from
founderblockgraphs_b200/synth.py (deterministic seeds).

    python tests/golden/make_golden_large.py [case ...]        # resumable: finished cases are kept

Per case: xGFA sha256 + length (the reference's own output_efg), sha256 over the u64
little-endian dumps of f / mml / backtrack / boundaries (elastic) or v / s / prev / boundaries
(non-elastic), the result-carrying stderr lines, and the reference's wall-clock per phase with
the thread count used (the full-size CPU figures BASELINE.md quotes).
"""
import hashlib
import json
import os
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from founderblockgraphs_b200 import synth  # noqa: E402

OUT = os.path.join(HERE, "large_cases.json")

# name -> how the MSA is made + how the reference is run.  gen = ("config", cfg, scale, seed) or
# ("msa", m, n, snp, gap_div, n_div, seed).  bench.py's workload generator is
# synth.make_msa(cfg..., seed = cfg number + 100 * rank + 7 * variant): the cfg*_full_v* cases
# below are exactly the MSAs bench.py builds on rank 0.
CASES = {
    "gapped_100x20000_p": dict(gen=("msa", 100, 20_000, 0.001, 4000, 0, 11), mode="elastic", paths=True, ignore=""),
    "gapped_200x50000_p": dict(gen=("msa", 200, 50_000, 0.001, 10_000, 0, 12), mode="elastic", paths=True, ignore=""),
    "gapless_300x20000_elastic": dict(gen=("msa", 300, 20_000, 0.002, 0, 0, 13), mode="elastic", paths=False, ignore=""),
    "cfg2_full_nonelastic_v0": dict(gen=("config", "cfg2", 1.0, 2), mode="nonelastic", paths=False, ignore=""),
    "cfg2_full_nonelastic_v1": dict(gen=("config", "cfg2", 1.0, 9), mode="nonelastic", paths=False, ignore=""),
    "cfg2_full_elastic_v0": dict(gen=("config", "cfg2", 1.0, 2), mode="elastic", paths=False, ignore=""),
    "cfg3_twin_0.02": dict(gen=("config", "cfg3", 0.02, 3), mode="elastic", paths=False, ignore=""),
    "cfg3_twin_0.1": dict(gen=("config", "cfg3", 0.1, 3), mode="elastic", paths=False, ignore=""),
    "cfg4_twin_0.01_N": dict(gen=("config", "cfg4", 0.01, 4), mode="elastic", paths=False, ignore="N"),
    "cfg4_twin_0.01_N_p": dict(gen=("config", "cfg4", 0.01, 4), mode="elastic", paths=True, ignore="N"),
    "cfg5_twin_0.01": dict(gen=("config", "cfg5", 0.01, 5), mode="elastic", paths=False, ignore=""),
    "cfg3_full_v0": dict(gen=("config", "cfg3", 1.0, 3), mode="elastic", paths=False, ignore=""),
    "cfg3_full_v1": dict(gen=("config", "cfg3", 1.0, 10), mode="elastic", paths=False, ignore=""),
}


def make_case_msa(gen):
    if gen[0] == "config":
        msa, _ = synth.make_config(gen[1], gen[2], seed=gen[3])
        return msa
    _, m, n, snp, gap_div, n_div, seed = gen
    return synth.make_msa(m, n, snp, gap_div, n_div, seed=seed)


def digest_u64(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<u8").tobytes()).hexdigest()


def stderr_results(err):
    keep = ("Input MSA", "MSA contains", "Computing optimal", "Optimal score", "Number of segments", "#nodes=", "total length",
            "#founders=", "#edges=", "No proper", "No valid")
    return [l for l in err.splitlines() if l.startswith(keep)]


def run_case(name, spec, threads):
    msa = make_case_msa(spec["gen"])
    m, n = msa.shape
    rec = dict(name=name, gen=list(spec["gen"]), m=int(m), n=int(n), mode=spec["mode"], paths=spec["paths"], ignore=spec["ignore"],
               msa_sha256=hashlib.sha256(msa.tobytes()).hexdigest())
    with tempfile.TemporaryDirectory(prefix="fbggold", dir=os.environ.get("FBG_GOLDEN_TMP", "/tmp")) as td:
        path = os.path.join(td, "in.fasta")
        O.write_fasta(path, [bytes(r) for r in msa])
        t0 = time.time()
        if spec["mode"] == "elastic":
            args = ["--threads", str(threads)]
            if spec["ignore"]:
                args += ["--ignore-chars", spec["ignore"]]
            if spec["paths"]:
                args += ["--output-paths"]
            res, err, xg = O.run_harness(path, "elastic", args, timeout=7200, want_output=True)
            rec["threads"] = threads
            for k in ("f", "mml", "backtrack", "boundaries"):
                rec[k + "_sha256"] = digest_u64(res[k])
            rec["n_blocks"] = int(len(res["boundaries"]))
            rec["score"] = int(res["mml"][-1])
            rec["xgfa_sha256"] = hashlib.sha256(xg).hexdigest()
            rec["xgfa_bytes"] = len(xg)
            rec["xgfa_by"] = "reference output_efg (fbg.cpp:1185-1301) through oracle/_ref/ref_harness"
        else:
            res, err, _ = O.run_harness(path, "nonelastic", ["--gap-limit", "1"], timeout=7200)
            rec["threads"] = 1  # segment() is single-threaded (fbg.cpp:526-735)
            for k in ("v", "s", "prev", "boundaries"):
                rec[k + "_sha256"] = digest_u64(res[k])
            b = np.asarray(res["boundaries"], dtype=np.uint64)
            rec["n_blocks"] = int(len(b))
            rec["score"] = int(res["s"][-1])
            # the reference crashes before it writes a non-elastic file: the xGFA digest is
            # the C restatement's writer (pinned on the 120 small goldens) on the REFERENCE's boundaries
            xp = os.path.join(td, "o.xgfa")
            O.write_xgfa(xp, msa, b)
            xg = open(xp, "rb").read()
            rec["xgfa_sha256"] = hashlib.sha256(xg).hexdigest()
            rec["xgfa_bytes"] = len(xg)
            rec["xgfa_by"] = "oracle writer on the reference's boundaries (the reference writes no non-elastic file, SURVEY F3)"
        rec["status"] = int(res.get("status", -1))
        rec["stderr"] = stderr_results(err)
        rec["ref_time_index_s"] = res.get("time_index_s")
        rec["ref_time_segment_s"] = res.get("time_segment_s")
        rec["ref_time_write_s"] = res.get("time_write_s", 0.0)
        rec["ref_wall_s"] = time.time() - t0
        rec["ref_host"] = "%d cores, build container" % (os.cpu_count() or 1)
    return rec


def main():
    O.build()
    assert O.have_reference(), "reference binaries missing: make -C oracle ref"
    done = {}
    if os.path.exists(OUT):
        with open(OUT) as fh:
            done = {r["name"]: r for r in json.load(fh)}
    want = sys.argv[1:] or list(CASES)
    threads = os.cpu_count() or 1
    for name in want:
        if name in done and not os.environ.get("FBG_GOLDEN_FORCE"):
            continue
        t0 = time.time()
        done[name] = run_case(name, CASES[name], threads)
        print("%-28s %7.1f s  blocks %d score %d xgfa %d B" % (name, time.time() - t0, done[name]["n_blocks"], done[name]["score"], done[name]["xgfa_bytes"]), flush=True)
        with open(OUT, "w") as fh:
            json.dump([done[k] for k in CASES if k in done], fh, indent=1)


if __name__ == "__main__":
    main()
