"""Time-bounded randomised parity run of the CUDA path against the oracle (SA / ISA / LCP, f,
DP tables, boundaries, graph arrays), beyond the fixed seeds of tests/:

    python tools/fuzz_gpu.py [seconds] [first_seed]

Shapes are drawn to hit the bucket-local sorter's corner cases: few founders with sparse
mutations (large buckets that split slowly, small minorities), periodic rows, identical rows,
unrelated rows (dictionary overflow), tiny and one-column inputs, gaps and ignore characters.
Exits 1 on the first mismatch and prints the seed."""
import os
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

import founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import synth
import test_gpu_parity as T


def draw(rng):
    kind = int(rng.integers(0, 9))
    alpha = bytes(rng.choice([b"ACGT", b"ACGT", b"AC", b"ACGTN", b"ACDEFGHIKLMNPQRSTVWY"]))
    if kind == 0:  # MSA-like: one ancestor, sparse SNPs, many rows
        m, n = int(rng.integers(2, 400)), int(rng.integers(50, 4000))
        return synth.make_msa(m, n, snp=float(rng.choice([0.0005, 0.002, 0.01, 0.05])), gap_div=int(rng.choice([0, 200, 1000])),
                              seed=int(rng.integers(1 << 30))), "msa-like"
    if kind == 1:  # few founders, gap runs, N's
        m, n = int(rng.integers(1, 40)), int(rng.integers(1, 1500))
        return synth.random_small_msa(rng, m, n, alphabet=alpha, gap_p=float(rng.choice([0, 0.05, 0.2])), n_p=float(rng.choice([0, 0.03]))), "founders"
    if kind == 2:  # periodic rows with a few mutations: buckets that stay large for many rounds
        m, n = int(rng.integers(2, 30)), int(rng.integers(100, 6000))
        unit = np.frombuffer(alpha, np.uint8)[rng.integers(0, len(alpha), size=int(rng.integers(1, 9)))]
        a = np.tile(np.resize(unit, n), (m, 1)).copy()
        k = int(rng.integers(0, 1 + m * n // 200))
        a[rng.integers(0, m, size=k), rng.integers(0, n, size=k)] = np.frombuffer(alpha, np.uint8)[rng.integers(0, len(alpha), size=k)]
        return a, "periodic"
    if kind == 3:  # identical rows
        m, n = int(rng.integers(2, 60)), int(rng.integers(1, 2500))
        return np.tile(np.frombuffer(alpha, np.uint8)[rng.integers(0, len(alpha), size=n)], (m, 1)).copy(), "identical"
    if kind == 4:  # unrelated rows
        m, n = int(rng.integers(1, 60)), int(rng.integers(1, 6000))
        return np.frombuffer(alpha, np.uint8)[rng.integers(0, len(alpha), size=(m, n))].copy(), "unrelated"
    if kind == 5:  # tiny
        m, n = int(rng.integers(1, 5)), int(rng.integers(1, 12))
        return synth.random_small_msa(rng, m, n, alphabet=alpha, gap_p=float(rng.choice([0, 0.3]))), "tiny"
    if kind == 6:  # many rows, short: buckets around the tile capacities
        m, n = int(rng.integers(500, 2500)), int(rng.integers(30, 200))
        base = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=n)]
        a = np.tile(base, (m, 1)).copy()
        mut = rng.random((m, n)) < float(rng.choice([0.0005, 0.003, 0.02]))
        a[mut] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=int(mut.sum()))]
        return a, "many-rows"
    if kind == 8:  # clades: many rows copied from a few founders with sparse mutations and gap runs -- buckets of 1000+ rows
        # whose minorities are large buckets themselves (the M list of the suffix sorter)
        m, n = int(rng.integers(1100, 5000)), int(rng.integers(40, 400))
        nf = int(rng.integers(1, 5))
        founders = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=(nf, n))]
        if nf > 1 and rng.random() < 0.7:  # founders that differ in a few columns only
            founders[1:] = founders[0]
            for q in range(1, nf):
                cols = rng.integers(0, n, size=int(rng.integers(1, 6)))
                founders[q, cols] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=len(cols))]
        a = founders[rng.integers(0, nf, size=m)].copy()
        mut = rng.random((m, n)) < float(rng.choice([0.0, 0.0003, 0.002, 0.01]))
        a[mut] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=int(mut.sum()))]
        if rng.random() < 0.5:
            for i in rng.integers(0, m, size=int(rng.integers(1, 40))):
                x = int(rng.integers(0, n))
                a[i, x:x + int(rng.integers(1, 12))] = 0x2D
        return a, "clades"
    m, n = int(rng.integers(2, 25)), int(rng.integers(200, 20000))  # long, few rows: many DP chunks / backtrack segments
    return synth.make_msa(m, n, snp=float(rng.choice([0.005, 0.03])), gap_div=int(rng.choice([0, 300])), seed=int(rng.integers(1 << 30))), "long"


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    ctx = pkg.Context(0)
    t0 = time.time()
    done = {}
    while time.time() - t0 < budget:
        rng = np.random.default_rng(seed)
        msa, kind = draw(rng)
        try:
            T._check_index(ctx, msa)
            gapless = not (msa == 45).any()
            mode = int(rng.integers(0, 3))
            if mode == 0 or not gapless:
                ign = b"N" if (msa == 78).any() and rng.random() < 0.5 else b""
                T._check_elastic(ctx, msa, ignore=ign, tricks_off=bool(rng.random() < 0.2))
            else:
                T._check_nonelastic(ctx, msa, 1 if mode == 1 else int(rng.integers(2, 6)))
        except Exception:
            print("MISMATCH seed %d kind %s shape %s" % (seed, kind, msa.shape), flush=True)
            traceback.print_exc()
            sys.exit(1)
        done[kind] = done.get(kind, 0) + 1
        if ctx.stats()["index_depth"]:
            done["(sufficient index)"] = done.get("(sufficient index)", 0) + 1
        seed += 1
    n_lazy = done.pop("(sufficient index)", 0)
    print("fuzz ok: %d cases in %.0f s (%d of them through the sufficient index), seeds up to %d, by kind %s" % (
        sum(done.values()), time.time() - t0, n_lazy, seed - 1, done), flush=True)


if __name__ == "__main__":
    main()
