"""Deterministic synthetic MSAs of the BASELINE.json shapes.

Alphabet ACGT; ancestral sequence uniform i.i.d.; each row = ancestor with
independent per-row SNPs (uniform substitution to a different base); optional
indel gap runs ('-') and 'N' runs.  Host-side numpy only; nothing here is on the
measured path.
"""
from __future__ import annotations

import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)

# name -> (m, n, snp, gap_lambda_div, n_lambda_div, elastic, ignore)
CONFIGS = {
    # cfg2: 1,000 x 30 kbp SARS-CoV-2-like, 0.1 % SNP, gapless, non-elastic
    "cfg2": dict(m=1000, n=30_000, snp=0.001, gap_div=0, n_div=0, elastic=False, ignore=b""),
    # cfg3: 5,000 x 100 kbp with indel gap runs, --elastic
    "cfg3": dict(m=5000, n=100_000, snp=0.001, gap_div=2000, n_div=0, elastic=True, ignore=b""),
    # cfg4: 500 x 5 Mbp bacterial pangenome, --elastic --ignore-chars=N
    "cfg4": dict(m=500, n=5_000_000, snp=0.005, gap_div=5000, n_div=1_000_000, elastic=True, ignore=b"N"),
    # cfg5: 100 x 50 Mbp chromosome scale, --elastic
    "cfg5": dict(m=100, n=50_000_000, snp=0.001, gap_div=10_000, n_div=0, elastic=True, ignore=b""),
}


def make_msa(m: int, n: int, snp: float = 0.001, gap_div: int = 0, n_div: int = 0,
             seed: int = 0, gap_mean: float = 10.0, gap_cap: int = 200) -> np.ndarray:
    """Returns an (m, n) uint8 array of ASCII bytes."""
    rng = np.random.Generator(np.random.PCG64(0xF0B60000 + seed))
    anc = rng.integers(0, 4, size=n, dtype=np.uint8)
    msa = np.empty((m, n), dtype=np.uint8)
    for i in range(m):
        row = anc.copy()
        k = rng.binomial(n, snp)
        if k:
            pos = rng.integers(0, n, size=k)
            row[pos] = (row[pos] + rng.integers(1, 4, size=k, dtype=np.uint8)) & 3
        row = _ACGT[row]
        if gap_div:
            runs = rng.poisson(n / gap_div)
            if runs:
                st = rng.integers(0, n, size=runs)
                ln = np.minimum(rng.geometric(1.0 / gap_mean, size=runs), gap_cap)
                for s, l in zip(st, ln):
                    row[s:s + l] = 0x2D
        if n_div:
            runs = rng.poisson(n / n_div)
            if runs:
                st = rng.integers(0, n, size=runs)
                ln = rng.integers(100, 1001, size=runs)
                for s, l in zip(st, ln):
                    seg = row[s:s + l]
                    seg[seg != 0x2D] = 0x4E
        msa[i] = row
    return msa


def make_config(name: str, scale: float = 1.0, seed: int | None = None) -> tuple[np.ndarray, dict]:
    """MSA for a named BASELINE config; `scale` < 1 shrinks the column count (CI twins)."""
    c = dict(CONFIGS[name])
    n = max(16, int(c["n"] * scale))
    msa = make_msa(c["m"], n, c["snp"], c["gap_div"], c["n_div"], seed=int(name[3:]) if seed is None else seed)
    c["n"] = n
    return msa, c


def random_small_msa(rng: np.random.Generator, m: int, n: int, alphabet: bytes = b"ACGT",
                     gap_p: float = 0.15, div: float = 0.2, n_p: float = 0.0) -> np.ndarray:
    """Small adversarial MSAs for parity tests: few founders, gap runs, optional N's."""
    alpha = np.frombuffer(alphabet, dtype=np.uint8)
    nf = max(1, int(rng.integers(1, max(2, m // 2 + 1))))
    founders = alpha[rng.integers(0, len(alpha), size=(nf, n))]
    msa = founders[rng.integers(0, nf, size=m)].copy()
    mut = rng.random((m, n)) < div * rng.random()
    msa[mut] = alpha[rng.integers(0, len(alpha), size=int(mut.sum()))]
    if gap_p > 0:
        for i in range(m):
            x = 0
            while x < n:
                if rng.random() < gap_p:
                    l = int(rng.integers(1, max(2, n // 3)))
                    msa[i, x:x + l] = 0x2D
                    x += l
                x += int(rng.integers(1, 6))
    if n_p > 0:
        nm = (rng.random((m, n)) < n_p) & (msa != 0x2D)
        msa[nm] = 0x4E
    return msa
