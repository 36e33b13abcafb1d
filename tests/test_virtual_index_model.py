"""CPU model of the "virtual index" (csrc/fcol.cu k2_from_dict): on a gapless
MSA whose K-mer buckets each lie in one column, f(x) follows from the sorted list of distinct
K-mers alone -- (column, LCP to the K-mer before) per bucket, runs of neighbouring buckets of one
column, two segmented minima, one maximum per column.  The model restates that computation in
numpy for ANY K and checks it against the oracle's f(x), which walks the full suffix array."""
import numpy as np
import pytest

from oracle import oracle as O


def f_from_kmer_list(msa: np.ndarray, K: int, tricks_off: bool):
    """Returns f, or None when some K-mer occurs in two columns (the GPU path then goes on sorting)."""
    m, n = msa.shape
    T, _, _, _ = O.text(msa)                       # rows, '#' after each, final 0
    N = len(T)
    pad = np.concatenate([T, np.zeros(K, np.uint8)])
    kmers = np.stack([pad[j:j + N] for j in range(K)], axis=1)     # [N, K] bytes, order preserving
    col = np.arange(N) % (n + 1)
    col[N - 1] = n + 1                             # the sentinel
    order = np.lexsort(kmers.T[::-1])
    sk, sc = kmers[order], col[order]
    head = np.ones(N, bool)
    head[1:] = (sk[1:] != sk[:-1]).any(axis=1)
    bucket = np.cumsum(head) - 1
    nd = bucket[-1] + 1
    bcol = sc[head]
    if (sc != bcol[bucket]).any():
        return None
    keys = sk[head]
    H = np.zeros(nd + 1, np.int64)                 # LCP to the bucket before; H[nd] = 0
    for t in range(1, nd):
        d = np.nonzero(keys[t] != keys[t - 1])[0]
        H[t] = d[0]
    f = np.arange(n, dtype=np.int64)
    t = 0
    while t < nd:                                  # runs of neighbouring buckets of one column
        e = t
        while e + 1 < nd and bcol[e + 1] == bcol[t]:
            e += 1
        c = int(bcol[t])
        if c < n and (tricks_off or c > 0):
            for u in range(t, e + 1):
                g = 1 + max(H[t:u + 1].min(), H[u + 1:e + 2].min())
                fi = (n if tricks_off else n - 1) if c + g > n else c + g - 1
                f[c] = max(f[c], fi)
        t = e + 1
    return f.astype(np.uint64)


def gapless_msa(rng, m, n, snp, alphabet=b"ACGT"):
    a = np.frombuffer(alphabet, np.uint8)
    base = a[rng.integers(0, len(a), size=n)]
    rows = np.tile(base, (m, 1))
    hit = rng.random(rows.shape) < snp
    rows[hit] = a[rng.integers(0, len(a), size=int(hit.sum()))]
    return rows


@pytest.mark.parametrize("seed", range(40))
def test_kmer_list_gives_f(seed):
    rng = np.random.default_rng(4000 + seed)
    m, n = int(rng.integers(1, 25)), int(rng.integers(2, 120))
    msa = gapless_msa(rng, m, n, float(rng.choice([0.0, 0.01, 0.05, 0.3])), bytes(rng.choice([b"ACGT", b"ACGT", b"ACDEFGHIKL"])))
    checked = 0
    for K in (3, 5, 8, 13, 21):
        for toff in (False, True):
            f = f_from_kmer_list(msa, K, toff)
            if f is None:
                continue
            assert np.array_equal(f, O.compute_f(msa, b"", toff)), (seed, K, toff)
            checked += 1
    # K = 21 on a short random text: almost always pure
    assert checked > 0 or n < 30


def test_repeat_across_columns_is_rejected_until_k_exceeds_it():
    rng = np.random.default_rng(9)
    msa = gapless_msa(rng, 8, 200, 0.01)
    msa[:, 120:130] = msa[:, 40:50]                # a 10-column repeat
    assert f_from_kmer_list(msa, 8, False) is None
    f = f_from_kmer_list(msa, 21, False)
    assert f is not None and np.array_equal(f, O.compute_f(msa, b"", False))
