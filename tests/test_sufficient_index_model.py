"""CPU model of the "sufficient index" (csrc/suffix.cu): on a gapless MSA the order
of the suffixes INSIDE a bucket whose members lie in one column, and the LCP values inside it,
cannot change f(x).  The model takes the true SA / LCP from the oracle, shuffles the order inside every
column-pure depth-D bucket, replaces the LCP inside by D, and evaluates the K2 formula
(g = 1 + max(min LCP[a..k], min LCP[k+1..b+1]) over runs of consecutive ranks of C_x, fbg.cpp:1633-1672)
on both: the results must agree with each other and with the oracle's f(x)."""
import numpy as np
import pytest

from oracle import oracle as O
from test_virtual_index_model import gapless_msa


def k2_formula(msa, SA, LCP, tricks_off):
    m, n = msa.shape
    N = len(SA)
    ISA = np.empty(N, np.int64)
    ISA[SA] = np.arange(N)
    lcp = np.append(LCP, 0)                                  # LCP[N] = 0
    f = np.arange(n, dtype=np.int64)
    for x in range(n):
        if x == 0 and not tricks_off:
            continue                                         # no row has a character before column 0
        ranks = np.sort(ISA[np.arange(m) * (n + 1) + x])
        a = 0
        while a < m:
            b = a
            while b + 1 < m and ranks[b + 1] == ranks[b] + 1:
                b += 1
            for j in range(a, b + 1):
                left = lcp[ranks[a]:ranks[j] + 1].min()
                right = lcp[ranks[j] + 1:ranks[b] + 2].min()
                g = 1 + max(left, right)
                fi = (n if tricks_off else n - 1) if x + g > n else x + g - 1
                f[x] = max(f[x], fi)
            a = b + 1
    return f.astype(np.uint64)


def scramble_pure_buckets(msa, T, SA, LCP, D, rng):
    """Order inside column-pure depth-D buckets shuffled, LCP inside them = D.  Returns None when
    some tied bucket spans two columns (prefix doubling would go on)."""
    m, n = msa.shape
    N = len(SA)
    SA2, LCP2 = SA.copy(), LCP.copy()
    col = np.arange(N) % (n + 1)
    col[N - 1] = n + 1
    k = 0
    while k < N:
        e = k
        while e + 1 < N and LCP[e + 1] >= D:
            e += 1
        if e > k:
            if len(set(col[SA[k:e + 1]])) != 1:
                return None
            SA2[k:e + 1] = rng.permutation(SA[k:e + 1])
            LCP2[k + 1:e + 1] = D
        k = e + 1
    return SA2, LCP2


@pytest.mark.parametrize("seed", range(25))
def test_order_inside_pure_buckets_is_irrelevant(seed):
    rng = np.random.default_rng(7000 + seed)
    m, n = int(rng.integers(2, 16)), int(rng.integers(5, 90))
    msa = gapless_msa(rng, m, n, float(rng.choice([0.0, 0.02, 0.1])))
    T, _, _, _ = O.text(msa)
    SA, _, LCP = O.sa_lcp(T)
    SA, LCP = SA.astype(np.int64), LCP.astype(np.int64)
    done = 0
    for toff in (False, True):
        f_ref = O.compute_f(msa, b"", toff)
        assert np.array_equal(k2_formula(msa, SA, LCP, toff), f_ref)        # the formula itself
        for D in (4, 8, 21, 42):
            sc = scramble_pure_buckets(msa, T, SA, LCP, D, rng)
            if sc is None:
                continue
            assert np.array_equal(k2_formula(msa, sc[0], sc[1], toff), f_ref), (seed, D, toff)
            done += 1
    assert done > 0
