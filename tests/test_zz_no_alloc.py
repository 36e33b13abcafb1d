"""GPU (-m gpu): fbg_radix_sort_pairs_noalloc, the entry point for callers that must not allocate
(scratch = a fixed __device__ array inside the library), against the existing fbg_radix_sort_pairs.
Integer work: the comparison is bit-exact.  Collected last on purpose."""
import numpy as np
import pytest

import founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import api

pytestmark = pytest.mark.gpu


def _pairs(seed, count):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 2**63, size=count, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=count, dtype=np.uint64)
    keys[: count // 3] &= np.uint64(0xFF00FF)  # many duplicates: stability matters
    return keys, np.arange(count, dtype=np.uint32)


def _noalloc_sort(ctx, stream, keys, vals, lo, hi, out_k, out_v, measure=False):
    import torch
    dk = torch.from_numpy(keys.view(np.int64)).cuda()
    dv = torch.from_numpy(vals.view(np.int32)).cuda()
    torch.cuda.synchronize()  # the library's stream does not wait for torch's
    free0 = torch.cuda.mem_get_info()[0]
    ctx.radix_sort_pairs_noalloc(dk.data_ptr(), dv.data_ptr(), out_k.data_ptr(), out_v.data_ptr(), len(keys), lo, hi)
    stream.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    if measure:
        assert free1 == free0, "device memory in use changed across a no-allocation call"
    n = len(keys)
    return out_k[:n].cpu().numpy().view(np.uint64), out_v[:n].cpu().numpy().view(np.uint32), dk, dv


def test_noalloc_radix_sort_matches_existing_entry_point(gpu_ctx):
    import torch
    ctx = gpu_ctx
    stream = torch.cuda.ExternalStream(ctx.stream_ptr(), device=torch.device("cuda", torch.cuda.current_device()))
    nmax = api.NOALLOC_MAX_PAIRS
    out_k = torch.full((nmax,), -1, dtype=torch.int64, device="cuda")
    out_v = torch.full((nmax,), -1, dtype=torch.int32, device="cuda")
    # two calls in a row on the shared scratch array: the maximum size with an even number of passes, then a
    # smaller input with an odd number of passes (stale scratch and stale output contents must not matter)
    for i, (count, lo, hi) in enumerate([(nmax, 0, 64), (70_001, 5, 29), (4097, 0, 8)]):
        keys, vals = _pairs(100 + i, count)
        want_k, want_v = ctx.radix_sort_pairs(keys, vals, lo, hi)  # the existing entry point (host buffers, arena scratch)
        got_k, got_v, dk, dv = _noalloc_sort(ctx, stream, keys, vals, lo, hi, out_k, out_v, measure=i > 0)
        assert np.array_equal(got_k, want_k) and np.array_equal(got_v, want_v)
        # the inputs are only read
        assert np.array_equal(dk.cpu().numpy().view(np.uint64), keys) and np.array_equal(dv.cpu().numpy().view(np.uint32), vals)
    # hi_bit == lo_bit: nothing to sort by, the stable order is the input order
    keys, vals = _pairs(7, 1000)
    got_k, got_v, _, _ = _noalloc_sort(ctx, stream, keys, vals, 12, 12, out_k, out_v)
    assert np.array_equal(got_k, keys) and np.array_equal(got_v, vals)


def test_noalloc_rejects_sizes_beyond_its_scratch(gpu_ctx):
    import torch
    a = torch.zeros(16, dtype=torch.int64, device="cuda")
    b = torch.zeros(16, dtype=torch.int64, device="cuda")
    with pytest.raises(pkg.api.FbgError) as e:
        gpu_ctx.radix_sort_pairs_noalloc(a.data_ptr(), a.data_ptr(), b.data_ptr(), b.data_ptr(), api.NOALLOC_MAX_PAIRS + 1, 0, 64)
    assert e.value.code == api.FBG_ERR_UNSUPPORTED and str(api.NOALLOC_MAX_PAIRS) in str(e.value)
    with pytest.raises(pkg.api.FbgError) as e:  # in place is not supported
        gpu_ctx.radix_sort_pairs_noalloc(a.data_ptr(), a.data_ptr(), a.data_ptr(), a.data_ptr(), 16, 0, 64)
    assert e.value.code == api.FBG_ERR_ARG
    gpu_ctx.radix_sort_pairs_noalloc(0, 0, 0, 0, 0, 0, 64)  # empty input: nothing to do
