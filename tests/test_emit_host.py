"""CPU run of the device-side xGFA writer's per-item code (csrc/emit_items.cuh, the code the
K5 kernels execute) against the oracle's writer: same item lengths, offsets and bytes."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from founderblockgraphs_b200 import synth
from oracle import oracle as O
from util import FIXTURES, ROOT

SRC = os.path.join(ROOT, "tests", "host", "emit_host.cpp")
HDR = os.path.join(ROOT, "founderblockgraphs_b200", "csrc", "emit_items.cuh")


@pytest.fixture(scope="module")
def emit_lib(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("emit") / "libemit_host.so")
    subprocess.check_call([os.environ.get("CXX", "g++"), "-O1", "-std=c++17", "-fPIC", "-shared", "-x", "c++", SRC, "-o", so])
    lib = C.CDLL(so)
    lib.emit_host_format.restype = C.c_uint64
    lib.emit_host_format.argtypes = [C.c_uint64] * 5 + [C.c_void_p] * 11 + [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_uint64]
    return lib


def records(msa, bnd):
    """What K0/K4 hold on the device: text, per-(block,row) label position and length."""
    m, n = msa.shape
    T, start, rowlen, _ = O.text(msa)
    nongap = (msa != 45)
    rank = np.zeros((m, n + 1), np.int64)
    rank[:, 1:] = np.cumsum(nongap, axis=1)
    nb = len(bnd)
    LEN = np.zeros((nb, m), np.uint32)
    POS = np.zeros((nb, m), np.uint32)
    a = 0
    for j, e in enumerate(bnd):
        b1 = min(int(e) + 1, n)
        LEN[j] = rank[:, b1] - rank[:, a]
        POS[j] = start.astype(np.int64) + rank[:, a]
        a = int(e) + 1
    return T, rowlen.astype(np.uint32), LEN, POS


def device_style_bytes(lib, msa, bnd, ids, lanes=32, reverse=False):
    m, n = msa.shape
    g = O.graph(msa, bnd)
    T, rowlen, LEN, POS = records(msa, bnd)
    nb, nn, ne = len(bnd), len(g["node_block"]), len(g["edge_src"])
    u32 = lambda x: np.ascontiguousarray(x, np.uint32)
    node_of = g["node_of"].copy()
    node_of[node_of == O.INVALID] = 0xFFFFFFFF
    arrs = [u32(bnd), u32(np.append(g["heights"], 0)), u32(g["node_block"]), u32(g["node_row"]), u32(g["edge_src"]), u32(g["edge_dst"]),
            u32(node_of), u32(LEN), u32(POS), u32(rowlen), np.ascontiguousarray(np.append(T, np.zeros(8, np.uint8)))]
    if ids is not None:
        idoff = u32(np.concatenate([[0], np.cumsum([len(i) for i in ids])]))
        idbytes = np.frombuffer(b"".join(ids) + b"\0", np.uint8).copy()
    else:
        idoff, idbytes = u32([0]), np.zeros(1, np.uint8)
    cap = 64 + 24 * (2 * nb + 2 * nn + 3 * ne) + int(LEN.sum()) + (m * (nb * 12 + 16) + int(idoff[-1]) if ids is not None else 0)
    out = np.zeros(cap, np.uint8)
    ptr = lambda x: x.ctypes.data
    nbytes = lib.emit_host_format(m, n, nb, nn, ne, *[ptr(x) for x in arrs], int(ids is not None), ptr(idoff), ptr(idbytes), lanes, int(reverse), ptr(out), cap)
    assert nbytes != 2 ** 64 - 1, "an item was given no length"
    assert nbytes <= cap
    return out[:nbytes].tobytes()


def oracle_bytes(tmp_path, msa, bnd, ids):
    p = str(tmp_path / "o.xgfa")
    O.write_xgfa(p, msa, bnd, ids)
    return open(p, "rb").read()


def elastic_boundaries(msa, ignore=b""):
    return O.dp_elastic(O.compute_f(msa, ignore, False))[2]


@pytest.mark.parametrize("seed", range(30))
def test_item_code_random(emit_lib, seed, tmp_path):
    rng = np.random.default_rng(900 + seed)
    m, n = int(rng.integers(1, 30)), int(rng.integers(1, 160))
    msa = synth.random_small_msa(rng, m, n, gap_p=float(rng.choice([0, 0.1, 0.4])), n_p=0.0, alphabet=bytes(rng.choice([b"AC", b"ACGT"])))
    if (msa == 45).all(axis=1).any():
        msa[(msa == 45).all(axis=1), 0] = 65  # the oracle's writer mirrors the reference's UB for all-gap rows
    bnd = elastic_boundaries(msa)
    ids = [b"row%d" % (i * 37) for i in range(m)] if seed % 3 else [b"" if i % 2 else b"a long header with spaces %d" % i for i in range(m)]
    for use_ids in (None, ids):
        ref = oracle_bytes(tmp_path, msa, bnd, use_ids)
        assert device_style_bytes(emit_lib, msa, bnd, use_ids, lanes=32, reverse=bool(seed & 1)) == ref
        assert device_style_bytes(emit_lib, msa, bnd, use_ids, lanes=1, reverse=not (seed & 1)) == ref


def test_item_code_arbitrary_blocks(emit_lib, tmp_path):
    """Any tiling of the columns is a valid input of the writer: single block, one block per column,
    blocks of all-gap columns (height 0), many-digit ids."""
    rng = np.random.default_rng(5)
    msa = synth.random_small_msa(rng, 12, 300, gap_p=0.2, n_p=0.0, alphabet=b"ACGT")
    msa[:, 100:110] = 45  # columns without a single character
    msa[:, 0] = 65
    n = msa.shape[1]
    ids = [b"s%d" % i for i in range(12)]
    for bnd in ([n], list(range(0, n - 1)) + [n], [99, 109, 200, n], [4, 9, 99, 104, 109, n - 1]):
        bnd = np.array(bnd, np.uint64)
        ref = oracle_bytes(tmp_path, msa, bnd, ids)
        assert device_style_bytes(emit_lib, msa, bnd, ids) == ref
        assert device_style_bytes(emit_lib, msa, bnd, None, reverse=True) == oracle_bytes(tmp_path, msa, bnd, None)


def test_item_code_reference_fixtures(emit_lib, tmp_path):
    from util import elastic_cases, flags_to_opts
    import founderblockgraphs_b200 as pkg
    done = 0
    for case in elastic_cases():
        if case["exit"] != 0 or "boundaries" not in case:
            continue
        o = flags_to_opts(case["flags"])
        msa, ids = pkg.read_input(os.path.join(FIXTURES, case["fixture"]), 1, True, o["paths"])
        if o["paths"] and len(ids) != msa.shape[0]:
            continue
        bnd = np.array(case["boundaries"], np.uint64)
        assert device_style_bytes(emit_lib, msa, bnd, ids if o["paths"] else None) == case["xgfa"], case["fixture"]
        done += 1
    assert done >= 10
