"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY.

ctypes front-end to oracle/_ref/libfbg_oracle.so (the plain-C CPU restatement in
oracle/fbg_oracle.c) plus runners for the reference binaries built by
oracle/Makefile (oracle/_ref/founderblockgraph, oracle/_ref/ref_harness).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module; the product package
(founderblockgraphs_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
LIB_PATH = os.path.join(REF_DIR, "libfbg_oracle.so")
REF_BIN = os.path.join(REF_DIR, "founderblockgraph")
HARNESS_BIN = os.path.join(REF_DIR, "ref_harness")
INVALID = np.uint64(0xFFFFFFFFFFFFFFFF)

_u8p = C.POINTER(C.c_uint8)
_u64p = C.POINTER(C.c_uint64)
_i64p = C.POINTER(C.c_int64)
_lib = None


def build(force: bool = False) -> None:
    """Compile the C restatement (and, when /root/reference exists, the reference)."""
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(HERE, "fbg_oracle.c")):
        subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference") and (force or not (os.path.exists(REF_BIN) and os.path.exists(HARNESS_BIN))):
        subprocess.check_call(["make", "-C", HERE, "ref", "-j8"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB_PATH)
        _lib.fbgo_text.restype = C.c_uint64
        _lib.fbgo_text.argtypes = [_u8p, C.c_uint64, C.c_uint64, _u8p, _u64p, _u64p, _u64p]
        _lib.fbgo_suffix_array.argtypes = [_u8p, C.c_uint64, _i64p, _i64p]
        _lib.fbgo_lcp.argtypes = [_u8p, C.c_uint64, _i64p, _i64p, _i64p]
        _lib.fbgo_f.argtypes = [_u8p, C.c_uint64, C.c_uint64, _u8p, C.c_uint64, C.c_int, _u64p]
        _lib.fbgo_counts.argtypes = [_u8p, C.c_uint64, C.c_uint64, _u8p, C.c_uint64, _u64p, _u64p]
        _lib.fbgo_dp_elastic.argtypes = [_u64p, C.c_uint64, _u64p, _u64p, _u64p, _u64p]
        _lib.fbgo_v.argtypes = [_u64p, C.c_uint64, _u64p]
        _lib.fbgo_dp_nonelastic.argtypes = [_u64p, C.c_uint64, C.c_int, _u64p, _u64p, _u64p, _u64p, _u64p]
        _lib.fbgo_graph.argtypes = [_u8p, C.c_uint64, C.c_uint64, _u64p, C.c_uint64] + [_u64p] * 9
        _lib.fbgo_write_xgfa.argtypes = [C.c_char_p, _u8p, C.c_uint64, C.c_uint64, _u64p, C.c_uint64, C.c_int, C.POINTER(C.c_char_p)]
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


# ---------------------------------------------------------------- FASTA (A1)

def read_fasta(path: str, gap_limit: int = 1, elastic: bool = True, output_paths: bool = False):
    """Restatement of read_input / check_sequence_length / check_gaps (fbg.cpp:103-201).

    Returns (rows, identifiers, stderr_lines). Bytes are kept verbatim (no CR stripping,
    Q5); the first line is assumed to be a header (fbg.cpp:160)."""
    with open(path, "rb") as fh:
        lines = fh.read().split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()  # std::getline does not yield a trailing empty line
    rows, ids, msgs = [], [], []
    if not lines:
        return rows, ids, msgs
    identifier = lines[0]
    if output_paths:
        ids.append(identifier[1:])
    entry = b""
    expected = None

    def keep(identifier, entry, expected):
        if len(entry) != expected:
            msgs.append("WARNING: length of the sequence “%s” does not match that of the first sequence; skipping. (%d vs. %d)"
                        % (identifier[1:].decode("latin1"), expected, len(entry)))
            return False
        if not elastic and gap_limit != 0:
            run = best = 0
            for c in entry:
                if c == 0x2D:
                    run += 1
                else:
                    best = max(best, run)
                    run = 0
            best = max(best, run)
            if best >= gap_limit:
                msgs.append("NOTICE: Sequence “%s” contained a gap run with %d characters." % (identifier[1:].decode("latin1"), best))
                return False
        return True

    for line in lines[1:]:
        if line[:1] == b">":
            if output_paths:
                ids.append(line[1:])
            if expected is None:
                expected = len(entry)
            if keep(identifier, entry, expected):
                rows.append(entry)
            entry = b""
            identifier = line
        else:
            entry += line
    if expected is None:
        expected = len(entry)
    if keep(identifier, entry, expected):
        rows.append(entry)
    return rows, ids, msgs


def msa_array(rows) -> np.ndarray:
    m, n = len(rows), len(rows[0])
    return np.frombuffer(b"".join(rows), dtype=np.uint8).reshape(m, n).copy()


# ------------------------------------------------------------ C restatement

def text(msa: np.ndarray):
    m, n = msa.shape
    cap = m * n + m + 1
    T = np.zeros(cap, np.uint8)
    start = np.zeros(m, np.uint64)
    rowlen = np.zeros(m, np.uint64)
    colof = np.zeros(cap, np.uint64)
    N = lib().fbgo_text(_p(np.ascontiguousarray(msa), _u8p), m, n, _p(T, _u8p), _p(start, _u64p), _p(rowlen, _u64p), _p(colof, _u64p))
    return T[:N], start, rowlen, colof[:N]


def sa_lcp(T: np.ndarray):
    N = len(T)
    T = np.ascontiguousarray(T)
    SA = np.zeros(N, np.int64)
    ISA = np.zeros(N, np.int64)
    LCP = np.zeros(N, np.int64)
    assert lib().fbgo_suffix_array(_p(T, _u8p), N, _p(SA, _i64p), _p(ISA, _i64p)) == 0
    lib().fbgo_lcp(_p(T, _u8p), N, _p(SA, _i64p), _p(ISA, _i64p), _p(LCP, _i64p))
    return SA, ISA, LCP


def compute_f(msa: np.ndarray, ignore: bytes = b"", disable_tricks: bool = False) -> np.ndarray:
    m, n = msa.shape
    f = np.zeros(n, np.uint64)
    ign = np.frombuffer(ignore, np.uint8).copy() if ignore else np.zeros(1, np.uint8)
    rc = lib().fbgo_f(_p(np.ascontiguousarray(msa), _u8p), m, n, _p(ign, _u8p), len(ignore), int(disable_tricks), _p(f, _u64p))
    assert rc == 0
    return f


def counts(msa: np.ndarray, ignore: bytes = b""):
    m, n = msa.shape
    g, t = C.c_uint64(), C.c_uint64()
    ign = np.frombuffer(ignore, np.uint8).copy() if ignore else np.zeros(1, np.uint8)
    lib().fbgo_counts(_p(np.ascontiguousarray(msa), _u8p), m, n, _p(ign, _u8p), len(ignore), C.byref(g), C.byref(t))
    return g.value, t.value


def dp_elastic(f: np.ndarray):
    n = len(f)
    f = np.ascontiguousarray(f, np.uint64)
    mml = np.zeros(n + 1, np.uint64)
    bt = np.zeros(n + 1, np.uint64)
    b = np.zeros(n + 2, np.uint64)
    nb = C.c_uint64()
    lib().fbgo_dp_elastic(_p(f, _u64p), n, _p(mml, _u64p), _p(bt, _u64p), _p(b, _u64p), C.byref(nb))
    return mml, bt, b[: nb.value].copy()


def v_from_f(f_notricks: np.ndarray) -> np.ndarray:
    n = len(f_notricks)
    f = np.ascontiguousarray(f_notricks, np.uint64)
    v = np.zeros(n, np.uint64)
    lib().fbgo_v(_p(f, _u64p), n, _p(v, _u64p))
    return v


def dp_nonelastic(v: np.ndarray, greedy: bool = False):
    """Returns (failed, s, prev, boundaries, score)."""
    n = len(v)
    v = np.ascontiguousarray(v, np.uint64)
    s = np.zeros(n, np.uint64)
    prev = np.zeros(n, np.uint64)
    b = np.zeros(n + 1, np.uint64)
    nb, score = C.c_uint64(), C.c_uint64()
    rc = lib().fbgo_dp_nonelastic(_p(v, _u64p), n, int(greedy), _p(s, _u64p), _p(prev, _u64p), _p(b, _u64p), C.byref(nb), C.byref(score))
    return bool(rc), s, prev, (b[: nb.value].copy() if rc == 0 else None), score.value


def graph(msa: np.ndarray, boundaries: np.ndarray) -> dict:
    m, n = msa.shape
    bnd = np.ascontiguousarray(boundaries, np.uint64)
    nb = len(bnd)
    cap = m * nb + 1
    node_of = np.zeros(cap, np.uint64)
    heights = np.zeros(nb + 1, np.uint64)
    node_block = np.zeros(cap, np.uint64)
    node_row = np.zeros(cap, np.uint64)
    es = np.zeros(cap, np.uint64)
    ed = np.zeros(cap, np.uint64)
    nn, ne, lt = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().fbgo_graph(_p(np.ascontiguousarray(msa), _u8p), m, n, _p(bnd, _u64p), nb, _p(node_of, _u64p), _p(heights, _u64p),
                     _p(node_block, _u64p), _p(node_row, _u64p), C.byref(nn), _p(es, _u64p), _p(ed, _u64p), C.byref(ne), C.byref(lt))
    return dict(node_of=node_of[: m * nb].reshape(nb, m).copy(), heights=heights[:nb].copy(),
                node_block=node_block[: nn.value].copy(), node_row=node_row[: nn.value].copy(),
                edge_src=es[: ne.value].copy(), edge_dst=ed[: ne.value].copy(), label_total=lt.value)


# --------------------------------------------------- --heuristic-subset

def efg_validate(msa: np.ndarray, boundaries: np.ndarray, ignore: bytes = b""):
    """efg_validate / efg_validate_node (fbg.cpp:3104-3292) on the graph of `boundaries`: a node that is neither
    source nor sink and holds no ignore character is INVALID when its label occurs in some edge string
    label(u) + label(v) anywhere but at the very start of u or of v with that node in the same block.  Returns the
    set of boundary indices to drop (block b invalid -> boundary b - 1, fbg.cpp:3270-3271).  Plain string search:
    small inputs only."""
    g = graph(msa, boundaries)
    b = [int(x) for x in boundaries]
    starts = [0] + [x + 1 for x in b[:-1]]
    n = msa.shape[1]

    def label(v):
        j, i = int(g["node_block"][v]), int(g["node_row"][v])
        seg = msa[i, starts[j]: min(b[j], n - 1) + 1]
        return bytes(seg[seg != 0x2D])

    nn = len(g["node_block"])
    labels = [label(v) for v in range(nn)]
    has_in, has_out = [False] * nn, [False] * nn
    edges = list(zip((int(x) for x in g["edge_src"]), (int(x) for x in g["edge_dst"])))
    for u, v in edges:
        has_out[u] = True
        has_in[v] = True
    drop = set()
    for w in range(nn):
        if not has_in[w] or not has_out[w]:   # is_source / is_sink (fbg.cpp:3116-3117)
            continue
        lab = labels[w]
        if ignore and any(c in lab for c in ignore):  # fbg.cpp:3119-3120
            continue
        bw = int(g["node_block"][w])
        bad = False
        for u, v in edges:
            s = labels[u] + labels[v]
            pos = s.find(lab)
            while pos >= 0 and not bad:
                if pos < len(labels[u]):
                    ok = pos == 0 and int(g["node_block"][u]) == bw
                else:
                    ok = pos == len(labels[u]) and int(g["node_block"][v]) == bw
                bad = not ok
                pos = s.find(lab, pos + 1)
            if bad:
                break
        if bad and bw > 0:
            drop.add(bw - 1)
    return drop


def heuristic_subset(msa: np.ndarray, rownum: int, ignore: bytes = b"", disable_tricks: bool = False):
    """The reference's --heuristic-subset=ROWNUM pipeline (fbg.cpp:3400-3433, 2158-2178, 3467-3500): f = max over
    chunks of ROWNUM consecutive rows of the chunk's own f, one DP on that f, then blocks are merged with their
    predecessor until every node passes efg_validate.  Returns (boundaries, [blocks removed per iteration], score,
    no_solution)."""
    m, n = msa.shape
    f = np.zeros(n, np.uint64)
    for r0 in range(0, m, rownum):
        f = np.maximum(f, compute_f(np.ascontiguousarray(msa[r0: r0 + rownum]), ignore, disable_tricks))
    if disable_tricks and int(f[0]) == n:     # fbg.cpp:2048-2053
        return None, [], None, True
    mml, _, b = dp_elastic(f)
    removed = []
    while True:
        drop = efg_validate(msa, b, ignore)
        removed.append(len(drop))
        if not drop:
            break
        b = np.array([x for i, x in enumerate(b) if i not in drop], np.uint64)
    return b, removed, int(mml[-1]), False


def write_xgfa(path: str, msa: np.ndarray, boundaries: np.ndarray, ids=None) -> None:
    m, n = msa.shape
    bnd = np.ascontiguousarray(boundaries, np.uint64)
    arr = (C.c_char_p * m)(*[bytes(i) for i in ids]) if ids is not None else None
    rc = lib().fbgo_write_xgfa(path.encode(), _p(np.ascontiguousarray(msa), _u8p), m, n, _p(bnd, _u64p), len(bnd), int(ids is not None), arr)
    assert rc == 0


# ------------------------------------------------------- reference runners

def have_reference() -> bool:
    return os.path.exists(REF_BIN) and os.path.exists(HARNESS_BIN)


def run_reference(fasta_path: str, args, timeout: float = 600.0):
    """Run the unmodified reference CLI on a scratch copy of the input (it writes
    <input>.plain and <input>.plain<gl>.cst next to the input, fbg.cpp:363-365).
    Returns (exit_code, xgfa_bytes_or_None, stderr_text)."""
    with tempfile.TemporaryDirectory(prefix="fbgref") as td:
        inp = os.path.join(td, "in.fasta")
        out = os.path.join(td, "out.xgfa")
        shutil.copyfile(fasta_path, inp)
        pr = subprocess.run([REF_BIN, "--input", inp, "--output", out] + list(args), capture_output=True, timeout=timeout, cwd=td)  # sdsl drops its cache files in the cwd
        data = open(out, "rb").read() if os.path.exists(out) else None
        return pr.returncode, data, pr.stderr.decode("utf8", "replace")


def run_harness(fasta_path: str, mode: str, args=(), timeout: float = 600.0, want_output: bool = False):
    """Run oracle/_ref/ref_harness; returns (dict of dumped vectors/scalars, stderr, xgfa bytes or None)."""
    with tempfile.TemporaryDirectory(prefix="fbgharn") as td:
        inp = os.path.join(td, "in.fasta")
        out = os.path.join(td, "out.xgfa")
        shutil.copyfile(fasta_path, inp)
        cmd = [HARNESS_BIN, "--mode", mode, "--input", inp] + list(args)
        if want_output:
            cmd += ["--output", out]
        pr = subprocess.run(cmd, capture_output=True, timeout=timeout, cwd=td)  # sdsl drops its cache files in the cwd
        res = {"returncode": pr.returncode}
        for line in pr.stdout.decode().splitlines():
            parts = line.split()
            if not parts:
                continue
            key, vals = parts[0], parts[1:]
            if key.startswith("time_"):
                res[key] = float(vals[0])
            elif key in ("m", "n", "status"):
                res[key] = int(vals[0])
            else:
                res[key] = np.array([int(v) for v in vals], dtype=np.uint64)
        data = open(out, "rb").read() if os.path.exists(out) else None
        return res, pr.stderr.decode("utf8", "replace"), data


def write_fasta(path: str, rows, names=None, width: int = 0) -> None:
    with open(path, "wb") as fh:
        for i, r in enumerate(rows):
            r = bytes(r)
            fh.write(b">" + (names[i] if names else b"h%d" % i) + b"\n")
            if width:
                for o in range(0, len(r), width):
                    fh.write(r[o:o + width] + b"\n")
            else:
                fh.write(r + b"\n")
