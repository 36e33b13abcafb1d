"""ctypes binding of libfbg_cuda.so (include/fbg_cuda.h) plus a host-side mirror of
the reference's function-level interface for the hot path.

The reference exposes the path only through its CLI; inside, main() calls
read_input -> load_cst -> segment_elastic_minmaxlength / segment /
segment2elasticValid -> output_efg (fbg.cpp:3349-3503).  The functions below keep
those names and argument meanings so the parity tests read like calls into the
reference; every one of them runs on the GPU through the C ABI.  There is no CPU
path: without the built library or without a B200 they raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libfbg_cuda.so")
CLI_PATH = os.path.join(_PKG, "founderblockgraph")

FBG_OK, FBG_ERR_CUDA, FBG_ERR_ARG, FBG_ERR_NOSEG, FBG_ERR_UNSUPPORTED = 0, -1, -2, -3, -4
FBG_NO_NODE = 0xFFFFFFFF
N_KERNELS = 14

_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)


class FbgGraph(C.Structure):
    _fields_ = [("n_blocks", C.c_uint64), ("n_nodes", C.c_uint64), ("n_edges", C.c_uint64), ("n_rows", C.c_uint64),
                ("label_total", C.c_uint64), ("max_height", C.c_uint64),
                ("block_height", _u64p), ("node_block", _u64p), ("node_row", _u64p),
                ("edge_src", _u64p), ("edge_dst", _u64p), ("node_of", _u32p), ("opaque", C.c_void_p)]


class FbgMsa(C.Structure):
    _fields_ = [("rows", _u8p), ("m", C.c_uint64), ("n", C.c_uint64), ("ids", C.POINTER(C.c_char_p)), ("n_ids", C.c_uint64),
                ("messages", C.c_char_p), ("opaque", C.c_void_p)]


class FbgStats(C.Structure):
    _fields_ = [("m", C.c_uint64), ("n", C.c_uint64), ("text_len", C.c_uint64), ("gaps", C.c_uint64), ("to_ignore", C.c_uint64),
                ("sa_rounds", C.c_uint64), ("arena_peak_bytes", C.c_uint64),
                ("kernel_ms", C.c_double * N_KERNELS), ("kernel_bytes", C.c_double * N_KERNELS), ("kernel_launches", C.c_uint64 * N_KERNELS),
                ("dp_passes", C.c_int64), ("kmer_dict", C.c_int64), ("index_depth", C.c_int64), ("index_virtual", C.c_int64)]


# every symbol include/fbg_cuda.h declares (checked by tests/test_cabi.py)
EXPORTS = [
    "fbg_kernel_name", "fbg_ctx_create", "fbg_ctx_destroy", "fbg_last_error", "fbg_set_options", "fbg_upload_msa",
    "fbg_upload_msa_device", "fbg_build_index", "fbg_compute_f", "fbg_segment", "fbg_get_boundaries", "fbg_build_graph",
    "fbg_graph_free", "fbg_run", "fbg_run_device", "fbg_get_stream", "fbg_get_stats", "fbg_reset_stats", "fbg_get_f", "fbg_get_text",
    "fbg_get_sa", "fbg_get_isa", "fbg_get_lcp", "fbg_get_dp", "fbg_radix_sort_pairs", "fbg_radix_sort_bench", "fbg_write_xgfa",
    "fbg_check_index", "fbg_set_dp_chunking", "fbg_dist_unique_id", "fbg_dist_init", "fbg_dist_shutdown", "fbg_plan_key_ranges", "fbg_plan_columns",
    "fbg_format_xgfa", "fbg_read_xgfa", "fbg_save_xgfa", "fbg_read_fasta", "fbg_msa_free", "fbg_radix_sort_pairs_noalloc", "fbg_device_count", "fbg_heuristic_f", "fbg_validate_blocks",
]
NOALLOC_MAX_PAIRS = 1 << 20  # FBG_NOALLOC_MAX_PAIRS

_lib = None


class FbgError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libfbg_cuda: %s (code %d)" % (msg, code))
        self.code = code


class NoSegmentation(FbgError):
    pass


def load_library():
    """Loads the in-tree CUDA library; fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(make -C founderblockgraphs_b200/csrc). There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.fbg_kernel_name.restype = C.c_char_p
    lib.fbg_kernel_name.argtypes = [C.c_int]
    lib.fbg_ctx_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    lib.fbg_ctx_destroy.argtypes = [C.c_void_p]
    lib.fbg_ctx_destroy.restype = None
    lib.fbg_last_error.restype = C.c_char_p
    lib.fbg_last_error.argtypes = [C.c_void_p]
    lib.fbg_set_options.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_char_p, C.c_size_t, C.c_int]
    lib.fbg_upload_msa.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64]
    lib.fbg_upload_msa_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64]
    lib.fbg_build_index.argtypes = [C.c_void_p]
    lib.fbg_compute_f.argtypes = [C.c_void_p]
    lib.fbg_segment.argtypes = [C.c_void_p, _u64p, _u64p]
    lib.fbg_get_boundaries.argtypes = [C.c_void_p, _u64p]
    lib.fbg_build_graph.argtypes = [C.c_void_p, C.c_int, C.POINTER(FbgGraph)]
    lib.fbg_graph_free.argtypes = [C.POINTER(FbgGraph)]
    lib.fbg_graph_free.restype = None
    lib.fbg_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, _u64p, _u64p, C.POINTER(FbgGraph)]
    lib.fbg_run_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, _u64p, _u64p, C.POINTER(FbgGraph)]
    lib.fbg_get_stream.argtypes = [C.c_void_p]
    lib.fbg_get_stream.restype = C.c_void_p
    lib.fbg_get_stats.argtypes = [C.c_void_p, C.POINTER(FbgStats)]
    lib.fbg_reset_stats.argtypes = [C.c_void_p]
    lib.fbg_reset_stats.restype = None
    for name in ("fbg_get_f", "fbg_get_sa", "fbg_get_isa", "fbg_get_lcp"):
        getattr(lib, name).argtypes = [C.c_void_p, _u64p]
    lib.fbg_get_text.argtypes = [C.c_void_p, _u8p]
    lib.fbg_get_dp.argtypes = [C.c_void_p, _u64p, _u64p, _u64p]
    lib.fbg_radix_sort_pairs.argtypes = [C.c_void_p, _u64p, _u32p, C.c_uint64, C.c_int, C.c_int]
    lib.fbg_radix_sort_bench.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.fbg_radix_sort_pairs_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int]
    lib.fbg_heuristic_f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, _u64p, _u64p]
    lib.fbg_validate_blocks.argtypes = [C.c_void_p, _u64p, _u64p, C.POINTER(FbgGraph)]
    lib.fbg_write_xgfa.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint64, _u64p, C.POINTER(FbgGraph), C.c_int, C.POINTER(C.c_char_p)]
    lib.fbg_format_xgfa.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(FbgGraph), _u64p]
    lib.fbg_read_xgfa.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]
    lib.fbg_save_xgfa.argtypes = [C.c_void_p, C.c_char_p]
    lib.fbg_read_fasta.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.POINTER(FbgMsa)]
    lib.fbg_msa_free.argtypes = [C.POINTER(FbgMsa)]
    lib.fbg_msa_free.restype = None
    lib.fbg_set_dp_chunking.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
    lib.fbg_check_index.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, _u32p]
    lib.fbg_dist_unique_id.argtypes = [_u8p]
    lib.fbg_dist_init.argtypes = [C.c_void_p, C.c_int, C.c_int, _u8p]
    lib.fbg_dist_shutdown.argtypes = [C.c_void_p]
    lib.fbg_plan_key_ranges.argtypes = [_u32p, C.c_uint32, C.c_int, _u32p]
    lib.fbg_plan_columns.argtypes = [C.c_uint64, C.c_int, C.c_int, _u64p, _u64p]
    _lib = lib
    return lib


def dist_unique_id() -> bytes:
    """ncclGetUniqueId through the library (rank 0); ship the bytes to the other ranks."""
    lib = load_library()
    buf = (C.c_uint8 * 128)()
    rc = lib.fbg_dist_unique_id(buf)
    if rc != FBG_OK:
        raise FbgError(rc, lib.fbg_last_error(None).decode())
    return bytes(buf)


def plan_key_ranges(hist: np.ndarray, world: int) -> np.ndarray:
    """Which k-mer histogram bins each rank sorts (host plan of the sharded K1)."""
    lib = load_library()
    h = np.ascontiguousarray(hist, np.uint32)
    out = np.zeros(world + 1, np.uint32)
    rc = lib.fbg_plan_key_ranges(h.ctypes.data_as(_u32p), len(h), world, out.ctypes.data_as(_u32p))
    if rc != FBG_OK:
        raise FbgError(rc, "bad arguments")
    return out


def plan_columns(n: int, world: int, rank: int):
    """Column range [x0, x1) of f(x) that `rank` computes in a sharded build."""
    lib = load_library()
    x0, x1 = C.c_uint64(), C.c_uint64()
    rc = lib.fbg_plan_columns(n, world, rank, C.byref(x0), C.byref(x1))
    if rc != FBG_OK:
        raise FbgError(rc, "bad arguments")
    return int(x0.value), int(x1.value)


def kernel_names():
    lib = load_library()
    return [lib.fbg_kernel_name(i).decode() for i in range(N_KERNELS)]


class Graph:
    """Host copy of an fbg_graph (arrays copied out, library memory released)."""

    def __init__(self, g: FbgGraph, keep_struct: bool = False):
        nb, nn, ne, m = g.n_blocks, g.n_nodes, g.n_edges, g.n_rows
        self.n_blocks, self.n_nodes, self.n_edges, self.n_rows = nb, nn, ne, m
        self.label_total, self.max_height = g.label_total, g.max_height

        def arr(ptr, cnt, dt):
            if not cnt or not ptr:
                return np.zeros(0, dt)
            return np.ctypeslib.as_array(ptr, shape=(cnt,)).astype(dt, copy=True)

        self.block_height = arr(g.block_height, nb, np.uint64)
        self.node_block = arr(g.node_block, nn, np.uint64)
        self.node_row = arr(g.node_row, nn, np.uint64)
        self.edge_src = arr(g.edge_src, ne, np.uint64)
        self.edge_dst = arr(g.edge_dst, ne, np.uint64)
        self.node_of = arr(g.node_of, nb * m, np.uint32).reshape(nb, m) if g.node_of else None


class Context:
    """One fbg_ctx (one GPU, one stream)."""

    def __init__(self, device: int = -1):
        self.lib = load_library()
        self.h = C.c_void_p()
        rc = self.lib.fbg_ctx_create(C.byref(self.h), device)
        if rc != FBG_OK:
            raise FbgError(rc, self.lib.fbg_last_error(None).decode())
        self._graph = None

    def dist_init(self, rank: int, world: int, unique_id: bytes):
        """Joins the NCCL communicator of a sharded build (collective over all ranks)."""
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(self.lib.fbg_dist_init(self.h, rank, world, buf))

    def dist_shutdown(self):
        self._check(self.lib.fbg_dist_shutdown(self.h))

    def close(self):
        if self.h:
            self.free_graph()
            self.lib.fbg_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc == FBG_OK:
            return
        msg = self.lib.fbg_last_error(self.h).decode()
        if rc == FBG_ERR_NOSEG:
            raise NoSegmentation(rc, msg)
        raise FbgError(rc, msg)

    def set_options(self, elastic=True, gap_limit=1, ignore_chars: bytes = b"", disable_tricks=False):
        self._check(self.lib.fbg_set_options(self.h, int(elastic), int(gap_limit), ignore_chars, len(ignore_chars), int(disable_tricks)))

    def upload(self, msa: np.ndarray):
        msa = np.ascontiguousarray(msa, dtype=np.uint8)
        self._msa = msa
        self._check(self.lib.fbg_upload_msa(self.h, msa.ctypes.data, msa.shape[0], msa.shape[1]))

    def upload_ptr(self, host_ptr: int, m: int, n: int):
        self._check(self.lib.fbg_upload_msa(self.h, host_ptr, m, n))

    def upload_device(self, dptr: int, m: int, n: int):
        self._check(self.lib.fbg_upload_msa_device(self.h, dptr, m, n))

    def build_index(self):
        self._check(self.lib.fbg_build_index(self.h))

    def compute_f(self) -> np.ndarray:
        self._check(self.lib.fbg_compute_f(self.h))
        return self.get_f()

    def get_f(self) -> np.ndarray:
        """f of the last compute_f / run, read back from the device."""
        st = self.stats()
        out = np.zeros(st["n"], np.uint64)
        self._check(self.lib.fbg_get_f(self.h, out.ctypes.data_as(_u64p)))
        return out

    def segment(self):
        nb, sc = C.c_uint64(), C.c_uint64()
        rc = self.lib.fbg_segment(self.h, C.byref(nb), C.byref(sc))
        self.score = sc.value
        self._check(rc)
        out = np.zeros(nb.value, np.uint64)
        self._check(self.lib.fbg_get_boundaries(self.h, out.ctypes.data_as(_u64p)))
        return out, sc.value

    def free_graph(self):
        if self._graph is not None:
            self.lib.fbg_graph_free(C.byref(self._graph))
            self._graph = None

    def build_graph(self, want_paths=False) -> Graph:
        self.free_graph()
        g = FbgGraph()
        self._check(self.lib.fbg_build_graph(self.h, int(want_paths), C.byref(g)))
        self._graph = g
        return Graph(g)

    def run_ptr(self, ptr: int, m: int, n: int, device: bool, want_paths=False):
        """Whole path in one C call; returns (n_blocks, score). The graph stays in the library."""
        self.free_graph()
        g = FbgGraph()
        nb, sc = C.c_uint64(), C.c_uint64()
        fn = self.lib.fbg_run_device if device else self.lib.fbg_run
        self._check(fn(self.h, ptr, m, n, int(want_paths), C.byref(nb), C.byref(sc), C.byref(g)))
        self._graph = g
        return nb.value, sc.value

    def run_xgfa_ptr(self, ptr: int, m: int, n: int, device: bool, ids=None, host_out: int = 0, host_cap: int = 0):
        """Whole path, MSA -> xGFA bytes: upload (host or device buffer), index, f, DP, graph and the
        bytes of the file formatted on the device (K5).  host_out: address of a host buffer that
        receives the bytes (end-to-end use); 0 leaves them in device memory.
        Returns (n_blocks, score, counts)."""
        self.free_graph()
        self._check((self.lib.fbg_upload_msa_device if device else self.lib.fbg_upload_msa)(self.h, ptr, m, n))
        nb, sc = C.c_uint64(), C.c_uint64()
        self._check(self.lib.fbg_segment(self.h, C.byref(nb), C.byref(sc)))
        info = self.format_xgfa(ids)
        if host_out:
            if info["n_bytes"] > host_cap:
                raise FbgError(FBG_ERR_ARG, "host buffer too small for the xGFA")
            self._check(self.lib.fbg_read_xgfa(self.h, 0, info["n_bytes"], host_out))
        return nb.value, sc.value, info

    def write_xgfa(self, path: str, msa: np.ndarray, boundaries: np.ndarray, ids=None):
        assert self._graph is not None
        msa = np.ascontiguousarray(msa, dtype=np.uint8)
        b = np.ascontiguousarray(boundaries, np.uint64)
        arr = (C.c_char_p * msa.shape[0])(*[bytes(i) for i in ids]) if ids is not None else None
        rc = self.lib.fbg_write_xgfa(path.encode(), msa.ctypes.data, msa.shape[0], msa.shape[1], b.ctypes.data_as(_u64p),
                                     C.byref(self._graph), int(ids is not None), arr)
        if rc != FBG_OK:
            raise FbgError(rc, "cannot write " + path)

    def format_xgfa(self, ids=None) -> dict:
        """K4 + K5: graph and xGFA bytes formatted on the device (nothing but counts comes back)."""
        g = FbgGraph()
        nbytes = C.c_uint64()
        arr = (C.c_char_p * len(ids))(*[bytes(i) for i in ids]) if ids is not None else None
        self._check(self.lib.fbg_format_xgfa(self.h, int(ids is not None), arr, C.byref(g), C.byref(nbytes)))
        self._xgfa_len = nbytes.value
        return dict(n_blocks=g.n_blocks, n_nodes=g.n_nodes, n_edges=g.n_edges, label_total=g.label_total,
                    max_height=g.max_height, n_bytes=nbytes.value)

    def read_xgfa(self, offset: int = 0, length: int | None = None) -> bytes:
        """Bytes [offset, offset + length) of the xGFA formatted by format_xgfa."""
        if length is None:
            length = self._xgfa_len - offset
        buf = np.zeros(max(length, 1), np.uint8)
        self._check(self.lib.fbg_read_xgfa(self.h, offset, length, buf.ctypes.data))
        return buf[:length].tobytes()

    def xgfa_bytes(self, ids=None) -> bytes:
        info = self.format_xgfa(ids)
        return self.read_xgfa(0, info["n_bytes"])

    def save_xgfa(self, path: str):
        self._check(self.lib.fbg_save_xgfa(self.h, path.encode()))

    def heuristic_subset(self, msa: np.ndarray, rownum: int):
        """--heuristic-subset=ROWNUM through the C ABI: f = max over row chunks (fbg_heuristic_f), one DP (fbg_segment), then
        the repair loop (fbg_validate_blocks until nothing is removed).  Options as set by set_options (elastic).
        Returns (boundaries, [boundaries removed per pass], score, f)."""
        msa = np.ascontiguousarray(msa, dtype=np.uint8)
        self._msa = msa
        self.free_graph()
        self._check(self.lib.fbg_heuristic_f(self.h, msa.ctypes.data, msa.shape[0], msa.shape[1], rownum, None, None))
        f = self.get_f()
        b, score = self.segment()
        removed = []
        while True:
            nr, nbl = C.c_uint64(), C.c_uint64()
            self._check(self.lib.fbg_validate_blocks(self.h, C.byref(nr), C.byref(nbl), None))
            removed.append(int(nr.value))
            if nr.value == 0:
                break
        out = np.zeros(nbl.value, np.uint64)
        self._check(self.lib.fbg_get_boundaries(self.h, out.ctypes.data_as(_u64p)))
        return out, removed, score, f

    def stream_ptr(self) -> int:
        return int(self.lib.fbg_get_stream(self.h) or 0)

    def stats(self) -> dict:
        st = FbgStats()
        self._check(self.lib.fbg_get_stats(self.h, C.byref(st)))
        names = kernel_names()
        return dict(m=st.m, n=st.n, text_len=st.text_len, gaps=st.gaps, to_ignore=st.to_ignore, sa_rounds=st.sa_rounds,
                    arena_peak_bytes=st.arena_peak_bytes, dp_passes=int(st.dp_passes), kmer_dict=int(st.kmer_dict), index_depth=int(st.index_depth), index_virtual=int(st.index_virtual),
                    kernels={names[i]: dict(ms=st.kernel_ms[i], bytes=st.kernel_bytes[i], launches=st.kernel_launches[i]) for i in range(N_KERNELS)})

    def reset_stats(self):
        self.lib.fbg_reset_stats(self.h)

    def _get_u64(self, fn, count):
        out = np.zeros(count, np.uint64)
        self._check(fn(self.h, out.ctypes.data_as(_u64p)))
        return out

    def text(self) -> np.ndarray:
        N = self.stats()["text_len"]
        out = np.zeros(N, np.uint8)
        self._check(self.lib.fbg_get_text(self.h, out.ctypes.data_as(_u8p)))
        return out

    def sa(self):
        return self._get_u64(self.lib.fbg_get_sa, self.stats()["text_len"])

    def isa(self):
        return self._get_u64(self.lib.fbg_get_isa, self.stats()["text_len"])

    def lcp(self):
        return self._get_u64(self.lib.fbg_get_lcp, self.stats()["text_len"])

    def dp(self, elastic: bool):
        n = self.stats()["n"]
        cnt = n + 1 if elastic else n
        a, b = np.zeros(cnt, np.uint64), np.zeros(cnt, np.uint64)
        v = np.zeros(n, np.uint64)
        self._check(self.lib.fbg_get_dp(self.h, a.ctypes.data_as(_u64p), b.ctypes.data_as(_u64p), None if elastic else v.ctypes.data_as(_u64p)))
        return (a, b) if elastic else (a, b, v)

    def set_dp_chunking(self, chunk: int = 0, warmup: int = 0):
        """Chunk / warm-up columns of the guess-and-verify DP (0, 0 = defaults); results never depend on it."""
        self._check(self.lib.fbg_set_dp_chunking(self.h, chunk, warmup))

    def check_index(self, nsamples: int = 100000, seed: int = 1) -> int:
        """Device-side spot check of SA/ISA/LCP by direct text comparison; returns the number of bad samples."""
        bad = C.c_uint32()
        self._check(self.lib.fbg_check_index(self.h, nsamples, seed, C.byref(bad)))
        return int(bad.value)

    def radix_sort_pairs(self, keys: np.ndarray, vals: np.ndarray, lo_bit=0, hi_bit=64):
        keys = np.ascontiguousarray(keys, np.uint64).copy()
        vals = np.ascontiguousarray(vals, np.uint32).copy()
        self._check(self.lib.fbg_radix_sort_pairs(self.h, keys.ctypes.data_as(_u64p), vals.ctypes.data_as(_u32p), len(keys), lo_bit, hi_bit))
        return keys, vals

    def radix_sort_pairs_noalloc(self, d_keys_in: int, d_vals_in: int, d_keys_out: int, d_vals_out: int, count: int, lo_bit=0, hi_bit=64):
        """The radix passes on DEVICE buffers (addresses) without any allocation: scratch is a fixed array inside the
        library, at most NOALLOC_MAX_PAIRS pairs.  Enqueued on the context's stream; the caller synchronises it."""
        self._check(self.lib.fbg_radix_sort_pairs_noalloc(self.h, d_keys_in, d_vals_in, d_keys_out, d_vals_out, count, lo_bit, hi_bit))

    def radix_sort_bench(self, count: int, lo_bit=0, hi_bit=64, reps=3) -> float:
        ms = C.c_double()
        self._check(self.lib.fbg_radix_sort_bench(self.h, count, lo_bit, hi_bit, reps, C.byref(ms)))
        return ms.value


# --------------------------------------------------------------------------
# Mirror of the reference's function-level interface (names as in fbg.cpp)

def read_input(input_path: str, gap_limit: int, elastic: bool, output_paths: bool = False):
    """read_input (fbg.cpp:151-201): returns (msa uint8[m,n], identifiers). Host code."""
    with open(input_path, "rb") as fh:
        lines = fh.read().split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    rows, ids = [], []
    if not lines:
        return np.zeros((0, 0), np.uint8), ids
    identifier, entry, expected = lines[0], [], None
    if output_paths:
        ids.append(identifier[1:])

    def keep(seq):
        if len(seq) != expected:
            return False
        if elastic or gap_limit == 0:
            return True
        run = best = 0
        for c in seq:
            if c == 0x2D:
                run += 1
            else:
                best, run = max(best, run), 0
        return max(best, run) < gap_limit

    for line in lines[1:]:
        if line[:1] == b">":
            if output_paths:
                ids.append(line[1:])
            seq = b"".join(entry)
            if expected is None:
                expected = len(seq)
            if keep(seq):
                rows.append(seq)
            entry, identifier = [], line
        else:
            entry.append(line)
    seq = b"".join(entry)
    if expected is None:
        expected = len(seq)
    if keep(seq):
        rows.append(seq)
    if not rows or expected == 0:
        return np.zeros((0, 0), np.uint8), ids
    return np.frombuffer(b"".join(rows), np.uint8).reshape(len(rows), expected).copy(), ids


def read_fasta(input_path: str, gap_limit: int, elastic: bool, output_paths: bool = False, threads: int = 0):
    """The library's own reader (host C++, several threads; what the executable uses): returns
    (msa uint8[m,n], identifiers, stderr lines of the skipped records)."""
    lib = load_library()
    out = FbgMsa()
    rc = lib.fbg_read_fasta(input_path.encode(), gap_limit, int(elastic), int(output_paths), threads, C.byref(out))
    if rc != FBG_OK:
        raise FbgError(rc, "cannot read " + input_path)
    try:
        m, n = out.m, out.n
        msa = np.ctypeslib.as_array(out.rows, shape=(m * n,)).reshape(m, n).copy() if m and n else np.zeros((m, n), np.uint8)
        ids = [out.ids[i] for i in range(out.n_ids)]
        msgs = out.messages.decode("utf-8", "replace").split("\n")[:-1]
    finally:
        lib.fbg_msa_free(C.byref(out))
    return msa, ids, msgs


def segment_elastic_minmaxlength(msa: np.ndarray, ignorechars: bytes = b"", disable_efg_tricks: bool = False, ctx: Context | None = None):
    """segment_elastic_minmaxlength (fbg.cpp:1836-2040): returns (out_indices, f, optimal_length)."""
    c = ctx or Context()
    c.set_options(elastic=True, gap_limit=1, ignore_chars=ignorechars, disable_tricks=disable_efg_tricks)
    c.upload(msa)
    f = c.compute_f()
    b, score = c.segment()
    return b, f, score


def segment(msa: np.ndarray, ctx: Context | None = None):
    """segment (fbg.cpp:526-735), gap_limit == 1: returns (boundaries, score, graph statistics)."""
    c = ctx or Context()
    c.set_options(elastic=False, gap_limit=1)
    c.upload(msa)
    b, score = c.segment()
    g = c.build_graph()
    return b, score, g


def segment2elasticValid(msa: np.ndarray, gap_limit: int = 0, ctx: Context | None = None):
    """segment2elasticValid (fbg.cpp:738-935), gap_limit != 1."""
    c = ctx or Context()
    c.set_options(elastic=False, gap_limit=gap_limit)
    c.upload(msa)
    b, score = c.segment()
    g = c.build_graph()
    return b, score, g


def output_efg(boundaries, msa: np.ndarray, output_paths: bool, identifiers, dst_path: str, ctx: Context, host_writer: bool = False):
    """output_efg (fbg.cpp:1185-1301) for the segmentation held by ctx: formatted on the device
    (K5) and streamed to dst_path, or -- host_writer -- graph arrays copied out and formatted by
    the host writer.  Same bytes."""
    if host_writer:
        ctx.build_graph(want_paths=output_paths)
        ctx.write_xgfa(dst_path, msa, boundaries, identifiers if output_paths else None)
    else:
        ctx.format_xgfa(identifiers if output_paths else None)
        ctx.save_xgfa(dst_path)


def main(argv, **kw):
    """Runs the compiled drop-in CLI (same flags as the reference executable)."""
    if not os.path.exists(CLI_PATH):
        raise ImportError(CLI_PATH + " is missing: build with make -C founderblockgraphs_b200/csrc")
    # the executable is a build artefact (not tracked): refuse one that predates its sources or the ABI header
    for src in (os.path.join(_PKG, "csrc", "host", "cli.cpp"), os.path.join(os.path.dirname(_PKG), "include", "fbg_cuda.h")):
        if os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(CLI_PATH):
            raise ImportError(CLI_PATH + " is older than " + src + ": rebuild with make -C founderblockgraphs_b200/csrc")
    return subprocess.run([CLI_PATH] + list(argv), **kw)
