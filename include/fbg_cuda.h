/* fbg_cuda.h -- C ABI of libfbg_cuda.so: the B200 (sm_100a) implementation of the
 * founderblockgraphs hot path (semi-repeat-free segmentation + founder-block
 * construction).
 *
 * The reference (algbio/founderblockgraphs @5c852c2) has no FFI of its own: the
 * whole path lives in one anonymous namespace of founderblockgraph.cpp ("fbg.cpp")
 * behind the `founderblockgraph` executable.  Each entry point
 * below therefore replaces the call main() makes at the cited line; the new host
 * executable (csrc/host/cli.cpp, same flags) and any other binding (ctypes, cgo,
 * JNI) call exactly these.  See INTEGRATION.md for the reference-side stub.
 *
 * Conventions: plain pointers and sizes, caller-owned host buffers unless stated,
 * no exceptions across the boundary, int status (0 = FBG_OK, < 0 = error with text
 * in fbg_last_error()).  One context per host thread; calls on one context are
 * serialised by the caller.  There is NO CPU fallback: every compute entry point
 * fails with FBG_ERR_CUDA when no sm_100 device is usable.  A call that returns an
 * error leaves the context usable: its device work has been waited for and its
 * scratch memory released; what the failed stage was producing is absent, the
 * uploaded MSA and the options stay (tests/test_gpu_parity.py,
 * test_context_usable_after_a_failed_stage).
 */
#ifndef FBG_CUDA_H
#define FBG_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FBG_OK 0
#define FBG_ERR_CUDA (-1)       /* CUDA runtime / device failure */
#define FBG_ERR_ARG (-2)        /* bad argument or call order */
#define FBG_ERR_NOSEG (-3)      /* "No proper segmentation exists." / "No valid segmentation found!" */
#define FBG_ERR_UNSUPPORTED (-4)  /* a size this build does not handle; fbg_last_error() says which */

/* Limits of this build, checked by fbg_upload_msa before any work is done (FBG_ERR_UNSUPPORTED):
 * rows (the per-column sorts of K2 / K4 run in shared memory; the reference has no row limit,
 * fbg.cpp:1610-1693), and text length N = non-gap cells + rows + 1 < 2^32 - 16 (32-bit suffix
 * indices; the reference switches to divsufsort64, construct_sa.hpp:183-190). */
#define FBG_MAX_ROWS 8192

#define FBG_NO_NODE UINT32_MAX  /* node_of[] entry of a row whose label in that block is empty */

typedef struct fbg_ctx fbg_ctx;

/* Result of fbg_build_graph; arrays are owned by the library until fbg_graph_free.  They live in
 * one page-locked host block (the device copies run at full PCIe speed); fbg_graph_free hands the
 * block back to a small process-wide pool, so a caller that builds graph after graph causes no
 * page-locking and no first-touch page faults in steady state. */
typedef struct fbg_graph {
	uint64_t n_blocks, n_nodes, n_edges, n_rows;
	uint64_t label_total;        /* "total length of node labels=" (fbg.cpp:695) */
	uint64_t max_height;         /* "#founders=" (fbg.cpp:701) */
	const uint64_t *block_height;/* [n_blocks] distinct non-empty labels: the B line (fbg.cpp:1210-1220) */
	const uint64_t *node_block;  /* [n_nodes] */
	const uint64_t *node_row;    /* [n_nodes] first-occurrence row; the host slices the label from its MSA copy */
	const uint64_t *edge_src;    /* [n_edges] grouped by destination block, ascending (src,dst) inside (fbg.cpp:1253) */
	const uint64_t *edge_dst;
	const uint32_t *node_of;     /* [n_blocks][n_rows] or NULL; FBG_NO_NODE = empty label (P lines, fbg.cpp:1263-1300) */
	void *opaque;
} fbg_graph;

/* Per-kernel accounting for the roofline: accumulated CUDA-event milliseconds,
 * algorithmic bytes (DESIGN.md states the per-unit figures) and launch counts. */
#define FBG_N_KERNELS 14
typedef struct fbg_stats {
	uint64_t m, n, text_len;      /* N = non-gap cells + m + 1 */
	uint64_t gaps, to_ignore;     /* fbg.cpp:1858-1859 */
	uint64_t sa_rounds;           /* prefix-doubling rounds incl. the k-mer round */
	uint64_t arena_peak_bytes;    /* device bytes in use: per-MSA arrays + peak of the scratch arena */
	double kernel_ms[FBG_N_KERNELS];
	double kernel_bytes[FBG_N_KERNELS];
	uint64_t kernel_launches[FBG_N_KERNELS];
	int64_t dp_passes;            /* K3: verify passes of the chunked DP (0 = one chunk, < 0 = that many passes were spent and the whole-range wavefront was needed) */
	int64_t kmer_dict;            /* K1 round 0: 1 = through the k-mer dictionary, 0 = by sorting all suffixes (multi-GPU build, too many distinct k-mers, 64-bit k-mers) */
	int64_t index_depth;          /* 0 = the full index; D > 0 = K1 stopped at depth D, where every bucket still tied lay in one column of a gapless MSA (the "sufficient index": same f, segmentation and graph) */
	int64_t index_virtual;        /* 1 = that stop came straight after the k-mer round with no ignore characters set: no SA / ISA / LCP array was written at all, f(x) was computed from the sorted list of distinct k-mers (fcol.cu k2_from_dict) */
} fbg_stats;

/* index -> name for the arrays above ("pack", "kmer", "radix_hist", ...) */
const char *fbg_kernel_name(int idx);

/* device < 0: current device. */
int fbg_ctx_create(fbg_ctx **out, int device);
void fbg_ctx_destroy(fbg_ctx *ctx);
const char *fbg_last_error(const fbg_ctx *ctx);  /* ctx may be NULL: error of the last failed fbg_ctx_create */

/* Mode flags exactly as main() reads them (fbg.cpp:3308-3340): --elastic,
 * --gap-limit, --ignore-chars, --disable-elastic-tricks. */
int fbg_set_options(fbg_ctx *ctx, int elastic, int64_t gap_limit, const uint8_t *ignore_chars, size_t n_ignore, int disable_tricks);

/* Rows already filtered by the host (read_input, fbg.cpp:151-201), row-major m x n
 * ASCII bytes, '-' = gap.  Host buffer (pinned recommended); copied before return
 * of the next synchronising call.  Replaces what load_cst consumes (fbg.cpp:361). */
int fbg_upload_msa(fbg_ctx *ctx, const uint8_t *rows_rowmajor, uint64_t m, uint64_t n);
/* Same, from a device buffer on the context's device (inputs resident in HBM). */
int fbg_upload_msa_device(fbg_ctx *ctx, const void *device_rows_rowmajor, uint64_t m, uint64_t n);

/* K0+K1: text, suffix array, inverse, LCP. Replaces load_cst -> sdsl::construct (fbg.cpp:3378, 428).
 * Called explicitly it builds the FULL index (true SA / ISA / LCP, what the read-backs below return).
 * The later stages need less and build it themselves when this call is left out: on a gapless MSA
 * (all of non-elastic --gap-limit=1) prefix doubling stops at the first depth at which every bucket
 * of tied suffixes lies in one column -- order and LCP inside such a bucket cannot change f(x)
 * (suffix.cu) -- see fbg_stats.index_depth.  FBG_FULL_INDEX=1 in the environment switches that off. */
int fbg_build_index(fbg_ctx *ctx);
/* K2: f(x). Replaces compute_f / compute_f_range (fbg.cpp:1923, 2278-2289); in non-elastic
 * mode computes f with tricks disabled, the quantity v[j] is derived from (fbg.cpp:552-611). */
int fbg_compute_f(fbg_ctx *ctx);
/* K3: DP + backtrack. Replaces fbg.cpp:1940-2039 (elastic), 622-663 (gap_limit == 1),
 * 827-865 (otherwise).  Runs the earlier stages if needed.  *score = mml[n] ("optimal
 * segment length") or s[n-1] ("Optimal score").  FBG_ERR_NOSEG when the reference would
 * print its no-segmentation message and exit 1. */
int fbg_segment(fbg_ctx *ctx, uint64_t *n_blocks, uint64_t *score);
/* out[n_blocks]: inclusive 0-based end column of every block; elastic mode keeps the
 * reference's quirk that the last element is n (fbg.cpp:2026-2039). */
int fbg_get_boundaries(fbg_ctx *ctx, uint64_t *out);
/* K4: nodes + edges (fbg.cpp:1222-1260 / 666-728). want_paths != 0 also fills node_of. */
int fbg_build_graph(fbg_ctx *ctx, int want_paths, fbg_graph *out);
void fbg_graph_free(fbg_graph *g);

/* K4 + K5: nodes and edges as above, and the xGFA bytes of output_efg
 * (fbg.cpp:1185-1301: M, X, B lines, per block S then L lines, P lines with output_paths)
 * formatted ON THE DEVICE; the graph arrays are not copied to the host.  ids = m header strings
 * (needed with output_paths).  counts (may be NULL) receives the scalar fields of fbg_graph
 * (n_blocks, n_nodes, n_edges, n_rows, label_total, max_height); its array pointers stay NULL
 * and it needs no fbg_graph_free.  The bytes stay in device memory until the next
 * fbg_format_xgfa / upload on this context; read or save them with the two calls below.
 * Byte-identical to fbg_write_xgfa. */
int fbg_format_xgfa(fbg_ctx *ctx, int output_paths, const char *const *ids, fbg_graph *counts, uint64_t *n_bytes);
/* dst[len] <- bytes [offset, offset + len) of the formatted xGFA. */
int fbg_read_xgfa(fbg_ctx *ctx, uint64_t offset, uint64_t len, uint8_t *dst);
/* Streams the formatted xGFA into a file (device -> two page-locked staging buffers -> write(),
 * copy and write overlapped).  Replaces the std::fstream of output_efg (fbg.cpp:1193). */
int fbg_save_xgfa(fbg_ctx *ctx, const char *path);

/* Whole path in one call: upload (host buffer) -> index -> f -> DP -> graph. */
int fbg_run(fbg_ctx *ctx, const uint8_t *rows_rowmajor, uint64_t m, uint64_t n, int want_paths,
            uint64_t *n_blocks, uint64_t *score, fbg_graph *out);
/* Same with the MSA already resident in device memory. */
int fbg_run_device(fbg_ctx *ctx, const void *device_rows, uint64_t m, uint64_t n, int want_paths,
                   uint64_t *n_blocks, uint64_t *score, fbg_graph *out);

/* ---- --heuristic-subset=ROWNUM (hidden option of the reference: memory relief at the price of optimality).
 * fbg_heuristic_f replaces the row-chunk loop of main() and its workers (fbg.cpp:3400-3433, 2158-2178): for every
 * chunk of ROWNUM consecutive rows the chunk's own index and f(x) are built (K0..K2, one chunk in device memory at a
 * time), f = the maximum over the chunks; then the whole MSA is uploaded and packed (no index of it is ever built)
 * and f installed, so that fbg_segment -> fbg_validate_blocks (until nothing is removed) -> fbg_build_graph /
 * fbg_format_xgfa follow.  Elastic options only.  chunk_gaps / chunk_ignore (may be NULL, else ceil(m / rownum)
 * entries) receive the two "MSA contains ..." counts the reference prints per chunk (fbg.cpp:1858-1859).
 * fbg_validate_blocks replaces ONE iteration of the repair loop (fbg.cpp:3467-3500: make_efg_external +
 * efg_validate 3199-3292): a node that is neither source nor sink, and holds no ignore character, is invalid when
 * its label occurs in an edge string label(u)label(v) anywhere but at the start of u or v with that node in its
 * own block; every block holding an invalid node is merged with its predecessor.  The edge strings are indexed by
 * the same suffix sorter as the MSA.  *n_removed = boundaries dropped in this pass, *n_blocks = blocks left; counts
 * (may be NULL) = scalar fields of the graph that was checked ("#nodes=", "total length of node labels=",
 * "#edges=", fbg.cpp:1119-1120, 1172).  Call until *n_removed == 0. */
int fbg_heuristic_f(fbg_ctx *ctx, const uint8_t *rows_rowmajor, uint64_t m, uint64_t n, uint64_t rownum, uint64_t *chunk_gaps, uint64_t *chunk_ignore);
int fbg_validate_blocks(fbg_ctx *ctx, uint64_t *n_removed, uint64_t *n_blocks, fbg_graph *counts);

/* ---- one MSA across the GPUs of a box (new: the reference is single-process; its only
 * parallelism is std::thread over column ranges, fbg.cpp:2278-2289).  One process per GPU;
 * every process uploads the same MSA and makes the same calls.  Suffix sorting is sharded by
 * k-mer key range, f(x) by contiguous column range; rank array and f are merged by NCCL
 * all-reduce(max) over NVLink.  Results (SA, ISA, LCP, f, segmentation, graph) are replicated
 * on every rank and bit-identical to the single-GPU build. */
int fbg_device_count(void);                 /* CUDA devices visible to this process (0 when there is no usable driver) */
int fbg_dist_unique_id(uint8_t *out128);   /* rank 0: ncclGetUniqueId; ship the 128 bytes to the other ranks */
int fbg_dist_init(fbg_ctx *ctx, int rank, int world, const uint8_t *id128);   /* collective */
int fbg_dist_shutdown(fbg_ctx *ctx);
/* The sharding plans as pure host functions (no device needed):
 * bounds[world+1]: rank r sorts the suffixes whose k-mer histogram bin is in [bounds[r], bounds[r+1]) */
int fbg_plan_key_ranges(const uint32_t *hist, uint32_t nbins, int world, uint32_t *bounds);
/* rank r computes f(x) for x in [*x0, *x1) */
int fbg_plan_columns(uint64_t n, int world, int rank, uint64_t *x0, uint64_t *x1);

/* K3 tuning, for the parity tests: columns per chunk and warm-up columns of the chunked
 * guess-and-verify DP (0, 0 = defaults 2048 / 1024).  Results never depend on it. */
int fbg_set_dp_chunking(fbg_ctx *ctx, uint64_t chunk, uint64_t warmup);

/* The CUDA stream (cudaStream_t) every kernel of this context is launched on, so a caller
 * can bracket the work with its own events. */
void *fbg_get_stream(fbg_ctx *ctx);

int fbg_get_stats(fbg_ctx *ctx, fbg_stats *out);
void fbg_reset_stats(fbg_ctx *ctx);

/* Parity / debug read-backs (host buffers sized by the caller). */
int fbg_get_f(fbg_ctx *ctx, uint64_t *out /* n */);
int fbg_get_text(fbg_ctx *ctx, uint8_t *out /* text_len */);
int fbg_get_sa(fbg_ctx *ctx, uint64_t *out /* text_len */);
int fbg_get_isa(fbg_ctx *ctx, uint64_t *out /* text_len */);
int fbg_get_lcp(fbg_ctx *ctx, uint64_t *out /* text_len, SA order: LCP[k] */);
int fbg_get_dp(fbg_ctx *ctx, uint64_t *a /* mml[n+1] or s[n] */, uint64_t *b /* backtrack[n+1] or prev[n] */, uint64_t *v /* n, non-elastic only, may be NULL */);

/* Size-independent property check of the index for inputs whose arrays are too large to read
 * back: nsamples pseudo-random ranks k; ISA[SA[k]] = k, the suffixes SA[k-1] and SA[k] share
 * exactly LCP[k] characters (direct text comparison on the device) and are in order.
 * *n_bad = samples that failed. */
int fbg_check_index(fbg_ctx *ctx, uint32_t nsamples, uint64_t seed, uint32_t *n_bad);

/* Standalone primitive, exported for the parity tests of the radix passes:
 * stable LSD sort of (key,value) pairs on bits [lo_bit,hi_bit). Host buffers. */
int fbg_radix_sort_pairs(fbg_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t count, int lo_bit, int hi_bit);
/* Device-resident variant used by bench.py to time one radix pass set in isolation:
 * sorts `count` pseudo-random 64-bit keys generated on the device; returns elapsed ms. */
int fbg_radix_sort_bench(fbg_ctx *ctx, uint64_t count, int lo_bit, int hi_bit, int reps, double *ms_per_sort);

/* The same radix passes for callers that cannot allocate: code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc.  This entry point has NO counterpart in the
 * reference.  It calls no cudaMalloc / cudaFree
 * and no allocating library routine: its scratch (one ping-pong buffer of pairs, the digit count
 * matrix, the digit totals) is a fixed __device__ array inside the library, FBG_NOALLOC_SCRATCH_BYTES
 * held by every GPU that loads the library until the process exits.
 *   - DEVICE buffers on the context's device, inputs only read, outputs distinct from the inputs;
 *     count <= FBG_NOALLOC_MAX_PAIRS, else FBG_ERR_UNSUPPORTED (fbg_last_error names the limit).
 *   - Stable LSD sort on bits [lo_bit, hi_bit), bit-identical to fbg_radix_sort_pairs.
 *   - Work is ENQUEUED on the context's stream (fbg_get_stream) and the call returns without
 *     synchronising: the outputs are ready when that stream has reached this point.
 *   - Every call on a GPU -- whatever the context or stream -- shares the one array, which keeps the
 *     last call's contents (a call clears what it needs): calls that overlap on the GPU race.
 *     Calls on one context are ordered by its stream; across contexts the caller orders them. */
#define FBG_NOALLOC_MAX_PAIRS (1ull << 20)
#define FBG_NOALLOC_SCRATCH_BYTES 12847104ull  /* 2^20 x (8 + 4) B pairs + 256 x 256 x 4 B counts + 256 x 8 B totals */
int fbg_radix_sort_pairs_noalloc(fbg_ctx *ctx, const uint64_t *d_keys_in, const uint32_t *d_vals_in, uint64_t *d_keys_out, uint32_t *d_vals_out,
                                 uint64_t count, int lo_bit, int hi_bit);

/* Host-side xGFA formatting, byte-identical to output_efg (fbg.cpp:1185-1301).
 * msa = the same filtered row-major rows; ids = m header strings (used when output_paths). */
int fbg_write_xgfa(const char *path, const uint8_t *rows_rowmajor, uint64_t m, uint64_t n,
                   const uint64_t *boundaries, const fbg_graph *g, int output_paths, const char *const *ids);

/* Host-side FASTA reader: read_input / check_sequence_length / check_gaps (fbg.cpp:103-201),
 * same acceptance rules; the file is mapped and the records are located, measured, filtered
 * and copied by several threads (threads <= 0: as many as the host offers, at most 16).
 * rows: m x n kept rows, row-major, what fbg_upload_msa takes.  ids: EVERY header without its
 * first byte when output_paths (fbg.cpp:161,169; may outnumber m), else NULL.
 * messages: the reference's WARNING / NOTICE stderr lines of the skipped records, in file
 * order, for the caller to print.  FBG_ERR_ARG when the file cannot be read.  Everything is
 * owned by the library until fbg_msa_free. */
typedef struct fbg_msa {
	uint8_t *rows;
	uint64_t m, n;
	const char *const *ids;
	uint64_t n_ids;
	const char *messages;
	void *opaque;
} fbg_msa;
int fbg_read_fasta(const char *path, uint64_t gap_limit, int elastic, int output_paths, int threads, fbg_msa *out);
void fbg_msa_free(fbg_msa *msa);

#ifdef __cplusplus
}
#endif
#endif /* FBG_CUDA_H */
