// capi.cu -- the extern "C" surface declared in include/fbg_cuda.h.
#include "../../include/fbg_cuda.h"
#include "pipeline.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace fbg;

struct fbg_ctx {
	Pipeline p;
	Pipeline *aux = nullptr; // --heuristic-subset: indexes the edge text of the repair loop (validate.cu); lives until the loop is done
	std::string err;
};

// everything a Pipeline owns on the device
static void release_pipeline(Pipeline &p) {
	if (p.dev.stream) cudaStreamSynchronize(p.dev.stream);
	p.free_all();
	if (p.scalars) { cudaFree(p.scalars); p.scalars = nullptr; }
	if (p.h_scalars) { cudaFreeHost(p.h_scalars); p.h_scalars = nullptr; }
	if (p.xgfa) { cudaFree(p.xgfa); p.xgfa = nullptr; }
	p.dev.destroy();
	if (p.dev.stream) { cudaStreamDestroy(p.dev.stream); p.dev.stream = nullptr; }
}
static void drop_aux(fbg_ctx *ctx) {
	if (!ctx->aux) return;
	release_pipeline(*ctx->aux);
	delete ctx->aux;
	ctx->aux = nullptr;
}

static std::string g_create_err;

namespace fbg {

// independent spot check of the index: random ranks k, direct text comparison
__global__ void check_index_kernel(const u8 *__restrict__ T, const u32 *__restrict__ SA, const u32 *__restrict__ ISA, const u32 *__restrict__ PLCP,
                                   u64 N, u64 seed, u32 nsamples, u32 *__restrict__ bad) {
	u32 i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nsamples || N < 2) return;
	u64 z = ((u64) i + seed) * 0x9E3779B97F4A7C15ull;
	z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
	u64 k = 1 + z % (N - 1);
	u64 a = SA[k - 1], b = SA[k];
	bool ok = a < N && b < N && ISA[a] == (u32) (k - 1) && ISA[b] == (u32) k;
	if (ok) {
		u64 l = PLCP[b];
		ok = a + l <= N && b + l <= N;
		for (u64 q = 0; ok && q < l; q++) ok = T[a + q] == T[b + q];
		// the two suffixes differ right after their common prefix, the smaller one first
		if (ok) ok = a + l < N && b + l < N ? T[a + l] < T[b + l] : a + l == N;
	}
	if (!ok) atomicAdd(bad, 1u);
}

// every per-MSA array lives in the persistent arena: dropping them is resetting the bump pointer
void Pipeline::free_all() {
	NG = IG = RK = NIA = start = rowlen = lastcol = nullptr;
	T = nullptr; msa = nullptr; BWT = nullptr; have_bwt = false;
	col = ISA = SA = PLCP = f = bnd = dp_a = dp_b = dp_v = nullptr;
	dev.pers_top = 0;
	have_msa = have_index = have_f = have_seg = have_xgfa = false;
	want_full_index = index_lazy = false; lazy_depth = 0;
	index_virtual = force_arrays = false; dict_col = dict_lcp = nullptr; dict_nd = 0;
	f_external = false;
}

} // namespace fbg

static const char *KNAMES[K_COUNT] = {"pack", "kmer", "radix_hist", "radix_scatter", "scan", "doubling_aux", "lcp", "fcol", "dp", "graph", "tile_sort", "comm", "emit", "edges"};

// A stage that throws leaves its scratch allocations on the bump arena and possibly an open timing record: put the
// context back into the state every entry point expects (scratch empty, nothing in flight), so that the caller can
// go on using it -- upload another MSA, change the options -- after an error.  The per-MSA arrays stay; whatever the
// failed stage was producing is simply not marked as present.
static void recover_after_error(fbg_ctx *ctx) {
	for (Pipeline *q : {&ctx->p, ctx->aux}) {
		if (!q) continue;
		if (q->dev.stream) cudaStreamSynchronize(q->dev.stream);
		q->dev.arena_top = 0;
		q->dev.open_id = -1;
		for (auto &r : q->dev.recs) { q->dev.pool.push_back(r.a); q->dev.pool.push_back(r.b); }
		q->dev.recs.clear();
	}
	cudaGetLastError();
}

#define FBG_TRY(ctx) try {
#define FBG_CATCH(ctx)                                                         \
	}                                                                          \
	catch (const fbg::unsupported_error &e) { (ctx)->err = e.what(); recover_after_error(ctx); return FBG_ERR_UNSUPPORTED; } \
	catch (const fbg::cuda_error &e) { (ctx)->err = e.what(); recover_after_error(ctx); return FBG_ERR_CUDA; } \
	catch (const std::exception &e) { (ctx)->err = e.what(); recover_after_error(ctx); return FBG_ERR_CUDA; }

extern "C" {

const char *fbg_kernel_name(int idx) { return (idx >= 0 && idx < K_COUNT) ? KNAMES[idx] : ""; }

int fbg_ctx_create(fbg_ctx **out, int device) {
	if (!out) return FBG_ERR_ARG;
	*out = nullptr;
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count == 0) {
		g_create_err = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU fallback)";
		cudaGetLastError();
		return FBG_ERR_CUDA;
	}
	if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) device = 0; }
	if (device >= count) { g_create_err = "device index out of range"; return FBG_ERR_ARG; }
	cudaDeviceProp prop;
	if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major < 10) {
		g_create_err = "device is not sm_100 class; this library is built for sm_100a only";
		cudaGetLastError();
		return FBG_ERR_CUDA;
	}
	fbg_ctx *c = new fbg_ctx();
	try {
		FBG_CUDA(cudaSetDevice(device));
		c->p.dev.device = device;
		FBG_CUDA(cudaStreamCreateWithFlags(&c->p.dev.stream, cudaStreamNonBlocking));
	} catch (const std::exception &ex) {
		g_create_err = ex.what();
		delete c;
		return FBG_ERR_CUDA;
	}
	*out = c;
	return FBG_OK;
}

void fbg_ctx_destroy(fbg_ctx *ctx) {
	if (!ctx) return;
	cudaSetDevice(ctx->p.dev.device);
	cudaStreamSynchronize(ctx->p.dev.stream);
	if (ctx->p.dist.comm) {
		ctx->p.dist.close_peers();
		if (ctx->p.dist.xchg) { cudaFree(ctx->p.dist.xchg); ctx->p.dist.xchg = nullptr; }
		ctx->p.dist.api.CommDestroy(ctx->p.dist.comm); ctx->p.dist.comm = nullptr;
	}
	drop_aux(ctx);
	ctx->p.free_all();
	if (ctx->p.scalars) cudaFree(ctx->p.scalars);
	if (ctx->p.h_scalars) cudaFreeHost(ctx->p.h_scalars);
	if (ctx->p.xgfa) cudaFree(ctx->p.xgfa);
	ctx->p.dev.destroy();
	if (ctx->p.dev.stream) cudaStreamDestroy(ctx->p.dev.stream);
	delete ctx;
}

const char *fbg_last_error(const fbg_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int fbg_set_options(fbg_ctx *ctx, int elastic, int64_t gap_limit, const uint8_t *ignore_chars, size_t n_ignore, int disable_tricks) {
	if (!ctx) return FBG_ERR_ARG;
	if (gap_limit < 0) { ctx->err = "Gap limit needs to be non-negative."; return FBG_ERR_ARG; }
	Options &o = ctx->p.opt;
	u32 old_bits[8];
	for (int q = 0; q < 8; q++) old_bits[q] = o.ignore_bits[q];
	o.elastic = elastic != 0; o.gap_limit = gap_limit; o.disable_tricks = disable_tricks != 0;
	for (int q = 0; q < 8; q++) o.ignore_bits[q] = 0;
	o.n_ignore = 0;
	// ignore characters only take part in the elastic path (fbg.cpp:3393)
	if (elastic && ignore_chars)
		for (size_t i = 0; i < n_ignore; i++) { u8 c = ignore_chars[i]; o.ignore_bits[c >> 5] |= 1u << (c & 31); o.n_ignore++; }
	ctx->p.have_f = ctx->p.have_seg = ctx->p.have_xgfa = false;
	// the ignore bitmaps and the "characters to ignore" count are made by K0: pack again when the set changed
	for (int q = 0; q < 8; q++) if (old_bits[q] != o.ignore_bits[q]) ctx->p.have_index = false;
	return FBG_OK;
}

static int upload_common(fbg_ctx *ctx, const void *rows, uint64_t m, uint64_t n, bool device) {
	if (!ctx || !rows || m == 0 || n == 0) { if (ctx) ctx->err = "empty MSA"; return FBG_ERR_ARG; }
	// limits of this build, reported before any work is done (DESIGN.md "Limits")
	if (m > FBG_MAX_ROWS) { ctx->err = "more than " + std::to_string(FBG_MAX_ROWS) + " rows: the per-column sorts of K2 / K4 run in shared memory"; return FBG_ERR_UNSUPPORTED; }
	if (n >= 0xFFFFFFF0ull) { ctx->err = "more than 2^32 - 16 columns"; return FBG_ERR_UNSUPPORTED; }
	FBG_TRY(ctx)
	FBG_CUDA(cudaSetDevice(ctx->p.dev.device));
	ctx->p.have_xgfa = false;
	if (device) k0_upload_device(ctx->p, (const u8 *) rows, m, n);
	else k0_upload_host(ctx->p, (const u8 *) rows, m, n);
	FBG_CATCH(ctx)
	return FBG_OK;
}
int fbg_upload_msa(fbg_ctx *ctx, const uint8_t *rows, uint64_t m, uint64_t n) { return upload_common(ctx, rows, m, n, false); }
int fbg_upload_msa_device(fbg_ctx *ctx, const void *rows, uint64_t m, uint64_t n) { return upload_common(ctx, rows, m, n, true); }

// full = the true SA / ISA / LCP are wanted (explicit fbg_build_index, read-backs); otherwise K1 may
// stop at the sufficient index (suffix.cu).  A sufficient index is replaced when the full one is asked for.
static int build_index_impl(fbg_ctx *ctx, bool full, bool need_arrays = true) {
	if (!ctx || !ctx->p.have_msa) { if (ctx) ctx->err = "no MSA uploaded"; return FBG_ERR_ARG; }
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	FBG_CUDA(cudaSetDevice(p.dev.device));
	if (p.have_index && !(full && p.index_lazy) && !(need_arrays && p.index_virtual)) return FBG_OK;
	p.want_full_index = full; // per call: a later rebuild on this MSA (new ignore set) may take the sufficient index again
	p.force_arrays = need_arrays; // need_arrays: a consumer that reads SA / ISA / PLCP follows
	if (!p.have_index) k0_pack(p);
	p.have_index = false;
	k1_suffix_array(p);
	k1_lcp(p);
	p.have_index = true;
	FBG_CATCH(ctx)
	return FBG_OK;
}

int fbg_build_index(fbg_ctx *ctx) {
	if (ctx && ctx->p.f_external) { ctx->err = "fbg_heuristic_f holds no index of the whole MSA: upload the MSA again for a full build"; return FBG_ERR_ARG; }
	return build_index_impl(ctx, true);
}

int fbg_compute_f(fbg_ctx *ctx) {
	if (!ctx) return FBG_ERR_ARG;
	// K2 reads the per-position arrays unless the k-mer table alone will do (fcol.cu k2_from_dict)
	int rc = build_index_impl(ctx, false, ctx && ctx->p.opt.n_ignore > 0);
	if (rc) return rc;
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	if (p.have_f) return FBG_OK;
	if (p.f_external) { ctx->err = "options changed after fbg_heuristic_f: call it again"; return FBG_ERR_ARG; }
	// non-elastic: v[j] is defined through f without the row-end tricks and without
	// ignore characters (fbg.cpp:552-611 has neither)
	struct Restore { bool &ref, val; ~Restore() { ref = val; } } restore{p.opt.disable_tricks, p.opt.disable_tricks}; // also when K2 throws
	if (!p.opt.elastic) p.opt.disable_tricks = true;
	k2_compute_f(p);
	FBG_CATCH(ctx)
	return FBG_OK;
}

int fbg_segment(fbg_ctx *ctx, uint64_t *n_blocks, uint64_t *score) {
	if (!ctx) return FBG_ERR_ARG;
	int rc = fbg_compute_f(ctx);
	if (rc) return rc;
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	if (!p.have_seg) {
		if (p.opt.elastic && p.opt.disable_tricks) {
			// fbg.cpp:1932-1937: f[0] == n -> "No valid segmentation found!"
			u32 f0 = 0;
			FBG_CUDA(cudaMemcpyAsync(&f0, p.f, sizeof(u32), cudaMemcpyDeviceToHost, p.dev.stream));
			FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
			if (f0 == p.n) { p.seg_status = 1; p.nb = 0; p.score = 0; p.have_seg = true; }
		}
		if (!p.have_seg) k3_segment(p);
	}
	if (n_blocks) *n_blocks = p.nb;
	if (score) *score = p.score;
	if (p.seg_status) { ctx->err = p.opt.elastic || p.opt.gap_limit != 1 ? "No valid segmentation found!" : "No proper segmentation exists."; return FBG_ERR_NOSEG; }
	FBG_CATCH(ctx)
	return FBG_OK;
}

static int copy_u32_as_u64(fbg_ctx *ctx, const u32 *src, u64 count, uint64_t *out) {
	FBG_TRY(ctx)
	std::vector<u32> tmp(count);
	if (count) FBG_CUDA(cudaMemcpyAsync(tmp.data(), src, count * sizeof(u32), cudaMemcpyDeviceToHost, ctx->p.dev.stream));
	FBG_CUDA(cudaStreamSynchronize(ctx->p.dev.stream));
	for (u64 i = 0; i < count; i++) out[i] = tmp[i];
	FBG_CATCH(ctx)
	return FBG_OK;
}

int fbg_get_boundaries(fbg_ctx *ctx, uint64_t *out) {
	if (!ctx || !out || !ctx->p.have_seg || ctx->p.seg_status) { if (ctx) ctx->err = "no segmentation available"; return FBG_ERR_ARG; }
	return copy_u32_as_u64(ctx, ctx->p.bnd, ctx->p.nb, out);
}

struct GraphHolder { Graph g; std::vector<u64> dummy; };

int fbg_build_graph(fbg_ctx *ctx, int want_paths, fbg_graph *out) {
	if (!ctx || !out || !ctx->p.have_seg || ctx->p.seg_status) { if (ctx) ctx->err = "no segmentation available"; return FBG_ERR_ARG; }
	std::memset(out, 0, sizeof *out);
	GraphHolder *h = new GraphHolder();
	try {
		k4_graph(ctx->p, &h->g, want_paths != 0);
	} catch (const fbg::unsupported_error &e) {
		ctx->err = e.what(); delete h; return FBG_ERR_UNSUPPORTED;
	} catch (const std::exception &e) {
		ctx->err = e.what(); cudaGetLastError(); delete h; return FBG_ERR_CUDA;
	}
	Graph &g = h->g;
	out->n_blocks = g.n_heights; out->n_nodes = g.n_nodes; out->n_edges = g.n_edges;
	out->n_rows = ctx->p.m; out->label_total = g.label_total;
	u64 mh = 0;
	for (u64 i = 0; i < g.n_heights; i++) mh = g.heights[i] > mh ? g.heights[i] : mh;
	out->max_height = mh;
	out->block_height = g.heights; out->node_block = g.node_block; out->node_row = g.node_row;
	out->edge_src = g.edge_src; out->edge_dst = g.edge_dst;
	out->node_of = g.n_node_of ? g.node_of : nullptr;
	out->opaque = h;
	return FBG_OK;
}

// K4 + K5 without a host copy of the graph arrays
int fbg_format_xgfa(fbg_ctx *ctx, int output_paths, const char *const *ids, fbg_graph *counts, uint64_t *n_bytes) {
	if (!ctx || !ctx->p.have_seg || ctx->p.seg_status) { if (ctx) ctx->err = "no segmentation available"; return FBG_ERR_ARG; }
	if (output_paths && !ids) { ctx->err = "output_paths needs the sequence headers"; return FBG_ERR_ARG; }
	if (counts) std::memset(counts, 0, sizeof *counts);
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	FBG_CUDA(cudaSetDevice(p.dev.device));
	p.have_xgfa = false;
	XgfaJob x;
	x.paths = output_paths != 0; x.ids = ids;
	k4_graph(p, nullptr, false, &x);
	if (counts) {
		counts->n_blocks = p.nb; counts->n_nodes = x.n_nodes; counts->n_edges = x.n_edges; counts->n_rows = p.m;
		counts->label_total = x.label_total; counts->max_height = x.max_height;
	}
	if (n_bytes) *n_bytes = x.n_bytes;
	p.dev.resolve();
	FBG_CATCH(ctx)
	return FBG_OK;
}

int fbg_read_xgfa(fbg_ctx *ctx, uint64_t offset, uint64_t len, uint8_t *dst) {
	if (!ctx || !ctx->p.have_xgfa || (len && !dst) || offset > ctx->p.xgfa_len || len > ctx->p.xgfa_len - offset) {
		if (ctx) ctx->err = ctx->p.have_xgfa ? "range outside the formatted xGFA" : "no formatted xGFA (call fbg_format_xgfa)";
		return FBG_ERR_ARG;
	}
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	FBG_CUDA(cudaSetDevice(p.dev.device));
	if (len) FBG_CUDA(cudaMemcpyAsync(dst, p.xgfa + offset, len, cudaMemcpyDeviceToHost, p.dev.stream));
	FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
	FBG_CATCH(ctx)
	return FBG_OK;
}

// device bytes -> file through two page-locked staging buffers: the copy of chunk k+1 runs
// while chunk k is handed to write()
int fbg_save_xgfa(fbg_ctx *ctx, const char *path) {
	if (!ctx || !path || !ctx->p.have_xgfa) { if (ctx) ctx->err = "no formatted xGFA (call fbg_format_xgfa)"; return FBG_ERR_ARG; }
	FILE *fp = fopen(path, "wb");
	if (!fp) { ctx->err = std::string("cannot open ") + path; return FBG_ERR_ARG; }
	setvbuf(fp, nullptr, _IONBF, 0);
	u64 chunk = 4u << 20;
	if (const char *e = getenv("FBG_SAVE_CHUNK")) { long long v = atoll(e); if (v >= 1) chunk = (u64) v; } // test hook
	PinnedPool::Block blk;
	cudaEvent_t ev[2] = {nullptr, nullptr};
	int rc = FBG_OK;
	try {
		Pipeline &p = ctx->p;
		FBG_CUDA(cudaSetDevice(p.dev.device));
		const u64 total = p.xgfa_len;
		if (total <= 2 * chunk) {
			// small file: page-locking the staging buffers would cost more than the copy
			std::vector<u8> tmp(total);
			if (total) FBG_CUDA(cudaMemcpyAsync(tmp.data(), p.xgfa, total, cudaMemcpyDeviceToHost, p.dev.stream));
			FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
			if (total && fwrite(tmp.data(), 1, total, fp) != total) { ctx->err = std::string("cannot write ") + path; rc = FBG_ERR_ARG; }
		} else {
			blk = PinnedPool::get().take(2 * chunk);
			u8 *buf[2] = {static_cast<u8 *>(blk.ptr), static_cast<u8 *>(blk.ptr) + chunk};
			FBG_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
			FBG_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
			const u64 nchunks = (total + chunk - 1) / chunk;
			auto issue = [&](u64 k) {
				const u64 off = k * chunk, len = total - off < chunk ? total - off : chunk;
				FBG_CUDA(cudaMemcpyAsync(buf[k & 1], p.xgfa + off, len, cudaMemcpyDeviceToHost, p.dev.stream));
				FBG_CUDA(cudaEventRecord(ev[k & 1], p.dev.stream));
			};
			issue(0);
			for (u64 k = 0; k < nchunks; k++) {
				if (k + 1 < nchunks) issue(k + 1); // into the buffer written out in the iteration before
				FBG_CUDA(cudaEventSynchronize(ev[k & 1]));
				const u64 off = k * chunk, len = total - off < chunk ? total - off : chunk;
				if (fwrite(buf[k & 1], 1, len, fp) != len) { ctx->err = std::string("cannot write ") + path; rc = FBG_ERR_ARG; break; }
			}
			FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
		}
	} catch (const std::exception &e) {
		ctx->err = e.what(); cudaGetLastError(); rc = FBG_ERR_CUDA;
	}
	if (ev[0]) cudaEventDestroy(ev[0]);
	if (ev[1]) cudaEventDestroy(ev[1]);
	PinnedPool::get().give(blk);
	if (fclose(fp) != 0 && rc == FBG_OK) { ctx->err = std::string("cannot write ") + path; rc = FBG_ERR_ARG; }
	return rc;
}

void fbg_graph_free(fbg_graph *g) {
	if (!g || !g->opaque) return;
	delete static_cast<GraphHolder *>(g->opaque);
	std::memset(g, 0, sizeof *g);
}

static int run_common(fbg_ctx *ctx, const void *rows, uint64_t m, uint64_t n, bool device, int want_paths, uint64_t *n_blocks, uint64_t *score, fbg_graph *out) {
	int rc = upload_common(ctx, rows, m, n, device);
	if (rc) return rc;
	rc = fbg_segment(ctx, n_blocks, score);
	if (rc) return rc;
	if (out) rc = fbg_build_graph(ctx, want_paths, out);
	if (rc) return rc;
	FBG_TRY(ctx)
	ctx->p.dev.resolve();
	FBG_CATCH(ctx)
	return FBG_OK;
}
int fbg_run(fbg_ctx *ctx, const uint8_t *rows, uint64_t m, uint64_t n, int want_paths, uint64_t *n_blocks, uint64_t *score, fbg_graph *out) {
	return run_common(ctx, rows, m, n, false, want_paths, n_blocks, score, out);
}
int fbg_run_device(fbg_ctx *ctx, const void *rows, uint64_t m, uint64_t n, int want_paths, uint64_t *n_blocks, uint64_t *score, fbg_graph *out) {
	return run_common(ctx, rows, m, n, true, want_paths, n_blocks, score, out);
}

namespace fbg {
__global__ void hs_max_kernel(u32 *__restrict__ acc, const u32 *__restrict__ f, u64 n, int first) {
	u64 x = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (x < n) { const u32 v = f[x]; acc[x] = first ? v : (v > acc[x] ? v : acc[x]); }
}
} // namespace fbg

// --heuristic-subset=ROWNUM, first half (fbg.cpp:3400-3433, worker 2158-2178): f(x) of every chunk of ROWNUM
// consecutive rows from the chunk's own index, f = their maximum.  Only one chunk's index is ever in device
// memory.  Then the WHOLE MSA is uploaded and packed (K0: text, bitmaps, rank tables -- what K4 / K5 read; no
// index) and f installed, so that fbg_segment, fbg_validate_blocks, fbg_build_graph and fbg_format_xgfa follow.
int fbg_heuristic_f(fbg_ctx *ctx, const uint8_t *rows, uint64_t m, uint64_t n, uint64_t rownum, uint64_t *chunk_gaps, uint64_t *chunk_ignore) {
	if (!ctx || !rows || m == 0 || n == 0 || rownum == 0) { if (ctx) ctx->err = "empty MSA or chunk size 0"; return FBG_ERR_ARG; }
	if (!ctx->p.opt.elastic) { ctx->err = "--heuristic-subset belongs to the elastic path (fbg.cpp:3390-3433)"; return FBG_ERR_ARG; }
	u32 *acc = nullptr;
	int rc = FBG_OK;
	drop_aux(ctx);
	{
		cudaSetDevice(ctx->p.dev.device);
		if (cudaMalloc(&acc, (n + 1) * sizeof(u32)) != cudaSuccess) { cudaGetLastError(); ctx->err = "cannot allocate the f accumulator"; return FBG_ERR_CUDA; }
	}
	u64 k = 0;
	for (u64 r0 = 0; r0 < m && rc == FBG_OK; r0 += rownum, k++) {
		const u64 mc = m - r0 < rownum ? m - r0 : rownum;
		rc = upload_common(ctx, rows + r0 * n, mc, n, false);
		if (rc == FBG_OK) rc = fbg_compute_f(ctx);
		if (rc != FBG_OK) break;
		Pipeline &p = ctx->p;
		if (chunk_gaps) chunk_gaps[k] = p.gaps;
		if (chunk_ignore) chunk_ignore[k] = p.toignore;
		hs_max_kernel<<<div_up(n, 256), 256, 0, p.dev.stream>>>(acc, p.f, n, r0 == 0 ? 1 : 0);
		if (cudaStreamSynchronize(p.dev.stream) != cudaSuccess) { ctx->err = "accumulating f failed"; cudaGetLastError(); rc = FBG_ERR_CUDA; }
	}
	if (rc == FBG_OK) rc = upload_common(ctx, rows, m, n, false);
	if (rc == FBG_OK) {
		try {
			Pipeline &p = ctx->p;
			k0_pack(p);
			p.f = p.dev.palloc<u32>(n + 1);
			FBG_CUDA(cudaMemcpyAsync(p.f, acc, n * sizeof(u32), cudaMemcpyDeviceToDevice, p.dev.stream));
			FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
			p.have_index = true; p.have_f = true; p.f_external = true; p.have_seg = false;
			p.sa_rounds = 0; p.kmer_dict = 0;
		} catch (const fbg::unsupported_error &e) { ctx->err = e.what(); rc = FBG_ERR_UNSUPPORTED; }
		catch (const std::exception &e) { ctx->err = e.what(); cudaGetLastError(); rc = FBG_ERR_CUDA; }
	}
	cudaFree(acc);
	return rc;
}

// --heuristic-subset, second half: ONE pass of the repair loop (fbg.cpp:3467-3500, efg_validate 3199-3292) over
// the graph of the current segmentation; blocks holding an invalid node are merged with their predecessor.
// *n_removed = boundaries dropped ("There are K blocks to remove"); call again until it is 0.
int fbg_validate_blocks(fbg_ctx *ctx, uint64_t *n_removed, uint64_t *n_blocks, fbg_graph *counts) {
	if (!ctx || !n_removed || !ctx->p.have_seg || ctx->p.seg_status) { if (ctx) ctx->err = "no segmentation available"; return FBG_ERR_ARG; }
	if (!ctx->p.opt.elastic) { ctx->err = "the repair loop belongs to the elastic path"; return FBG_ERR_ARG; }
	if (counts) std::memset(counts, 0, sizeof *counts);
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	FBG_CUDA(cudaSetDevice(p.dev.device));
	if (!ctx->aux) ctx->aux = new Pipeline();
	ValidateJob v;
	v.aux = ctx->aux;
	p.have_xgfa = false;
	const u64 nb_before = p.nb;
	k4_graph(p, nullptr, false, nullptr, &v);
	*n_removed = v.n_removed;
	if (n_blocks) *n_blocks = p.nb;
	if (counts) { counts->n_blocks = nb_before; counts->n_nodes = v.n_nodes; counts->n_edges = v.n_edges; counts->n_rows = p.m; counts->label_total = v.label_total; }
	p.dev.resolve();
	if (v.n_removed == 0) drop_aux(ctx); // the loop is over: give the edge index's memory back
	FBG_CATCH(ctx)
	return FBG_OK;
}

int fbg_device_count(void) {
	int count = 0;
	if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
	return count;
}

int fbg_dist_unique_id(uint8_t *out128) {
	if (!out128) return FBG_ERR_ARG;
	static NcclApi api;
	if (!api.load(g_create_err)) return FBG_ERR_UNSUPPORTED;
	NcclUniqueId id;
	int rc = api.GetUniqueId(&id);
	if (rc != 0) { g_create_err = std::string("ncclGetUniqueId: ") + api.GetErrorString(rc); return FBG_ERR_CUDA; }
	std::memcpy(out128, id.internal, 128);
	return FBG_OK;
}

int fbg_dist_init(fbg_ctx *ctx, int rank, int world, const uint8_t *id128) {
	if (!ctx || !id128 || world < 1 || rank < 0 || rank >= world) { if (ctx) ctx->err = "bad rank / world size"; return FBG_ERR_ARG; }
	FBG_TRY(ctx)
	Dist &ds = ctx->p.dist;
	if (ds.comm) { ctx->err = "context already joined a communicator"; return FBG_ERR_ARG; }
	FBG_CUDA(cudaSetDevice(ctx->p.dev.device));
	if (world == 1) { ds.rank = 0; ds.world = 1; return FBG_OK; }
	if (!ds.api.load(ctx->err)) return FBG_ERR_UNSUPPORTED;
	NcclUniqueId id;
	std::memcpy(id.internal, id128, 128);
	ds.check(ds.api.CommInitRank(&ds.comm, world, id, rank), "ncclCommInitRank");
	ds.rank = rank; ds.world = world;
	ctx->p.have_index = ctx->p.have_f = ctx->p.have_seg = false;
	FBG_CATCH(ctx)
	return FBG_OK;
}

int fbg_dist_shutdown(fbg_ctx *ctx) {
	if (!ctx) return FBG_ERR_ARG;
	Dist &ds = ctx->p.dist;
	if (ds.comm) {
		cudaSetDevice(ctx->p.dev.device);
		cudaStreamSynchronize(ctx->p.dev.stream);
		ds.close_peers();
		if (ds.xchg) { cudaFree(ds.xchg); ds.xchg = nullptr; }
		ds.api.CommDestroy(ds.comm);
		ds.comm = nullptr;
	}
	ds.rank = 0; ds.world = 1;
	ctx->p.have_index = ctx->p.have_f = ctx->p.have_seg = false;
	return FBG_OK;
}

int fbg_plan_key_ranges(const uint32_t *hist, uint32_t nbins, int world, uint32_t *bounds) {
	if (!hist || !bounds || world < 1 || nbins == 0) return FBG_ERR_ARG;
	plan_key_ranges(hist, nbins, world, bounds);
	return FBG_OK;
}

int fbg_plan_columns(uint64_t n, int world, int rank, uint64_t *x0, uint64_t *x1) {
	if (!x0 || !x1 || world < 1 || rank < 0 || rank >= world) return FBG_ERR_ARG;
	plan_columns(n, world, rank, x0, x1);
	return FBG_OK;
}

int fbg_set_dp_chunking(fbg_ctx *ctx, uint64_t chunk, uint64_t warmup) {
	if (!ctx || (chunk && chunk < 16) || warmup > chunk) { if (ctx) ctx->err = "chunk must be 0 or >= 16, warm-up <= chunk"; return FBG_ERR_ARG; }
	ctx->p.dp_chunk = chunk; ctx->p.dp_warm = warmup;
	ctx->p.have_seg = false;
	return FBG_OK;
}

void *fbg_get_stream(fbg_ctx *ctx) { return ctx ? (void *) ctx->p.dev.stream : nullptr; }

int fbg_get_stats(fbg_ctx *ctx, fbg_stats *out) {
	if (!ctx || !out) return FBG_ERR_ARG;
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
	p.dev.resolve();
	std::memset(out, 0, sizeof *out);
	out->m = p.m; out->n = p.n; out->text_len = p.N; out->gaps = p.gaps; out->to_ignore = p.toignore;
	out->sa_rounds = p.sa_rounds; out->arena_peak_bytes = p.dev.arena_peak + p.dev.pers_top; out->dp_passes = p.dp_passes; out->kmer_dict = p.kmer_dict;
	out->index_depth = p.index_lazy ? (int64_t) p.lazy_depth : 0;
	out->index_virtual = p.index_virtual ? 1 : 0;
	for (int k = 0; k < K_COUNT; k++) {
		out->kernel_ms[k] = p.dev.stats[k].ms; out->kernel_bytes[k] = p.dev.stats[k].bytes; out->kernel_launches[k] = p.dev.stats[k].launches;
	}
	FBG_CATCH(ctx)
	return FBG_OK;
}

void fbg_reset_stats(fbg_ctx *ctx) {
	if (!ctx) return;
	try { cudaStreamSynchronize(ctx->p.dev.stream); ctx->p.dev.resolve(); } catch (...) {}
	ctx->p.dev.reset_stats();
}

int fbg_get_f(fbg_ctx *ctx, uint64_t *out) {
	if (!ctx || !out || !ctx->p.have_f) { if (ctx) ctx->err = "f not computed"; return FBG_ERR_ARG; }
	return copy_u32_as_u64(ctx, ctx->p.f, ctx->p.n, out);
}
int fbg_get_text(fbg_ctx *ctx, uint8_t *out) {
	if (!ctx || !out || !ctx->p.have_index) { if (ctx) ctx->err = "index not built"; return FBG_ERR_ARG; }
	FBG_TRY(ctx)
	FBG_CUDA(cudaMemcpyAsync(out, ctx->p.T, ctx->p.N, cudaMemcpyDeviceToHost, ctx->p.dev.stream));
	FBG_CUDA(cudaStreamSynchronize(ctx->p.dev.stream));
	FBG_CATCH(ctx)
	return FBG_OK;
}
int fbg_get_sa(fbg_ctx *ctx, uint64_t *out) {
	if (!ctx || !out || !ctx->p.have_index || ctx->p.f_external) { if (ctx) ctx->err = ctx->p.f_external ? "no index of the whole MSA in --heuristic-subset mode" : "index not built"; return FBG_ERR_ARG; }
	if (int rc = build_index_impl(ctx, true)) return rc;
	return copy_u32_as_u64(ctx, ctx->p.SA, ctx->p.N, out);
}
int fbg_get_isa(fbg_ctx *ctx, uint64_t *out) {
	if (!ctx || !out || !ctx->p.have_index || ctx->p.f_external) { if (ctx) ctx->err = ctx->p.f_external ? "no index of the whole MSA in --heuristic-subset mode" : "index not built"; return FBG_ERR_ARG; }
	if (int rc = build_index_impl(ctx, true)) return rc;
	return copy_u32_as_u64(ctx, ctx->p.ISA, ctx->p.N, out);
}
int fbg_get_lcp(fbg_ctx *ctx, uint64_t *out) {
	if (!ctx || !out || !ctx->p.have_index || ctx->p.f_external) { if (ctx) ctx->err = ctx->p.f_external ? "no index of the whole MSA in --heuristic-subset mode" : "index not built"; return FBG_ERR_ARG; }
	if (int rc = build_index_impl(ctx, true)) return rc;
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	std::vector<u32> plcp(p.N), sa(p.N);
	FBG_CUDA(cudaMemcpyAsync(plcp.data(), p.PLCP, p.N * sizeof(u32), cudaMemcpyDeviceToHost, p.dev.stream));
	FBG_CUDA(cudaMemcpyAsync(sa.data(), p.SA, p.N * sizeof(u32), cudaMemcpyDeviceToHost, p.dev.stream));
	FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
	for (u64 k = 0; k < p.N; k++) out[k] = plcp[sa[k]];
	FBG_CATCH(ctx)
	return FBG_OK;
}
int fbg_get_dp(fbg_ctx *ctx, uint64_t *a, uint64_t *b, uint64_t *v) {
	if (!ctx || !ctx->p.have_seg || !ctx->p.dp_a) { if (ctx) ctx->err = "DP not computed"; return FBG_ERR_ARG; }
	Pipeline &p = ctx->p;
	u64 cnt = p.opt.elastic ? p.n + 1 : p.n;
	int rc = FBG_OK;
	if (a) rc = copy_u32_as_u64(ctx, p.dp_a, cnt, a);
	if (!rc && b) rc = copy_u32_as_u64(ctx, p.dp_b, cnt, b);
	if (!rc && v && !p.opt.elastic) rc = copy_u32_as_u64(ctx, p.dp_v, p.n, v);
	return rc;
}

int fbg_check_index(fbg_ctx *ctx, uint32_t nsamples, uint64_t seed, uint32_t *n_bad) {
	if (!ctx || !n_bad || !ctx->p.have_index || ctx->p.f_external) { if (ctx) ctx->err = "index not built"; return FBG_ERR_ARG; }
	if (int rc = build_index_impl(ctx, true)) return rc;
	FBG_TRY(ctx)
	Pipeline &p = ctx->p;
	FBG_CUDA(cudaSetDevice(p.dev.device));
	u32 *bad = reinterpret_cast<u32 *>(p.scalars + 40);
	FBG_CUDA(cudaMemsetAsync(bad, 0, sizeof(u64), p.dev.stream));
	if (nsamples) check_index_kernel<<<div_up(nsamples, 128), 128, 0, p.dev.stream>>>(p.T, p.SA, p.ISA, p.PLCP, p.N, seed, nsamples, bad);
	FBG_LAUNCH_CHECK();
	FBG_CUDA(cudaMemcpyAsync(n_bad, bad, sizeof(u32), cudaMemcpyDeviceToHost, p.dev.stream));
	FBG_CUDA(cudaStreamSynchronize(p.dev.stream));
	FBG_CATCH(ctx)
	return FBG_OK;
}

int fbg_radix_sort_pairs(fbg_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t count, int lo_bit, int hi_bit) {
	if (!ctx || !keys || !vals || lo_bit < 0 || hi_bit > 64 || hi_bit < lo_bit) return FBG_ERR_ARG;
	FBG_TRY(ctx)
	Dev &d = ctx->p.dev;
	FBG_CUDA(cudaSetDevice(d.device));
	u64 ntiles = (count + RS_TILE - 1) / RS_TILE;
	d.reserve(count * 24 + ntiles * RS_BINS * 12 + (8u << 20));
	u64 mk = d.mark();
	RadixBuffers rb;
	rb.k[0] = d.alloc<u64>(count); rb.k[1] = d.alloc<u64>(count);
	rb.v[0] = d.alloc<u32>(count); rb.v[1] = d.alloc<u32>(count);
	rb.cur = 0;
	FBG_CUDA(cudaMemcpyAsync(rb.k[0], keys, count * 8, cudaMemcpyHostToDevice, d.stream));
	FBG_CUDA(cudaMemcpyAsync(rb.v[0], vals, count * 4, cudaMemcpyHostToDevice, d.stream));
	int cur = radix_sort<u64, true>(d, rb, count, lo_bit, hi_bit);
	FBG_CUDA(cudaMemcpyAsync(keys, rb.k[cur], count * 8, cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaMemcpyAsync(vals, rb.v[cur], count * 4, cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	d.release(mk);
	FBG_CATCH(ctx)
	return FBG_OK;
}

__global__ void bench_fill_kernel(u64 *k, u32 *v, u64 n, u64 seed) {
	u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	u64 z = (i + seed) * 0x9E3779B97F4A7C15ull;
	z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
	k[i] = z; v[i] = (u32) i;
}

int fbg_radix_sort_bench(fbg_ctx *ctx, uint64_t count, int lo_bit, int hi_bit, int reps, double *ms_per_sort) {
	if (!ctx || !ms_per_sort || reps < 1) return FBG_ERR_ARG;
	FBG_TRY(ctx)
	Dev &d = ctx->p.dev;
	FBG_CUDA(cudaSetDevice(d.device));
	u64 ntiles = (count + RS_TILE - 1) / RS_TILE;
	d.reserve(count * 24 + ntiles * RS_BINS * 12 + (8u << 20));
	u64 mk = d.mark();
	RadixBuffers rb;
	rb.k[0] = d.alloc<u64>(count); rb.k[1] = d.alloc<u64>(count);
	rb.v[0] = d.alloc<u32>(count); rb.v[1] = d.alloc<u32>(count);
	cudaEvent_t a, b;
	FBG_CUDA(cudaEventCreate(&a)); FBG_CUDA(cudaEventCreate(&b));
	double total = 0;
	for (int r = 0; r < reps + 1; r++) {
		rb.cur = 0;
		bench_fill_kernel<<<div_up(count, 256), 256, 0, d.stream>>>((u64 *) rb.k[0], rb.v[0], count, (u64) r * 7919);
		FBG_CUDA(cudaEventRecord(a, d.stream));
		radix_sort<u64, true>(d, rb, count, lo_bit, hi_bit);
		FBG_CUDA(cudaEventRecord(b, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		float ms;
		FBG_CUDA(cudaEventElapsedTime(&ms, a, b));
		if (r > 0) total += ms; // first repetition is the warm-up
	}
	cudaEventDestroy(a); cudaEventDestroy(b);
	*ms_per_sort = total / reps;
	d.release(mk);
	FBG_CATCH(ctx)
	return FBG_OK;
}


} // extern "C"

// ---------------------------------------------------------------------------------------------
// fbg_radix_sort_pairs_noalloc: the radix passes for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  The scratch the passes need -- one
// ping-pong (key, value) buffer, the per-tile digit count matrix, the digit totals -- lives in a
// file-scope __device__ array of fixed size, so every GPU that loads the library holds
// FBG_NOALLOC_SCRATCH_BYTES (about 12.3 MB) until the process exits.
//
// In HOST code the array's name is only a host-side stand-in for the symbol, not a device address:
// its address is taken once per device, with that device current, through cudaGetSymbolAddress
// (noalloc_scratch_for below) and that pointer is what the kernels get.
namespace {

constexpr u64 NA_MAX_PAIRS = FBG_NOALLOC_MAX_PAIRS;
constexpr u64 NA_MAX_TILES = (NA_MAX_PAIRS + RS_TILE - 1) / RS_TILE;
constexpr u64 NA_OFF_KEYS = 0;                                            // u64[NA_MAX_PAIRS]
constexpr u64 NA_OFF_VALS = NA_OFF_KEYS + NA_MAX_PAIRS * sizeof(u64);      // u32[NA_MAX_PAIRS]
constexpr u64 NA_OFF_COUNTS = NA_OFF_VALS + NA_MAX_PAIRS * sizeof(u32);    // u32[RS_BINS][NA_MAX_TILES]
constexpr u64 NA_OFF_TOTALS = NA_OFF_COUNTS + (u64) RS_BINS * NA_MAX_TILES * sizeof(u32); // u64[RS_BINS]
constexpr u64 NA_BYTES = NA_OFF_TOTALS + (u64) RS_BINS * sizeof(u64);
static_assert(NA_BYTES == FBG_NOALLOC_SCRATCH_BYTES, "include/fbg_cuda.h states the size of the scratch array");
static_assert(NA_OFF_VALS % 256 == 0 && NA_OFF_COUNTS % 256 == 0 && NA_OFF_TOTALS % 256 == 0, "scratch sections stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
unsigned char *g_noalloc_ptr[NA_MAX_DEVICES] = {};

// Device address of the scratch array on `device` (which must be current); set-up work of the
// first call on a device: the symbol's address and the shared-memory opt-in of the scatter kernel.
unsigned char *noalloc_scratch_for(int device) {
	if (device < 0 || device >= NA_MAX_DEVICES) throw cuda_error("device index beyond the no-allocation entry point's table");
	std::lock_guard<std::mutex> lk(g_noalloc_mu);
	if (!g_noalloc_ptr[device]) {
		void *ptr = nullptr;
		FBG_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
		if (!ptr) throw cuda_error("cudaGetSymbolAddress returned a null address for the scratch array");
		FBG_CUDA(cudaFuncSetAttribute(rs_scatter_kernel<u64, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sizeof(RsSmem<u64>)));
		g_noalloc_ptr[device] = static_cast<unsigned char *>(ptr);
	}
	return g_noalloc_ptr[device];
}

} // namespace

extern "C" int fbg_radix_sort_pairs_noalloc(fbg_ctx *ctx, const uint64_t *d_keys_in, const uint32_t *d_vals_in, uint64_t *d_keys_out, uint32_t *d_vals_out,
                                            uint64_t count, int lo_bit, int hi_bit) {
	if (!ctx) return FBG_ERR_ARG;
	if (lo_bit < 0 || hi_bit > 64 || hi_bit < lo_bit) { ctx->err = "bit range must satisfy 0 <= lo_bit <= hi_bit <= 64"; return FBG_ERR_ARG; }
	if (count > NA_MAX_PAIRS) {
		ctx->err = "fbg_radix_sort_pairs_noalloc sorts at most " + std::to_string(NA_MAX_PAIRS) + " pairs (FBG_NOALLOC_MAX_PAIRS; its scratch is a fixed device array), got " +
		           std::to_string(count) + ": use fbg_radix_sort_pairs";
		return FBG_ERR_UNSUPPORTED;
	}
	if (count == 0) return FBG_OK;
	if (!d_keys_in || !d_vals_in || !d_keys_out || !d_vals_out) { ctx->err = "null device buffer"; return FBG_ERR_ARG; }
	if ((const void *) d_keys_in == (const void *) d_keys_out || (const void *) d_vals_in == (const void *) d_vals_out) { ctx->err = "input and output buffers must be distinct"; return FBG_ERR_ARG; }
	FBG_TRY(ctx)
	Dev &d = ctx->p.dev;
	FBG_CUDA(cudaSetDevice(d.device));
	unsigned char *sc = noalloc_scratch_for(d.device);
	u64 *tkeys = reinterpret_cast<u64 *>(sc + NA_OFF_KEYS);
	u32 *tvals = reinterpret_cast<u32 *>(sc + NA_OFF_VALS);
	u32 *counts = reinterpret_cast<u32 *>(sc + NA_OFF_COUNTS);
	u64 *totals = reinterpret_cast<u64 *>(sc + NA_OFF_TOTALS);
	const int total = hi_bit - lo_bit;
	const int passes = (total + 7) / 8;
	if (passes == 0) { // nothing to sort by: the stable order is the input order
		FBG_CUDA(cudaMemcpyAsync(d_keys_out, d_keys_in, count * sizeof(u64), cudaMemcpyDeviceToDevice, d.stream));
		FBG_CUDA(cudaMemcpyAsync(d_vals_out, d_vals_in, count * sizeof(u32), cudaMemcpyDeviceToDevice, d.stream));
		return FBG_OK;
	}
	const u32 ntiles = div_up(count, RS_TILE);
	// The array keeps what the last call left.  Every count and total a pass reads is written by that
	// pass first (rs_hist_kernel stores all RS_BINS x ntiles entries, rs_binscan_kernel all totals);
	// they are cleared all the same, so that no call can ever depend on an earlier one.
	FBG_CUDA(cudaMemsetAsync(counts, 0, (u64) RS_BINS * ntiles * sizeof(u32), d.stream));
	FBG_CUDA(cudaMemsetAsync(totals, 0, (u64) RS_BINS * sizeof(u64), d.stream));
	// ping-pong between the scratch buffer and the caller's output so that the last pass lands in the output;
	// the input is only read
	const u64 *kin = d_keys_in;
	const u32 *vin = d_vals_in;
	int bit = lo_bit;
	for (int p = 0; p < passes; p++) {
		const int bits = (hi_bit - bit + (passes - p) - 1) / (passes - p); // even split, like radix_sort
		const u32 mask = (1u << bits) - 1;
		const bool to_out = ((passes - 1 - p) & 1) == 0;
		u64 *kout = to_out ? d_keys_out : tkeys;
		u32 *vout = to_out ? d_vals_out : tvals;
		rs_hist_kernel<u64><<<ntiles, RS_HTHREADS, 0, d.stream>>>(kin, count, bit, mask, ntiles, counts);
		rs_binscan_kernel<<<RS_BINS, 1024, 0, d.stream>>>(counts, ntiles, totals);
		rs_scatter_kernel<u64, true><<<ntiles, RS_THREADS, sizeof(RsSmem<u64>), d.stream>>>(kin, vin, kout, vout, count, bit, mask, ntiles, counts, totals);
		FBG_LAUNCH_CHECK();
		kin = kout; vin = vout;
		bit += bits;
	}
	FBG_CATCH(ctx)
	return FBG_OK;
}
