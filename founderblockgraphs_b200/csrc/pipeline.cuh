// pipeline.cuh -- device-resident state of one founder-block-graph build and the
// phase entry points (K0..K4).  Data layout in HBM (all indices u32, N < 2^32):
//
//   msa    m x pitch bytes, row-major, pitch = round_up(n,32), pad = '-'
//   NG,IG  [W][m] u32   non-gap / ignore-char bitmaps per 32-column tile (tile-major)
//   RK     [W][m] u32   rank_i(32 t): non-gap chars of row i before tile t
//   NIA    [W][m] u32   first ignore column of row i in tiles > t (U32_MAX if none)
//   start  [m+1], rowlen[m], lastcol[m]
//   T      [N]   u8     gap-stripped rows, '#' after each, final 0 sentinel
//   col    [N]   u32    text position -> MSA column ('#': n)
//   ISA    [N]   u32    suffix rank of every text position
//   SA     [N]   u32    suffix array (LCP construction; LCP[k+1] at run ends in K2)
//   PLCP   [N]   u32    LCP[ISA[p]]      (text order)
//   f      [n]   u32    minimal right extension
//   bnd    [nb]  u32    inclusive block end columns (elastic: last = n)
//   xgfa   bytes        the formatted graph (K5), only when asked for
#pragma once
#include <mutex>
#include <vector>
#include "prims.cuh"
#include "dist.cuh"

namespace fbg {

struct Options {
	bool elastic = true;
	i64 gap_limit = 1;
	bool disable_tricks = false;
	u32 ignore_bits[8] = {0, 0, 0, 0, 0, 0, 0, 0}; // 256-bit membership set
	int n_ignore = 0;
};

struct IgnoreSet { u32 bits[8]; };

// Page-locked host blocks for the graph arrays handed out through the C ABI: the device
// copies land in them at full PCIe speed, and a context that builds graph after graph gets the
// same blocks back (fbg_graph_free returns them), so there is no cudaHostAlloc and no
// first-touch page fault in steady state.  Process-wide because fbg_graph_free has no context.
struct PinnedPool {
	struct Block { void *ptr = nullptr; u64 bytes = 0; };
	static constexpr size_t KEEP = 4; // free blocks kept for reuse
	std::mutex mu;
	std::vector<Block> free_blocks;
	static PinnedPool &get() { static PinnedPool p; return p; }
	Block take(u64 bytes) {
		if (bytes == 0) bytes = 1;
		{
			std::lock_guard<std::mutex> lk(mu);
			size_t best = free_blocks.size();
			for (size_t i = 0; i < free_blocks.size(); i++)
				if (free_blocks[i].bytes >= bytes && (best == free_blocks.size() || free_blocks[i].bytes < free_blocks[best].bytes)) best = i;
			if (best != free_blocks.size()) { Block b = free_blocks[best]; free_blocks.erase(free_blocks.begin() + best); return b; }
		}
		Block b;
		b.bytes = (bytes + (bytes >> 2) + 4095) & ~4095ull; // some slack: graphs of one workload differ a little in size
		if (cudaHostAlloc(&b.ptr, b.bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); throw cuda_error("cannot page-lock host memory for the graph arrays"); }
		return b;
	}
	void give(Block b) {
		if (!b.ptr) return;
		Block drop;
		{
			std::lock_guard<std::mutex> lk(mu);
			free_blocks.push_back(b);
			if (free_blocks.size() > KEEP) { // let the smallest go
				size_t s = 0;
				for (size_t i = 1; i < free_blocks.size(); i++) if (free_blocks[i].bytes < free_blocks[s].bytes) s = i;
				drop = free_blocks[s]; free_blocks.erase(free_blocks.begin() + s);
			}
		}
		if (drop.ptr) cudaFreeHost(drop.ptr);
	}
};

struct Graph {
	// host-side result of K4: views into one page-locked block
	PinnedPool::Block block;
	u64 *heights = nullptr;     u64 n_heights = 0;  // per block
	u64 *node_block = nullptr;  u64 n_nodes = 0;    // per node
	u64 *node_row = nullptr;                        // representative (first-occurrence) row
	u64 *edge_src = nullptr, *edge_dst = nullptr; u64 n_edges = 0;
	u32 *node_of = nullptr;     u64 n_node_of = 0;  // [nb][m], U32_MAX = empty label (only kept when requested)
	u64 label_total = 0;
	Graph() = default;
	Graph(const Graph &) = delete;
	Graph &operator=(const Graph &) = delete;
	~Graph() { PinnedPool::get().give(block); }
};

// device-side view of the K4 result while it is still in the scratch arena (input of K5)
struct GraphDev {
	u64 m = 0, nb = 0, nn = 0, ne = 0;
	const u32 *heights = nullptr;    // [nb+1], heights[nb] = 0
	const u32 *node_block = nullptr, *node_row = nullptr; // [nn]
	const u32 *es = nullptr, *ed = nullptr;               // [ne], grouped by destination block
	const u32 *EOFF = nullptr;       // [nb] edges arriving in the blocks before j
	const u32 *node_of = nullptr;    // [nb][m]
	const u32 *LEN = nullptr, *POS = nullptr;             // [nb][m] label length / text position
	u64 label_total = 0;
};

// K5 request: format the xGFA of the graph on the device (emit.cu); the bytes stay in
// Pipeline::xgfa until they are read or saved
struct XgfaJob {
	bool paths = false;
	const char *const *ids = nullptr; // m header strings when paths
	// results
	u64 n_nodes = 0, n_edges = 0, label_total = 0, max_height = 0, n_bytes = 0;
};

struct Pipeline;
// repair loop of --heuristic-subset (validate.cu): one efg_validate pass on the graph K4 has just built
struct ValidateJob {
	Pipeline *aux = nullptr;   // auxiliary pipeline that indexes the edge text (own arenas, own stream)
	u64 n_removed = 0;         // boundaries dropped by this pass
	u64 edge_text_len = 0;     // characters of the edge text it indexed
	u64 n_nodes = 0, n_edges = 0, label_total = 0; // of the graph that was checked ("#nodes=", "total length of node labels=", "#edges=", fbg.cpp:1119-1120,1172)
};

struct Pipeline {
	Dev dev;
	Dist dist;
	Options opt;
	u64 m = 0, n = 0, pitch = 0, W = 0, N = 0;
	u8 *msa = nullptr;
	u32 *NG = nullptr, *IG = nullptr, *RK = nullptr, *NIA = nullptr;
	u32 *start = nullptr, *rowlen = nullptr, *lastcol = nullptr;
	u8 *T = nullptr;
	u32 *col = nullptr, *ISA = nullptr, *SA = nullptr, *PLCP = nullptr;
	u8 *BWT = nullptr;     // [N] T[SA[k] - 1], written with the SA entries of a single-GPU full index (suffix.cu SaOut)
	bool have_bwt = false;
	u32 *f = nullptr, *bnd = nullptr;
	u32 *dp_a = nullptr, *dp_b = nullptr; // mml/backtrack or s/prev (+ v)
	u32 *dp_v = nullptr;
	u64 *scalars = nullptr;               // small device scratch for totals
	u64 *h_scalars = nullptr;             // pinned mirror
	u64 nb = 0, score = 0;
	u64 gaps = 0, toignore = 0;
	bool keep_sa = false;
	int sa_rounds = 0;
	u64 dp_chunk = 0, dp_warm = 0; // 0 = defaults (segdp.cu); settable for the parity tests
	int kmer_dict = 0;             // round 0 of the last index build went through the k-mer dictionary
	int dp_passes = 0;             // verify passes of the last DP (0 = single chunk, < 0 = fell back to the whole-range wavefront)
	bool have_msa = false, have_index = false, have_f = false, have_seg = false;
	int seg_status = 0; // 0 ok, 1 "no segmentation"
	// K1 may stop early on gapless MSAs (suffix.cu, "sufficient index"): SA / ISA / PLCP are then exact
	// everywhere except inside the column-pure buckets left at depth lazy_depth.  fbg_build_index
	// and the read-back calls set want_full_index and get the true arrays.
	bool want_full_index = false, index_lazy = false;
	u64 lazy_depth = 0;
	// ... and when the k-mer round itself is column-pure (and no ignore characters are set) K2 needs
	// no per-position array at all: the sorted list of DISTINCT k-mers with their column and the
	// LCP to the k-mer before is the whole index ("virtual"; fcol.cu k2_from_dict)
	bool index_virtual = false, force_arrays = false;
	// --heuristic-subset (fbg_heuristic_f): f is the maximum over row chunks and there is no index of the whole MSA
	bool f_external = false;
	bool no_peer = false; // multi-GPU: the peer-mode attempt of this build was given up, use the key-range form (suffix.cu)
	u32 *dict_col = nullptr, *dict_lcp = nullptr; // [dict_nd], [dict_nd + 1] in rank order
	u64 dict_nd = 0;
	// K5: formatted xGFA bytes (own allocation, grown on demand, kept across builds)
	u8 *xgfa = nullptr;
	u64 xgfa_cap = 0, xgfa_len = 0;
	bool have_xgfa = false;

	void free_all();
};

// phases (each enqueues on p.dev.stream; those that need a host decision synchronise)
void k0_upload_host(Pipeline &p, const u8 *rows, u64 m, u64 n);
void k0_upload_device(Pipeline &p, const u8 *drows, u64 m, u64 n);
void k0_pack(Pipeline &p);
void k1_suffix_array(Pipeline &p);
void k1_lcp(Pipeline &p);
void k2_compute_f(Pipeline &p);
void k3_segment(Pipeline &p);
// g: host copy of the graph arrays (may be null); x: also format the xGFA on the device (may be null)
void k4_graph(Pipeline &p, Graph *g, bool want_node_of, XgfaJob *x = nullptr, ValidateJob *v = nullptr);
void k6_validate(Pipeline &p, const GraphDev &gd, ValidateJob &v);
u64 k6_scratch_need(u64 m, u64 nb);
void k5_emit_xgfa(Pipeline &p, const GraphDev &gd, XgfaJob &x);
// scratch bytes K5 needs on top of K4's (R = m * nb records)
u64 k5_scratch_need(u64 m, u64 nb, bool paths, u64 id_bytes);

// bytes of the persistent arena one m x n MSA can need (N <= m*n + m + 1)
static inline u64 persistent_need(u64 m, u64 n) {
	u64 pitch = round_up(n, 32), W = pitch / 32, N = m * n + m + 1;
	return m * pitch + 16 * W * m + 12 * (m + 1) + (N + 64) + 4 * 4 * (N + 1) + 6 * 4 * (n + 2) + 64 * 256 + N + (N + 64) + (1u << 19) + 64;
}

} // namespace fbg
