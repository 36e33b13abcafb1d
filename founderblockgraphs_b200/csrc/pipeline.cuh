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
#pragma once
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

struct Graph {
	// host-side result of K4
	std::vector<u64> heights;    // per block
	std::vector<u64> node_block; // per node
	std::vector<u64> node_row;   // representative (first-occurrence) row
	std::vector<u64> edge_src, edge_dst;
	std::vector<u32> node_of;    // [nb][m], U32_MAX = empty label (only kept when requested)
	u64 label_total = 0;
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
void k4_graph(Pipeline &p, Graph &g, bool want_node_of);

// bytes of the persistent arena one m x n MSA can need (N <= m*n + m + 1)
static inline u64 persistent_need(u64 m, u64 n) {
	u64 pitch = round_up(n, 32), W = pitch / 32, N = m * n + m + 1;
	return m * pitch + 16 * W * m + 12 * (m + 1) + (N + 64) + 4 * 4 * (N + 1) + 6 * 4 * (n + 2) + 64 * 256;
}

} // namespace fbg
