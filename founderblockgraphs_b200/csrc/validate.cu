// validate.cu -- the repair loop of --heuristic-subset: one pass of efg_validate
// (fbg.cpp:3199-3292, efg_validate_node 3104-3170) over the graph of the current segmentation.
//
// The reference indexes the concatenation of all edge strings  label(u) label(v) '#'  with an sdsl csa_wt
// (fbg.cpp:3222-3246), looks every node label up in it by backward search (3125) and walks the occurrences
// (3130-3166): a node that is neither source nor sink and holds no ignore character is INVALID when some
// occurrence does not start exactly at the first character of u or of v with that node in the node's own
// block; its block b is then merged with its predecessor (boundary b - 1 dropped, 3270-3271).
//
// Here the edge text is written on the device from K4's arrays (labels are slices of the gap-free text T),
// its suffix array comes from the same K1 that indexes the MSA -- run on an auxiliary pipeline whose text is
// the edge text -- and one warp per node finds its SA interval by two binary searches (32 characters per
// comparison step) and classifies the occurrences.  Boundaries are then compacted on the device.
#include "pipeline.cuh"
#include <cstring>

namespace fbg {

// (position in T, length) of every node's label; edge string lengths
__global__ void k6_node_labels_kernel(const u32 *__restrict__ node_block, const u32 *__restrict__ node_row, const u32 *__restrict__ LEN, const u32 *__restrict__ POS,
                                      u64 m, u32 nn, u32 *__restrict__ npos, u32 *__restrict__ nlen) {
	u32 w = blockIdx.x * blockDim.x + threadIdx.x;
	if (w >= nn) return;
	const u64 rec = (u64) node_block[w] * m + node_row[w];
	npos[w] = POS[rec]; nlen[w] = LEN[rec];
}
__global__ void k6_edge_len_kernel(const u32 *__restrict__ es, const u32 *__restrict__ ed, const u32 *__restrict__ nlen, u32 ne, u64 *__restrict__ elen,
                                   u8 *__restrict__ has_in, u8 *__restrict__ has_out) {
	u32 e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e >= ne) return;
	const u32 u = es[e], v = ed[e];
	elen[e] = (u64) nlen[u] + nlen[v] + 1;   // label(u) label(v) '#'  (fbg.cpp:3238)
	has_out[u] = 1; has_in[v] = 1;           // is_sink / is_source (fbg.cpp:3252-3257)
}
// one warp per edge copies the two labels out of T
__global__ void __launch_bounds__(256) k6_edge_text_kernel(const u8 *__restrict__ T, const u32 *__restrict__ es, const u32 *__restrict__ ed, const u32 *__restrict__ npos,
                                                           const u32 *__restrict__ nlen, const u64 *__restrict__ estart, u32 ne, u64 NE, u8 *__restrict__ E) {
	const u64 wg = ((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const u32 l = lane_id();
	if (wg == 0 && l == 0) E[NE - 1] = 0; // the sentinel sdsl appends (construct.hpp:142-147)
	if (wg >= ne) return;
	const u32 u = es[wg], v = ed[wg];
	const u64 o = estart[wg];
	const u32 lu = nlen[u], lv = nlen[v];
	for (u32 c = l; c < lu; c += 32) E[o + c] = T[(u64) npos[u] + c];
	for (u32 c = l; c < lv; c += 32) E[o + lu + c] = T[(u64) npos[v] + c];
	if (l == 0) E[o + lu + lv] = (u8) '#';
}

struct ValArgs {
	const u8 *T, *E; const u32 *SA; u64 NE;
	const u32 *npos, *nlen, *node_block, *es, *ed;
	const u64 *estart; u32 nn, ne;
	const u8 *has_in, *has_out;
	IgnoreSet ign; int has_ignore;
	u32 *drop; // [nb] 1 = boundary to drop
};

// Compares the label L (len characters at T + lp) with the suffix of E at q, warp-wide: < 0 label smaller,
// 0 label is a prefix of the suffix, > 0 label greater
__device__ __forceinline__ int k6_cmp(const ValArgs &a, u64 lp, u32 len, u64 q) {
	const u32 l = lane_id();
	for (u32 c0 = 0; c0 < len; c0 += 32) {
		const u32 c = c0 + l;
		int d = 0;
		if (c < len) {
			const int x = a.T[lp + c];
			const int y = q + c < a.NE ? a.E[q + c] : 0;
			d = x - y;
		}
		const u32 mm = __ballot_sync(0xffffffffu, d != 0);
		if (mm) return __shfl_sync(0xffffffffu, d, __ffs(mm) - 1);
	}
	return 0;
}

__global__ void __launch_bounds__(256) k6_validate_kernel(ValArgs a) {
	const u32 w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const u32 l = lane_id();
	if (w >= a.nn) return;
	if (!a.has_in[w] || !a.has_out[w]) return;               // fbg.cpp:3116-3117
	const u64 lp = a.npos[w];
	const u32 len = a.nlen[w];
	if (a.has_ignore) {                                      // fbg.cpp:3119-3120
		bool any = false;
		for (u32 c = l; c < len; c += 32) { const u8 ch = a.T[lp + c]; any = any || ((a.ign.bits[ch >> 5] >> (ch & 31)) & 1u); }
		if (__any_sync(0xffffffffu, any)) return;
	}
	// SA interval of the label: first rank whose suffix is >= L, first rank whose suffix is > L and lacks the prefix
	u64 lo = 0, hi = a.NE;
	while (lo < hi) { const u64 mid = (lo + hi) >> 1; if (k6_cmp(a, lp, len, a.SA[mid]) > 0) lo = mid + 1; else hi = mid; }
	const u64 first = lo;
	hi = a.NE;
	while (lo < hi) { const u64 mid = (lo + hi) >> 1; if (k6_cmp(a, lp, len, a.SA[mid]) >= 0) lo = mid + 1; else hi = mid; }
	const u64 last = lo; // exclusive
	const u32 bw = a.node_block[w];
	bool bad = false;
	for (u64 k0 = first; k0 < last && !bad; k0 += 32) {        // fbg.cpp:3130-3166
		const u64 k = k0 + l;
		bool inv = false;
		if (k < last) {
			const u64 occ = a.SA[k];
			u32 x = 0, y = a.ne; // the edge whose string holds occ: last e with estart[e] <= occ
			while (y - x > 1) { const u32 z = (x + y) >> 1; if (a.estart[z] <= occ) x = z; else y = z; }
			const u64 off = occ - a.estart[x];
			const u32 u = a.es[x], v = a.ed[x];
			const u32 lu = a.nlen[u];
			const bool ok = off < lu ? (off == 0 && a.node_block[u] == bw) : (off == lu && a.node_block[v] == bw);
			inv = !ok;
		}
		bad = __any_sync(0xffffffffu, inv);
	}
	if (bad && bw > 0 && l == 0) a.drop[bw - 1] = 1u;           // fbg.cpp:3270-3271
}

__global__ void k6_keep_flags_kernel(const u32 *__restrict__ drop, u32 nb, u32 *__restrict__ keep) {
	u32 j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j < nb) keep[j] = drop[j] ? 0u : 1u;
}
__global__ void k6_compact_bnd_kernel(const u32 *__restrict__ bnd, const u32 *__restrict__ keep, const u32 *__restrict__ pos, u32 nb, u32 *__restrict__ out) {
	u32 j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j < nb && keep[j]) out[pos[j]] = bnd[j];
}

u64 k6_scratch_need(u64 m, u64 nb) {
	const u64 R = m * nb;
	return R * 40 + (nb + 1) * 16 + (R / SCAN_TILE + 2) * 48 + (4u << 20);
}

// One pass of the repair loop on the graph K4 has just built (gd views K4's arrays in p's scratch arena).
// Leaves the surviving boundaries in p.bnd / p.nb and the number dropped in v.n_removed.
void k6_validate(Pipeline &p, const GraphDev &gd, ValidateJob &v) {
	Dev &d = p.dev;
	const u64 m = gd.m, nb = gd.nb, nn = gd.nn, ne = gd.ne;
	v.n_removed = 0; v.edge_text_len = 0;
	if (ne == 0 || nb < 2 || nn == 0) return; // no node with both an edge in and an edge out
	u64 mk = d.mark();
	u32 *npos = d.alloc<u32>(nn), *nlen = d.alloc<u32>(nn);
	u64 *estart = d.alloc<u64>(ne + 1);
	u8 *has_in = d.alloc<u8>(nn), *has_out = d.alloc<u8>(nn);
	u32 *drop = d.alloc<u32>(nb), *keep = d.alloc<u32>(nb), *pos = d.alloc<u32>(nb), *nbnd = d.alloc<u32>(nb);
	u64 *tot64 = p.scalars + 36;
	u32 *tot32 = reinterpret_cast<u32 *>(p.scalars + 37);
	FBG_CUDA(cudaMemsetAsync(has_in, 0, nn, d.stream));
	FBG_CUDA(cudaMemsetAsync(has_out, 0, nn, d.stream));
	FBG_CUDA(cudaMemsetAsync(drop, 0, nb * sizeof(u32), d.stream));
	d.begin(K_GRAPH);
	k6_node_labels_kernel<<<div_up(nn, 256), 256, 0, d.stream>>>(gd.node_block, gd.node_row, gd.LEN, gd.POS, m, (u32) nn, npos, nlen);
	k6_edge_len_kernel<<<div_up(ne, 256), 256, 0, d.stream>>>(gd.es, gd.ed, nlen, (u32) ne, estart, has_in, has_out);
	FBG_LAUNCH_CHECK();
	d.end((double) nn * 16.0 + (double) ne * 24.0);
	d.note_launches(K_GRAPH, 1);
	device_scan<OpSumU64, true>(d, estart, estart, ne, tot64);
	FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 36, tot64, sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	const u64 NE = p.h_scalars[36] + 1; // + the final sentinel
	v.edge_text_len = NE;
	if (NE >= 0xFFFFFFF0ull) throw unsupported_error("the edge strings of this graph hold " + std::to_string(NE) + " characters: more than 32-bit suffix indices address");

	// the auxiliary pipeline: its "text" is the edge text, K1 gives the suffix array
	Pipeline &a = *v.aux;
	a.dev.device = d.device;
	if (!a.dev.stream) FBG_CUDA(cudaStreamCreateWithFlags(&a.dev.stream, cudaStreamNonBlocking));
	if (!a.scalars) FBG_CUDA(cudaMalloc(&a.scalars, 64 * sizeof(u64)));
	if (!a.h_scalars) FBG_CUDA(cudaMallocHost(&a.h_scalars, 64 * sizeof(u64)));
	a.free_all();
	const u64 pers = (NE + 64) * 2 + 3 * 4 * (NE + 1) + (1u << 16);
	if (pers > a.dev.pers_cap) FBG_CUDA(cudaStreamSynchronize(a.dev.stream));
	a.dev.reserve_pers(pers);
	a.m = 1; a.n = NE; a.N = NE; a.gaps = 1; // (gaps != 0: never the sufficient index)
	a.want_full_index = true;
	a.T = a.dev.palloc<u8>(NE + 64);
	FBG_CUDA(cudaMemsetAsync(a.T + NE, 0, 64, d.stream));
	d.begin(K_GRAPH);
	k6_edge_text_kernel<<<div_up(ne * 32, 256), 256, 0, d.stream>>>(p.T, gd.es, gd.ed, npos, nlen, estart, (u32) ne, NE, a.T);
	FBG_LAUNCH_CHECK();
	d.end((double) NE * 2.0);
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	a.have_msa = true;
	k1_suffix_array(a);
	FBG_CUDA(cudaStreamSynchronize(a.dev.stream));
	a.dev.resolve();
	for (int k = 0; k < K_COUNT; k++) { // the index of the edge text is part of this build's accounting
		d.stats[k].ms += a.dev.stats[k].ms; d.stats[k].bytes += a.dev.stats[k].bytes; d.stats[k].launches += a.dev.stats[k].launches;
	}
	a.dev.reset_stats();

	ValArgs va;
	va.T = p.T; va.E = a.T; va.SA = a.SA; va.NE = NE;
	va.npos = npos; va.nlen = nlen; va.node_block = gd.node_block; va.es = gd.es; va.ed = gd.ed;
	va.estart = estart; va.nn = (u32) nn; va.ne = (u32) ne;
	va.has_in = has_in; va.has_out = has_out;
	for (int q = 0; q < 8; q++) va.ign.bits[q] = p.opt.ignore_bits[q];
	va.has_ignore = p.opt.n_ignore > 0 ? 1 : 0;
	va.drop = drop;
	d.begin(K_GRAPH);
	k6_validate_kernel<<<div_up(nn * 32, 256), 256, 0, d.stream>>>(va);
	k6_keep_flags_kernel<<<div_up(nb, 256), 256, 0, d.stream>>>(drop, (u32) nb, keep);
	FBG_LAUNCH_CHECK();
	d.end((double) nn * 64.0);
	d.note_launches(K_GRAPH, 1);
	device_scan<OpSumU32, true>(d, keep, pos, nb, tot32);
	k6_compact_bnd_kernel<<<div_up(nb, 256), 256, 0, d.stream>>>(p.bnd, keep, pos, (u32) nb, nbnd);
	FBG_LAUNCH_CHECK();
	FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 37, tot32, sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	const u64 kept = *reinterpret_cast<u32 *>(p.h_scalars + 37);
	if (kept < nb) {
		FBG_CUDA(cudaMemcpyAsync(p.bnd, nbnd, kept * sizeof(u32), cudaMemcpyDeviceToDevice, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
	}
	v.n_removed = nb - kept;
	p.nb = kept;
	d.release(mk);
}

} // namespace fbg
