// graph.cu -- K4: founder nodes and inter-block edges from the segmentation.
//
// Replaces the std::unordered_map<std::string,...> / std::set<pair> logic of
// output_efg (fbg.cpp:1222-1260) and the node/edge statistics of segment()
// (fbg.cpp:666-728).  A row's label in block [a,b] is the contiguous text range
// T[start_i + rank_i(a) .. start_i + rank_i(b+1)), so:
//   1. 128-bit polynomial hash (+ length) of every (block,row) label
//   2. per block: bitonic sort of (hash prefix,row) in shared memory, group heads
//      = first-occurrence rows, every row learns its representative row
//   3. exactness guard: every row's label is compared byte-wise with its
//      representative's; any mismatch (a hash collision) re-runs with a new seed
//   4. node ids = exclusive scan of head flags in (block,row) order  -> the
//      reference's "first occurrence in row order inside the block" numbering
//   5. edges (node of row i in block j-1, node in block j): per-block sort/unique,
//      emitted in ascending (src,dst) order like std::set<pair>
#include "pipeline.cuh"
#include "blocksort.cuh"
#include <cstring>

namespace fbg {

struct GraphArgs {
	const u32 *bnd, *NG, *RK, *start, *rowlen;
	const u8 *T;
	u64 m, n, nb;
	u64 seed;
};

__device__ __forceinline__ u32 rank_at(const GraphArgs &a, u64 i, u64 x) {
	if (x >= a.n) return a.rowlen[i];
	u64 t = x >> 5;
	u32 below = (1u << (u32) (x & 31)) - 1u;
	return a.RK[t * a.m + i] + __popc(a.NG[t * a.m + i] & below);
}
__device__ __forceinline__ void block_cols(const GraphArgs &a, u64 j, u64 &ca, u64 &cb1) {
	ca = j ? (u64) a.bnd[j - 1] + 1 : 0;
	u64 e = a.bnd[j];
	cb1 = e >= a.n ? a.n : e + 1; // exclusive end, clamped like std::string::substr
}

// thread per (row, block), block index fastest: neighbouring threads read neighbouring text
__global__ void k4_hash_kernel(GraphArgs a, u64 *__restrict__ H1, u64 *__restrict__ H2, u32 *__restrict__ LEN, u32 *__restrict__ POS) {
	u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (gid >= a.m * a.nb) return;
	u64 i = gid / a.nb, j = gid % a.nb;
	u64 ca, cb1;
	block_cols(a, j, ca, cb1);
	u32 r0 = rank_at(a, i, ca), r1 = rank_at(a, i, cb1);
	u64 p0 = (u64) a.start[i] + r0;
	u32 len = r1 - r0;
	u64 h1 = 0x9E3779B97F4A7C15ull ^ a.seed, h2 = 0xC2B2AE3D27D4EB4Full + a.seed * 0x100000001B3ull;
	for (u32 q = 0; q < len; q++) {
		u64 c = a.T[p0 + q];
		h1 = h1 * 0x100000001B3ull + c + 1;
		h2 = (h2 ^ c) * 0xD6E8FEB86659FD93ull;
		h2 ^= h2 >> 29;
	}
	h1 ^= h1 >> 31; h1 *= 0xBF58476D1CE4E5B9ull; h1 ^= h1 >> 29;
	u64 o = j * a.m + i;
	H1[o] = h1; H2[o] = h2 + len; LEN[o] = len; POS[o] = (u32) p0;
}

// per block group: REP[j*m+i] = first row with the same label (U32_MAX: empty), HEAD[j*m+i] = 1 for first occurrences
__global__ void __launch_bounds__(1024) k4_group_kernel(u64 m, u64 nb, u32 mpad, u32 tpg, u32 groups,
                                                        const u64 *__restrict__ H1, const u64 *__restrict__ H2, const u32 *__restrict__ LEN,
                                                        u32 *__restrict__ REP, u32 *__restrict__ HEAD, u32 *__restrict__ heights) {
	extern __shared__ __align__(16) unsigned char k4_raw[];
	const u32 grp = threadIdx.x / tpg, tg = threadIdx.x % tpg;
	unsigned char *basep = k4_raw + (u64) grp * mpad * 12;
	u64 *key = reinterpret_cast<u64 *>(basep);
	u32 *hd = reinterpret_cast<u32 *>(basep + (u64) mpad * 8);
	for (u64 jb = (u64) blockIdx.x * groups; jb < nb; jb += (u64) gridDim.x * groups) {
		const u64 j = jb + grp;
		const bool live = j < nb;
		for (u32 i = tg; i < mpad; i += tpg) {
			u64 kk = ~0ull;
			if (live && i < m && LEN[j * m + i] != 0) kk = ((H1[j * m + i] >> 24) << 24) | i;
			key[i] = kk;
		}
		__syncthreads();
		group_bitonic_sort(key, mpad, tg, tpg);
		// head detection on the full (H1,H2,LEN) triple; hd[q] = q for heads else 0
		for (u32 q = tg; q < mpad; q += tpg) {
			u64 kq = key[q];
			u32 h = 0;
			if (kq != ~0ull && q > 0) {
				u64 o = j * m + (kq & 0xFFFFFFu), op = j * m + (key[q - 1] & 0xFFFFFFu);
				bool same = H1[o] == H1[op] && H2[o] == H2[op] && LEN[o] == LEN[op];
				h = same ? 0 : q;
			}
			hd[q] = h;
		}
		__syncthreads();
		// inclusive max-scan of hd (Hillis-Steele)
		for (u32 o = 1; o < mpad; o <<= 1) {
			u32 nv[8]; u32 c = 0;
			for (u32 q = tg; q < mpad; q += tpg, c++) { u32 v = hd[q]; if (q >= o) { u32 w = hd[q - o]; v = w > v ? w : v; } nv[c] = v; }
			__syncthreads();
			c = 0;
			for (u32 q = tg; q < mpad; q += tpg, c++) hd[q] = nv[c];
			__syncthreads();
		}
		u32 nheads = 0;
		for (u32 q = tg; q < mpad; q += tpg) {
			u64 kq = key[q];
			if (kq == ~0ull) continue;
			u32 i = (u32) (kq & 0xFFFFFFu);
			u32 hq = hd[q];
			u32 rep = (u32) (key[hq] & 0xFFFFFFu);
			REP[j * m + i] = rep;
			HEAD[j * m + i] = (hq == q) ? 1u : 0u;
			nheads += (hq == q);
		}
		if (live) for (u32 i = tg; i < m; i += tpg) if (LEN[j * m + i] == 0) { REP[j * m + i] = U32_MAX; HEAD[j * m + i] = 0; }
		nheads = warp_reduce_sum(nheads);
		if (lane_id() == 0 && nheads && live) atomicAdd(&heights[j], nheads);
		__syncthreads();
	}
}

// exactness guard: label(i) must equal label(rep) byte for byte
__global__ void k4_verify_kernel(GraphArgs a, const u32 *__restrict__ REP, const u32 *__restrict__ LEN, const u32 *__restrict__ POS, u32 *__restrict__ bad) {
	u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (gid >= a.m * a.nb) return;
	u64 j = gid / a.m, i = gid % a.m;
	u32 rep = REP[gid];
	if (rep == U32_MAX || rep == i) return;
	u64 o2 = j * a.m + rep;
	u32 len = LEN[gid];
	bool ok = len == LEN[o2] && rep < i;
	u64 p = POS[gid], q = POS[o2];
	for (u32 c = 0; ok && c < len; c++) ok = a.T[p + c] == a.T[q + c];
	if (!ok) *bad = 1;
}

// node id of every (block,row); node arrays for heads; label length total
__global__ void k4_nodes_kernel(u64 m, u64 nb, const u32 *__restrict__ REP, const u32 *__restrict__ HEAD, const u32 *__restrict__ NIDX, const u32 *__restrict__ LEN,
                                u32 *__restrict__ node_of, u32 *__restrict__ node_block, u32 *__restrict__ node_row, u64 *__restrict__ label_total) {
	u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	u64 lt = 0;
	if (gid < m * nb) {
		u64 j = gid / m, i = gid % m;
		u32 rep = REP[gid];
		u32 id = U32_MAX;
		if (rep != U32_MAX) id = NIDX[j * m + rep];
		node_of[gid] = id;
		if (HEAD[gid]) { node_block[id] = (u32) j; node_row[id] = (u32) i; lt = LEN[gid]; }
	}
	lt = warp_reduce_sum(lt);
	if (lane_id() == 0 && lt) atomicAdd((unsigned long long *) label_total, (unsigned long long) lt);
}

// per block j >= 1: sorted unique (src,dst) pairs -> PAIRS[j*m + q], EC[j] = count
__global__ void __launch_bounds__(1024) k4_edges_kernel(u64 m, u64 nb, u32 mpad, u32 tpg, u32 groups, const u32 *__restrict__ node_of,
                                                        u64 *__restrict__ PAIRS, u32 *__restrict__ EC) {
	extern __shared__ __align__(16) unsigned char k4_raw[];
	const u32 grp = threadIdx.x / tpg, tg = threadIdx.x % tpg;
	unsigned char *basep = k4_raw + (u64) grp * mpad * 12;
	u64 *key = reinterpret_cast<u64 *>(basep);
	u32 *hd = reinterpret_cast<u32 *>(basep + (u64) mpad * 8);
	for (u64 jb = 1 + (u64) blockIdx.x * groups; jb < nb; jb += (u64) gridDim.x * groups) {
		const u64 j = jb + grp;
		const bool live = j < nb;
		for (u32 i = tg; i < mpad; i += tpg) {
			u64 kk = ~0ull;
			if (live && i < m) {
				u32 s = node_of[(j - 1) * m + i], t = node_of[j * m + i];
				if (s != U32_MAX && t != U32_MAX) kk = ((u64) s << 32) | t;
			}
			key[i] = kk;
		}
		__syncthreads();
		group_bitonic_sort(key, mpad, tg, tpg);
		// unique flags -> exclusive positions (Hillis-Steele inclusive sum)
		for (u32 q = tg; q < mpad; q += tpg) {
			u64 kq = key[q];
			hd[q] = (kq != ~0ull && (q == 0 || key[q - 1] != kq)) ? 1u : 0u;
		}
		__syncthreads();
		for (u32 o = 1; o < mpad; o <<= 1) {
			u32 nv[8]; u32 c = 0;
			for (u32 q = tg; q < mpad; q += tpg, c++) { u32 v = hd[q]; if (q >= o) v += hd[q - o]; nv[c] = v; }
			__syncthreads();
			c = 0;
			for (u32 q = tg; q < mpad; q += tpg, c++) hd[q] = nv[c];
			__syncthreads();
		}
		for (u32 q = tg; q < mpad; q += tpg) {
			u64 kq = key[q];
			bool uniq = kq != ~0ull && (q == 0 || key[q - 1] != kq);
			if (uniq && live) PAIRS[j * m + hd[q] - 1] = kq;
		}
		if (tg == 0 && live) EC[j] = hd[mpad - 1];
		__syncthreads();
	}
}

__global__ void k4_edge_gather_kernel(u64 m, u64 nb, const u64 *__restrict__ PAIRS, const u32 *__restrict__ EC, const u32 *__restrict__ EOFF, u32 *__restrict__ es, u32 *__restrict__ ed) {
	u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (gid >= m * nb) return;
	u64 j = gid / m, q = gid % m;
	if (q >= EC[j]) return;
	u64 pr = PAIRS[gid];
	u64 o = (u64) EOFF[j] + q;
	es[o] = (u32) (pr >> 32); ed[o] = (u32) pr;
}

__global__ void k4_widen_kernel(const u32 *__restrict__ src, u64 *__restrict__ dst, u64 count) {
	u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < count) dst[i] = src[i];
}

// the C ABI hands out 64-bit arrays: widen on the device, one copy into the page-locked block
// (stream order keeps the staging buffer safe between two arrays; the caller synchronises once)
static void d2h(u64 *dst, const u32 *src, u64 count, u64 *wide, cudaStream_t s) {
	if (!count) return;
	k4_widen_kernel<<<div_up(count, 256), 256, 0, s>>>(src, wide, count);
	FBG_CUDA(cudaMemcpyAsync(dst, wide, count * sizeof(u64), cudaMemcpyDeviceToHost, s));
}

void k4_graph(Pipeline &p, Graph *gp, bool want_node_of, XgfaJob *x, ValidateJob *v) {
	Dev &d = p.dev;
	const u64 m = p.m, nb = p.nb, R = m * nb;
	if (R >= 0xFFFFFFF0ull) throw unsupported_error("more than 2^32 (row,block) records");
	GroupCfg gc = group_cfg(m);
	if (gc.mpad > 8192) throw unsupported_error("more than 8192 rows is not supported by the shared-memory dedup kernel");
	u64 id_bytes = 0;
	if (x && x->paths) for (u64 i = 0; i < m; i++) id_bytes += strlen(x->ids[i]);
	d.ensure(R * 72 + (nb + 1) * 12 + (R / SCAN_TILE + 2) * 24 + (4u << 20) + (x ? k5_scratch_need(m, nb, x->paths, id_bytes) : 0) + (v ? k6_scratch_need(m, nb) : 0));
	u64 mk = d.mark();
	u64 *H1 = d.alloc<u64>(R), *H2 = d.alloc<u64>(R);
	u32 *LEN = d.alloc<u32>(R), *POS = d.alloc<u32>(R), *REP = d.alloc<u32>(R), *HEAD = d.alloc<u32>(R), *NIDX = d.alloc<u32>(R);
	u32 *node_of = d.alloc<u32>(R), *node_block = d.alloc<u32>(R), *node_row = d.alloc<u32>(R);
	u32 *heights = d.alloc<u32>(nb + 1), *EC = d.alloc<u32>(nb + 1), *EOFF = d.alloc<u32>(nb + 1);
	u64 *PAIRS = d.alloc<u64>(R);
	u32 *es = d.alloc<u32>(R), *ed = d.alloc<u32>(R);
	u64 *wide = d.alloc<u64>(R > nb ? R : nb + 1);
	u32 *bad = reinterpret_cast<u32 *>(p.scalars + 32);
	u32 *tot_nodes = reinterpret_cast<u32 *>(p.scalars + 33);
	u32 *tot_edges = reinterpret_cast<u32 *>(p.scalars + 34);
	u64 *label_total = p.scalars + 35;

	GraphArgs a;
	a.bnd = p.bnd; a.NG = p.NG; a.RK = p.RK; a.start = p.start; a.rowlen = p.rowlen; a.T = p.T;
	a.m = m; a.n = p.n; a.nb = nb;
	size_t smem = (size_t) gc.groups * gc.mpad * 12;
	FBG_CUDA(cudaFuncSetAttribute(k4_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	FBG_CUDA(cudaFuncSetAttribute(k4_edges_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	u64 bgroups = (nb + gc.groups - 1) / gc.groups;
	u32 grid = (u32) (bgroups < (u64) NUM_SMS * 8 ? bgroups : (u64) NUM_SMS * 8);
	u32 gR = div_up(R, 256);
	for (int attempt = 0;; attempt++) {
		a.seed = 0x5851F42D4C957F2Dull * (u64) attempt;
		FBG_CUDA(cudaMemsetAsync(p.scalars + 32, 0, 8 * sizeof(u64), d.stream));
		FBG_CUDA(cudaMemsetAsync(heights, 0, (nb + 1) * sizeof(u32), d.stream));
		d.begin(K_GRAPH);
		k4_hash_kernel<<<gR, 256, 0, d.stream>>>(a, H1, H2, LEN, POS);
		k4_group_kernel<<<grid, gc.threads, smem, d.stream>>>(m, nb, gc.mpad, gc.tpg, gc.groups, H1, H2, LEN, REP, HEAD, heights);
		k4_verify_kernel<<<gR, 256, 0, d.stream>>>(a, REP, LEN, POS, bad);
		FBG_LAUNCH_CHECK();
		d.end((double) p.N * 2.0 + (double) R * 60.0);
		d.note_launches(K_GRAPH, 2);
		FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 32, p.scalars + 32, sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		if (*reinterpret_cast<u32 *>(p.h_scalars + 32) == 0) break;
		if (attempt >= 3) throw cuda_error("label hashing failed its exactness check four times");
	}
	device_scan<OpSumU32, true>(d, HEAD, NIDX, R, tot_nodes);
	FBG_CUDA(cudaMemsetAsync(EC, 0, (nb + 1) * sizeof(u32), d.stream));
	d.begin(K_EDGES);
	k4_nodes_kernel<<<gR, 256, 0, d.stream>>>(m, nb, REP, HEAD, NIDX, LEN, node_of, node_block, node_row, label_total);
	if (nb > 1) k4_edges_kernel<<<grid, gc.threads, smem, d.stream>>>(m, nb, gc.mpad, gc.tpg, gc.groups, node_of, PAIRS, EC);
	FBG_LAUNCH_CHECK();
	d.end((double) R * 40.0);
	d.note_launches(K_EDGES, 1);
	device_scan<OpSumU32, true>(d, EC, EOFF, nb, tot_edges);
	d.begin(K_EDGES);
	k4_edge_gather_kernel<<<gR, 256, 0, d.stream>>>(m, nb, PAIRS, EC, EOFF, es, ed);
	FBG_LAUNCH_CHECK();
	d.end((double) R * 12.0);
	FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 32, p.scalars + 32, 8 * sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	u64 nn = *reinterpret_cast<u32 *>(p.h_scalars + 33), ne = *reinterpret_cast<u32 *>(p.h_scalars + 34);
	const u64 label_sum = p.h_scalars[35];
	if (gp) {
		Graph &g = *gp;
		g.label_total = label_sum;
		const u64 words = nb + 2 * nn + 2 * ne;
		g.block = PinnedPool::get().take(words * sizeof(u64) + (want_node_of ? R * sizeof(u32) : 0) + 64);
		u64 *hp = static_cast<u64 *>(g.block.ptr);
		g.heights = hp; g.n_heights = nb; hp += nb;
		g.node_block = hp; hp += nn;
		g.node_row = hp; g.n_nodes = nn; hp += nn;
		g.edge_src = hp; hp += ne;
		g.edge_dst = hp; g.n_edges = ne; hp += ne;
		g.node_of = want_node_of ? reinterpret_cast<u32 *>(hp) : nullptr; g.n_node_of = want_node_of ? R : 0;
		d2h(g.heights, heights, nb, wide, d.stream);
		d2h(g.node_block, node_block, nn, wide, d.stream);
		d2h(g.node_row, node_row, nn, wide, d.stream);
		d2h(g.edge_src, es, ne, wide, d.stream);
		d2h(g.edge_dst, ed, ne, wide, d.stream);
		if (want_node_of && R) FBG_CUDA(cudaMemcpyAsync(g.node_of, node_of, R * sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
	}
	if (x) {
		GraphDev gd;
		gd.m = m; gd.nb = nb; gd.nn = nn; gd.ne = ne;
		gd.heights = heights; gd.node_block = node_block; gd.node_row = node_row; gd.es = es; gd.ed = ed;
		gd.EOFF = EOFF; gd.node_of = node_of; gd.LEN = LEN; gd.POS = POS; gd.label_total = label_sum;
		k5_emit_xgfa(p, gd, *x);
	}
	if (v) {
		GraphDev gd;
		gd.m = m; gd.nb = nb; gd.nn = nn; gd.ne = ne;
		gd.heights = heights; gd.node_block = node_block; gd.node_row = node_row; gd.es = es; gd.ed = ed;
		gd.EOFF = EOFF; gd.node_of = node_of; gd.LEN = LEN; gd.POS = POS; gd.label_total = label_sum;
		v->n_nodes = nn; v->n_edges = ne; v->label_total = label_sum;
		k6_validate(p, gd, *v);
	}
	d.release(mk);
}

} // namespace fbg
