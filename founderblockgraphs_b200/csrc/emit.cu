// emit.cu -- K5: the xGFA bytes of output_efg (fbg.cpp:1185-1301) formatted on the device
// (SURVEY 8(f) f1).  The graph arrays of K4 never leave HBM: the file is a sequence of ITEMS
// whose byte lengths are known per item, one exclusive scan turns lengths into offsets, and
// every item is then written at its offset independently.
//
//   items, in file order                                         bytes
//   (prefix, written by thread 0)  "M\t<m>\t<n>\nX\t1"           from the host, <= 64
//   X items   i in [0, nb-1)       "\t<bnd[i]+2>"                fbg.cpp:1204-1207
//   sep1                           "\nB\t"
//   B items   j in [0, nb)         ["\t"]"<height[j]>"           fbg.cpp:1210-1220
//   sep2                           "\n"
//   per block j: S items of its nodes "S\t<id>\t<label>\n", then L items of the edges arriving
//             in it "L\t<a>\t+\t<b>\t+\t0M\n"                    fbg.cpp:1228-1260
//   per row i (with -p): head "P\t<header_i>", one item per block "<sep><id>+" (sep = tab for
//             the first non-empty block of the row, ',' after), tail "\t*\n"   fbg.cpp:1263-1300
//
// Nodes are numbered block by block and the edge list is grouped by destination block, so the
// item index of node v (block j) is v + EOFF[j] and that of edge e (into block j) is NS[j+1] + e
// (NS = first node of a block, EOFF = first edge of a block): no search, no sort.
// Labels are copied from the gap-free text T (POS/LEN of the representative row's record), one
// warp per node.  The per-item code lives in emit_items.cuh (plain C++, also run on the CPU by
// tests/host/emit_host.cpp).  Same bytes as host/xgfa_writer.cpp, which stays as the host-side writer.
#include "pipeline.cuh"
#include "emit_items.cuh"
#include <cstring>
#include <string>
#include <vector>

namespace fbg {

template <bool WRITE> __global__ void k5_head_kernel(EmitArgs a) {
	const u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t < a.nb) emit_head<WRITE>(a, t);
}
// S items: lengths by one thread per node, bytes by one warp per node
template <bool WRITE> __global__ void k5_nodes_kernel(EmitArgs a) {
	const u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	const u64 v = WRITE ? gid >> 5 : gid;
	if (v < a.nn) emit_node<WRITE>(a, v, lane_id(), 32);
}
template <bool WRITE> __global__ void k5_edges_kernel(EmitArgs a) {
	const u64 e = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (e < a.ne) emit_edge<WRITE>(a, e);
}
template <bool WRITE> __global__ void k5_paths_kernel(EmitArgs a) {
	const u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (gid < a.m * (a.nb + 2)) emit_path<WRITE>(a, gid);
}

u64 k5_scratch_need(u64 m, u64 nb, bool paths, u64 id_bytes) {
	const u64 R = m * nb;
	const u64 ni = emit_item_count(m, nb, R, R, paths);
	return ni * 8 + (ni / SCAN_TILE + 2) * 24 + (nb + 2) * 8 + (nb / SCAN_TILE + 2) * 32 + id_bytes + (m + 2) * 4 + (1u << 16);
}

void k5_emit_xgfa(Pipeline &p, const GraphDev &gd, XgfaJob &x) {
	Dev &d = p.dev;
	const u64 m = gd.m, nb = gd.nb, nn = gd.nn, ne = gd.ne;
	EmitArgs a;
	std::memset(&a, 0, sizeof a);
	a.m = m; a.n = p.n; a.nb = nb; a.nn = nn; a.ne = ne;
	emit_layout(a, x.paths);
	{
		std::string pre = "M\t" + std::to_string(m) + "\t" + std::to_string(p.n) + "\nX\t1"; // fbg.cpp:1201,1204
		a.prelen = (u32) pre.size();
		std::memcpy(a.pre, pre.data(), pre.size());
	}
	const u64 mk = d.mark();
	u32 *NS = d.alloc<u32>(nb + 1), *hmax = d.alloc<u32>(nb + 1);
	u64 *OFF = d.alloc<u64>(a.ni + 1);
	u64 *total = p.scalars + 44;
	u32 *maxh = reinterpret_cast<u32 *>(p.scalars + 45);
	a.bnd = p.bnd; a.heights = gd.heights; a.NS = NS; a.EOFF = gd.EOFF; a.node_block = gd.node_block; a.node_row = gd.node_row;
	a.es = gd.es; a.ed = gd.ed; a.node_of = gd.node_of; a.LEN = gd.LEN; a.POS = gd.POS; a.rowlen = p.rowlen; a.T = p.T; a.OFF = OFF;
	std::vector<u32> idoff;
	std::vector<u8> idbytes;
	if (x.paths) {
		idoff.resize(m + 1);
		for (u64 i = 0; i < m; i++) {
			idoff[i] = (u32) idbytes.size();
			const char *s = x.ids[i];
			idbytes.insert(idbytes.end(), s, s + strlen(s));
			if (idbytes.size() >= 0xFFFFFFF0ull) throw cuda_error("sequence headers exceed 4 GiB");
		}
		idoff[m] = (u32) idbytes.size();
		u32 *d_off = d.alloc<u32>(m + 1);
		u8 *d_ids = d.alloc<u8>(idbytes.size() + 1);
		FBG_CUDA(cudaMemcpyAsync(d_off, idoff.data(), (m + 1) * sizeof(u32), cudaMemcpyHostToDevice, d.stream));
		if (!idbytes.empty()) FBG_CUDA(cudaMemcpyAsync(d_ids, idbytes.data(), idbytes.size(), cudaMemcpyHostToDevice, d.stream));
		a.idoff = d_off; a.idbytes = d_ids;
	}
	// first node of every block (NS[nb] = nn) and the tallest block ("#founders=", fbg.cpp:701)
	device_scan<OpSumU32, true>(d, gd.heights, NS, nb + 1);
	device_scan<OpMaxU32, false>(d, gd.heights, hmax, nb, maxh);
	const u32 TH = 256;
	const u64 np = x.paths ? m * (nb + 2) : 0;
	d.begin(K_EMIT);
	k5_head_kernel<false><<<div_up(nb, TH), TH, 0, d.stream>>>(a);
	if (nn) k5_nodes_kernel<false><<<div_up(nn, TH), TH, 0, d.stream>>>(a);
	if (ne) k5_edges_kernel<false><<<div_up(ne, TH), TH, 0, d.stream>>>(a);
	if (np) k5_paths_kernel<false><<<div_up(np, TH), TH, 0, d.stream>>>(a);
	FBG_LAUNCH_CHECK();
	d.end((double) a.ni * 8.0 + (double) (nn + ne) * 12.0 + (double) np * 4.0);
	d.note_launches(K_EMIT, 3);
	device_scan<OpSumU64, true>(d, OFF, OFF, a.ni, total);
	FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 44, p.scalars + 44, 2 * sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream)); // also keeps idoff / idbytes alive until their copies are done
	const u64 bytes = p.h_scalars[44] + a.prelen;
	x.max_height = *reinterpret_cast<u32 *>(p.h_scalars + 45);
	if (bytes > p.xgfa_cap) {
		if (p.xgfa) FBG_CUDA(cudaFree(p.xgfa));
		p.xgfa = nullptr; p.xgfa_cap = 0;
		const u64 cap = bytes + (bytes >> 3) + 4096;
		FBG_CUDA(cudaMalloc(&p.xgfa, cap));
		p.xgfa_cap = cap;
	}
	a.out = p.xgfa;
	d.begin(K_EMIT);
	k5_head_kernel<true><<<div_up(nb, TH), TH, 0, d.stream>>>(a);
	if (nn) k5_nodes_kernel<true><<<div_up(nn * 32, TH), TH, 0, d.stream>>>(a);
	if (ne) k5_edges_kernel<true><<<div_up(ne, TH), TH, 0, d.stream>>>(a);
	if (np) k5_paths_kernel<true><<<div_up(np, TH), TH, 0, d.stream>>>(a);
	FBG_LAUNCH_CHECK();
	d.end((double) bytes + (double) gd.label_total + (double) a.ni * 8.0 + (double) (nn + ne) * 16.0 + (double) np * 4.0);
	d.note_launches(K_EMIT, 3);
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	d.release(mk);
	p.xgfa_len = bytes;
	p.have_xgfa = true;
	x.n_nodes = nn; x.n_edges = ne; x.label_total = gd.label_total; x.n_bytes = bytes;
}

} // namespace fbg
