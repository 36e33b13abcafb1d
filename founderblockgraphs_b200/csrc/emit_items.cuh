// emit_items.cuh -- K5: the per-item formatting code of the device-side xGFA writer
// (emit.cu), kept free of CUDA so that tests/host/emit_host.cpp can run the very same index
// arithmetic and byte writes on the CPU, one "thread" at a time, against the reference bytes.
// Item layout and file grammar: header of emit.cu.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define FBG_EHD __host__ __device__ __forceinline__
#else
#define FBG_EHD inline
#endif

namespace fbg {

struct EmitArgs {
	uint64_t m, n, nb, nn, ne;
	uint64_t iB, iG, iP, ni; // first B item, first graph item, first P item, number of items
	const uint32_t *bnd, *heights, *NS, *EOFF, *node_block, *node_row, *es, *ed, *node_of, *LEN, *POS, *rowlen, *idoff;
	const uint8_t *T, *idbytes;
	uint64_t *OFF;  // lengths, then (after the exclusive scan) offsets relative to the end of the prefix
	uint8_t *out;
	uint32_t prelen;
	uint8_t pre[64];
};

FBG_EHD uint64_t emit_item_count(uint64_t m, uint64_t nb, uint64_t nn, uint64_t ne, bool paths) {
	return (nb - 1) + 1 + nb + 1 + nn + ne + (paths ? m * (nb + 2) : 0);
}
FBG_EHD void emit_layout(EmitArgs &a, bool paths) {
	a.iB = a.nb; a.iG = 2 * a.nb + 1; a.iP = a.iG + a.nn + a.ne;
	a.ni = emit_item_count(a.m, a.nb, a.nn, a.ne, paths);
}

FBG_EHD uint32_t emit_ndigits(uint64_t v) {
	uint32_t d = 1;
	while (v >= 10) { v /= 10; d++; }
	return d;
}
FBG_EHD void emit_put_num(uint8_t *o, uint64_t v, uint32_t nd) {
	for (uint32_t k = nd; k-- > 0;) { o[k] = (uint8_t) ('0' + v % 10); v /= 10; }
}

// X item t and B item t; t == 0 also writes the two separators and the prefix
template <bool WRITE> FBG_EHD void emit_head(const EmitArgs &a, uint64_t t) {
	uint8_t *base = a.out + a.prelen;
	if (t + 1 < a.nb) {
		const uint64_t v = (uint64_t) a.bnd[t] + 2;
		const uint32_t nd = emit_ndigits(v);
		if (!WRITE) a.OFF[t] = 1 + nd;
		else { uint8_t *o = base + a.OFF[t]; o[0] = '\t'; emit_put_num(o + 1, v, nd); }
	}
	{
		const uint64_t v = a.heights[t];
		const uint32_t nd = emit_ndigits(v), tab = t ? 1u : 0u;
		if (!WRITE) a.OFF[a.iB + t] = tab + nd;
		else { uint8_t *o = base + a.OFF[a.iB + t]; if (tab) o[0] = '\t'; emit_put_num(o + tab, v, nd); }
	}
	if (t == 0) {
		if (!WRITE) { a.OFF[a.iB - 1] = 3; a.OFF[a.iG - 1] = 1; }
		else {
			for (uint32_t q = 0; q < a.prelen; q++) a.out[q] = a.pre[q];
			uint8_t *o = base + a.OFF[a.iB - 1];
			o[0] = '\n'; o[1] = 'B'; o[2] = '\t';
			base[a.OFF[a.iG - 1]] = '\n';
		}
	}
}

// S item of node v; the WRITE form is run by `lanes` cooperating threads (lane = 0..lanes-1)
template <bool WRITE> FBG_EHD void emit_node(const EmitArgs &a, uint64_t v, uint32_t lane, uint32_t lanes) {
	const uint64_t j = a.node_block[v], rec = j * a.m + a.node_row[v];
	const uint32_t len = a.LEN[rec], nd = emit_ndigits(v);
	const uint64_t idx = a.iG + v + a.EOFF[j];
	if (!WRITE) { a.OFF[idx] = (uint64_t) 4 + nd + len; return; }
	uint8_t *o = a.out + a.prelen + a.OFF[idx];
	if (lane == 0) { o[0] = 'S'; o[1] = '\t'; emit_put_num(o + 2, v, nd); o[2 + nd] = '\t'; o[3 + nd + len] = '\n'; }
	const uint8_t *src = a.T + a.POS[rec];
	uint8_t *dst = o + 3 + nd;
	for (uint32_t q = lane; q < len; q += lanes) dst[q] = src[q];
}

// L item of edge e
template <bool WRITE> FBG_EHD void emit_edge(const EmitArgs &a, uint64_t e) {
	const uint64_t s = a.es[e], t = a.ed[e];
	const uint64_t j = a.node_block[t];
	const uint32_t ds = emit_ndigits(s), dt = emit_ndigits(t);
	const uint64_t idx = a.iG + a.NS[j + 1] + e;
	if (!WRITE) { a.OFF[idx] = (uint64_t) 11 + ds + dt; return; }
	uint8_t *o = a.out + a.prelen + a.OFF[idx];
	o[0] = 'L'; o[1] = '\t'; emit_put_num(o + 2, s, ds); o += 2 + ds;
	o[0] = '\t'; o[1] = '+'; o[2] = '\t'; emit_put_num(o + 3, t, dt); o += 3 + dt;
	o[0] = '\t'; o[1] = '+'; o[2] = '\t'; o[3] = '0'; o[4] = 'M'; o[5] = '\n';
}

// P items: gid = i * (nb + 2) + k; k = 0 head, 1..nb blocks, nb + 1 tail
template <bool WRITE> FBG_EHD void emit_path(const EmitArgs &a, uint64_t gid) {
	const uint64_t per = a.nb + 2;
	const uint64_t i = gid / per, k = gid % per;
	const uint64_t idx = a.iP + gid;
	uint8_t *o = WRITE ? a.out + a.prelen + a.OFF[idx] : nullptr;
	if (k == 0) {
		const uint32_t b = a.idoff[i], len = a.idoff[i + 1] - b;
		if (!WRITE) { a.OFF[idx] = (uint64_t) 2 + len; return; }
		o[0] = 'P'; o[1] = '\t';
		for (uint32_t q = 0; q < len; q++) o[2 + q] = a.idbytes[b + q];
	} else if (k <= a.nb) {
		const uint32_t id = a.node_of[(k - 1) * a.m + i];
		if (id == 0xFFFFFFFFu) { if (!WRITE) a.OFF[idx] = 0; return; }
		const uint32_t nd = emit_ndigits(id);
		if (!WRITE) { a.OFF[idx] = (uint64_t) 2 + nd; return; }
		// the first non-empty block of the row starts where the head ends
		const bool first = a.OFF[idx] == a.OFF[a.iP + i * per + 1];
		o[0] = first ? '\t' : ',';
		emit_put_num(o + 1, id, nd);
		o[1 + nd] = '+';
	} else {
		// a row without a single non-gap character has no items: "P\t<id>\t" + "" + "\t*\n"
		const uint32_t extra = a.rowlen[i] == 0 ? 1u : 0u;
		if (!WRITE) { a.OFF[idx] = (uint64_t) 3 + extra; return; }
		if (extra) *o++ = '\t';
		o[0] = '\t'; o[1] = '*'; o[2] = '\n';
	}
}

} // namespace fbg
