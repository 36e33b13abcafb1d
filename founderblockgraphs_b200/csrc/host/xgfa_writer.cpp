// xgfa_writer.cpp -- host-side xGFA formatting, byte-identical to the reference's
// output_efg (fbg.cpp:1185-1301): "M", "X", "B" header lines, then per block the S
// lines of the nodes first seen in that block followed by the L lines of the edges
// arriving in it, then optional P lines.  The device (K4) supplies node order,
// representative rows and the sorted edge set; labels are sliced from the caller's
// own copy of the MSA.
#include "../../../include/fbg_cuda.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

namespace {

// growable formatting buffer
struct Buf {
	std::vector<char> buf;
	size_t used = 0;
	explicit Buf(size_t cap = 1u << 20) : buf(cap) {}
	void need(size_t k) { if (used + k > buf.size()) buf.resize((used + k) * 2); }
	void ch(char c) { need(1); buf[used++] = c; }
	void str(const char *s, size_t k) { need(k); memcpy(buf.data() + used, s, k); used += k; }
	void lit(const char *s) { str(s, strlen(s)); }
	void num(uint64_t v) {
		char tmp[24]; int k = 0;
		do { tmp[k++] = (char) ('0' + v % 10); v /= 10; } while (v);
		need(k);
		while (k) buf[used++] = tmp[--k];
	}
	bool write_to(FILE *fp) { bool ok = used == 0 || fwrite(buf.data(), 1, used, fp) == used; used = 0; return ok; }
};

// S and L lines of the blocks [j0, j1) (fbg.cpp:1228-1260)
void format_blocks(Buf &o, uint64_t j0, uint64_t j1, const uint8_t *rows, uint64_t n, const uint64_t *boundaries, const fbg_graph *g,
                   const uint64_t *nstart, const uint64_t *estart) {
	for (uint64_t j = j0; j < j1; j++) {
		const uint64_t a = j ? boundaries[j - 1] + 1 : 0, b = boundaries[j] < n ? boundaries[j] : n - 1;
		for (uint64_t node = nstart[j]; node < nstart[j + 1]; node++) {
			o.lit("S\t"); o.num(node); o.ch('\t');
			const uint8_t *row = rows + g->node_row[node] * n;
			o.need(b - a + 2);
			for (uint64_t x = a; x <= b; x++) if (row[x] != '-') o.buf[o.used++] = (char) row[x];
			o.ch('\n');
		}
		for (uint64_t edge = estart[j]; edge < estart[j + 1]; edge++) {
			o.lit("L\t"); o.num(g->edge_src[edge]); o.lit("\t+\t"); o.num(g->edge_dst[edge]); o.lit("\t+\t0M\n");
		}
	}
}

} // namespace

extern "C" int fbg_write_xgfa(const char *path, const uint8_t *rows, uint64_t m, uint64_t n,
                              const uint64_t *boundaries, const fbg_graph *g, int output_paths, const char *const *ids) {
	if (!path || !rows || !boundaries || !g) return FBG_ERR_ARG;
	if (output_paths && (!ids || !g->node_of)) return FBG_ERR_ARG;
	FILE *fp = fopen(path, "wb");
	if (!fp) return FBG_ERR_ARG;
	Buf o(1u << 22);
	bool ok = true;
	const uint64_t nb = g->n_blocks;
	o.lit("M\t"); o.num(m); o.ch('\t'); o.num(n); o.ch('\n');                    // fbg.cpp:1201
	o.lit("X\t1");                                                                // 1204-1207
	for (uint64_t i = 0; i + 1 < nb; i++) { o.ch('\t'); o.num(boundaries[i] + 2); }
	o.ch('\n');
	o.lit("B\t");                                                                 // 1210-1220
	for (uint64_t j = 0; j < nb; j++) { if (j) o.ch('\t'); o.num(g->block_height[j]); }
	o.ch('\n');
	ok = o.write_to(fp) && ok;
	// Per block: first node, first edge (nodes are numbered block by block, edges are sorted by
	// the block they arrive in) and a running estimate of the bytes the block will take; the
	// blocks are then formatted by several threads, batch after batch, and written in order.
	std::vector<uint64_t> nstart(nb + 1), estart(nb + 1), weight(nb + 1);
	{
		uint64_t node = 0, edge = 0, wsum = 0, prevstart = 0;
		for (uint64_t j = 0; j < nb; j++) {
			nstart[j] = node; estart[j] = edge; weight[j] = wsum;
			while (node < g->n_nodes && g->node_block[node] == j) node++;
			while (edge < g->n_edges && g->node_block[g->edge_dst[edge]] == j) edge++;
			const uint64_t len = (boundaries[j] < n ? boundaries[j] : n - 1) + 1 - prevstart;
			wsum += (node - nstart[j]) * (len + 12) + (edge - estart[j]) * 24;
			prevstart = boundaries[j] + 1;
		}
		nstart[nb] = node; estart[nb] = edge; weight[nb] = wsum;
	}
	unsigned T = std::thread::hardware_concurrency();
	if (T > 16) T = 16;
	if (T < 1 || weight[nb] < (8u << 20)) T = 1;
	uint64_t batch_bytes = (uint64_t) T * (32u << 20);
	// test hooks: force the number of threads / the batch size whatever the output size
	if (const char *e = getenv("FBG_WRITER_THREADS")) { int v = atoi(e); if (v >= 1 && v <= 64) T = (unsigned) v; }
	if (const char *e = getenv("FBG_WRITER_BATCH")) { long long v = atoll(e); if (v >= 1) batch_bytes = (uint64_t) v; }
	std::vector<Buf> parts(T);
	for (uint64_t j0 = 0; j0 < nb && ok;) {
		// batch [j0, jend): at most batch_bytes of estimated output, at least one block
		uint64_t jend = (uint64_t) (std::upper_bound(weight.begin() + j0, weight.begin() + nb + 1, weight[j0] + batch_bytes) - weight.begin()) - 1;
		if (jend <= j0) jend = j0 + 1;
		if (jend > nb) jend = nb;
		std::vector<uint64_t> cut(T + 1);
		cut[0] = j0; cut[T] = jend;
		for (unsigned t = 1; t < T; t++) {
			const uint64_t target = weight[j0] + (weight[jend] - weight[j0]) * t / T;
			uint64_t c = (uint64_t) (std::lower_bound(weight.begin() + j0, weight.begin() + jend, target) - weight.begin());
			cut[t] = c < cut[t - 1] ? cut[t - 1] : c;
		}
		if (T == 1) format_blocks(parts[0], j0, jend, rows, n, boundaries, g, nstart.data(), estart.data());
		else {
			std::vector<std::thread> th;
			for (unsigned t = 0; t < T; t++)
				th.emplace_back([&, t]() { format_blocks(parts[t], cut[t], cut[t + 1], rows, n, boundaries, g, nstart.data(), estart.data()); });
			for (auto &x : th) x.join();
		}
		for (unsigned t = 0; t < T; t++) ok = parts[t].write_to(fp) && ok;
		j0 = jend;
	}
	if (output_paths) {                                                           // 1263-1300
		for (uint64_t i = 0; i < m; i++) {
			o.lit("P\t"); o.lit(ids[i]); o.ch('\t');
			bool first = true;
			for (uint64_t j = 0; j < nb; j++) {
				uint32_t id = g->node_of[j * m + i];
				if (id == FBG_NO_NODE) continue;
				if (!first) o.ch(',');
				o.num(id); o.ch('+');
				first = false;
			}
			o.lit("\t*\n");
			if (o.used > (4u << 20)) ok = o.write_to(fp) && ok;
		}
	}
	ok = o.write_to(fp) && ok;
	int rc = (!ok || ferror(fp)) ? FBG_ERR_ARG : FBG_OK;
	fclose(fp);
	return rc;
}
