// xgfa_writer.cpp -- host-side xGFA formatting, byte-identical to the reference's
// output_efg (fbg.cpp:1185-1301): "M", "X", "B" header lines, then per block the S
// lines of the nodes first seen in that block followed by the L lines of the edges
// arriving in it, then optional P lines.  The device (K4) supplies node order,
// representative rows and the sorted edge set; labels are sliced from the caller's
// own copy of the MSA.
#include "../../../include/fbg_cuda.h"
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

struct Out {
	FILE *fp;
	std::vector<char> buf;
	size_t used = 0;
	explicit Out(FILE *f) : fp(f), buf(1u << 22) {}
	void flush() { if (used) fwrite(buf.data(), 1, used, fp); used = 0; }
	void need(size_t k) { if (used + k > buf.size()) { flush(); if (k > buf.size()) buf.resize(k * 2); } }
	void ch(char c) { need(1); buf[used++] = c; }
	void str(const char *s, size_t k) { need(k); memcpy(buf.data() + used, s, k); used += k; }
	void lit(const char *s) { str(s, strlen(s)); }
	void num(uint64_t v) {
		char tmp[24]; int k = 0;
		do { tmp[k++] = (char) ('0' + v % 10); v /= 10; } while (v);
		need(k);
		while (k) buf[used++] = tmp[--k];
	}
};

} // namespace

extern "C" int fbg_write_xgfa(const char *path, const uint8_t *rows, uint64_t m, uint64_t n,
                              const uint64_t *boundaries, const fbg_graph *g, int output_paths, const char *const *ids) {
	if (!path || !rows || !boundaries || !g) return FBG_ERR_ARG;
	if (output_paths && (!ids || !g->node_of)) return FBG_ERR_ARG;
	FILE *fp = fopen(path, "wb");
	if (!fp) return FBG_ERR_ARG;
	Out o(fp);
	const uint64_t nb = g->n_blocks;
	o.lit("M\t"); o.num(m); o.ch('\t'); o.num(n); o.ch('\n');                    // fbg.cpp:1201
	o.lit("X\t1");                                                                // 1204-1207
	for (uint64_t i = 0; i + 1 < nb; i++) { o.ch('\t'); o.num(boundaries[i] + 2); }
	o.ch('\n');
	o.lit("B\t");                                                                 // 1210-1220
	for (uint64_t j = 0; j < nb; j++) { if (j) o.ch('\t'); o.num(g->block_height[j]); }
	o.ch('\n');
	uint64_t node = 0, edge = 0, prevstart = 0;
	for (uint64_t j = 0; j < nb; j++) {                                           // 1228-1260
		const uint64_t a = prevstart, b = boundaries[j] < n ? boundaries[j] : n - 1;
		for (; node < g->n_nodes && g->node_block[node] == j; node++) {
			o.lit("S\t"); o.num(node); o.ch('\t');
			const uint8_t *row = rows + g->node_row[node] * n;
			o.need(b - a + 2);
			for (uint64_t x = a; x <= b; x++) if (row[x] != '-') o.buf[o.used++] = (char) row[x];
			o.ch('\n');
		}
		for (; edge < g->n_edges && g->node_block[g->edge_dst[edge]] == j; edge++) {
			o.lit("L\t"); o.num(g->edge_src[edge]); o.lit("\t+\t"); o.num(g->edge_dst[edge]); o.lit("\t+\t0M\n");
		}
		prevstart = boundaries[j] + 1;
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
		}
	}
	o.flush();
	int rc = ferror(fp) ? FBG_ERR_ARG : FBG_OK;
	fclose(fp);
	return rc;
}
