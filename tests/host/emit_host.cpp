// Host run of the device-side xGFA writer's per-item code (csrc/emit_items.cuh): the same
// index arithmetic and byte writes the K5 kernels execute, one "thread" at a time, with a
// sequential exclusive scan in the place of the device scan.  tests/test_emit_host.py feeds
// it the oracle's graph arrays and compares the bytes with the oracle's writer.
#include "../../founderblockgraphs_b200/csrc/emit_items.cuh"
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

extern "C" uint64_t emit_host_format(uint64_t m, uint64_t n, uint64_t nb, uint64_t nn, uint64_t ne, const uint32_t *bnd, const uint32_t *heights /* nb+1, last 0 */,
                                     const uint32_t *node_block, const uint32_t *node_row, const uint32_t *es, const uint32_t *ed, const uint32_t *node_of,
                                     const uint32_t *LEN, const uint32_t *POS, const uint32_t *rowlen, const uint8_t *T, int paths, const uint32_t *idoff,
                                     const uint8_t *idbytes, uint32_t lanes, int reverse_order, uint8_t *out, uint64_t out_cap) {
	using namespace fbg;
	EmitArgs a;
	std::memset(&a, 0, sizeof a);
	a.m = m; a.n = n; a.nb = nb; a.nn = nn; a.ne = ne;
	emit_layout(a, paths != 0);
	std::string pre = "M\t" + std::to_string(m) + "\t" + std::to_string(n) + "\nX\t1";
	a.prelen = (uint32_t) pre.size();
	std::memcpy(a.pre, pre.data(), pre.size());
	// what K4 leaves behind: first node / first arriving edge of every block
	std::vector<uint32_t> NS(nb + 1, 0), EOFF(nb + 1, 0), EC(nb + 1, 0);
	for (uint64_t j = 0; j < nb; j++) NS[j + 1] = NS[j] + heights[j];
	for (uint64_t e = 0; e < ne; e++) EC[node_block[ed[e]]]++;
	for (uint64_t j = 0; j < nb; j++) EOFF[j + 1] = EOFF[j] + EC[j];
	std::vector<uint64_t> OFF(a.ni + 1, 0xDEADBEEFDEADBEEFull);
	a.bnd = bnd; a.heights = heights; a.NS = NS.data(); a.EOFF = EOFF.data(); a.node_block = node_block; a.node_row = node_row;
	a.es = es; a.ed = ed; a.node_of = node_of; a.LEN = LEN; a.POS = POS; a.rowlen = rowlen; a.idoff = idoff; a.T = T; a.idbytes = idbytes;
	a.OFF = OFF.data();
	const uint64_t np = paths ? m * (nb + 2) : 0;
	for (uint64_t t = 0; t < nb; t++) emit_head<false>(a, t);
	for (uint64_t v = 0; v < nn; v++) emit_node<false>(a, v, 0, 1);
	for (uint64_t e = 0; e < ne; e++) emit_edge<false>(a, e);
	for (uint64_t g = 0; g < np; g++) emit_path<false>(a, g);
	uint64_t run = 0;
	for (uint64_t i = 0; i < a.ni; i++) {
		if (OFF[i] == 0xDEADBEEFDEADBEEFull) return ~0ull; // an item nobody gave a length
		uint64_t l = OFF[i]; OFF[i] = run; run += l;
	}
	const uint64_t bytes = run + a.prelen;
	if (bytes > out_cap) return bytes;
	std::memset(out, 0xEE, out_cap);
	a.out = out;
	// "threads" in ascending or descending order: the result must not depend on it
	auto idx = [&](uint64_t i, uint64_t cnt) { return reverse_order ? cnt - 1 - i : i; };
	for (uint64_t t = 0; t < nb; t++) emit_head<true>(a, idx(t, nb));
	for (uint64_t v = 0; v < nn; v++) for (uint32_t l = 0; l < lanes; l++) emit_node<true>(a, idx(v, nn), reverse_order ? lanes - 1 - l : l, lanes);
	for (uint64_t e = 0; e < ne; e++) emit_edge<true>(a, idx(e, ne));
	for (uint64_t g = 0; g < np; g++) emit_path<true>(a, idx(g, np));
	return bytes;
}
