// Host check of the sorting network in csrc/tilesort.cuh: the same index arithmetic
// the CUDA kernel runs, executed one "thread" at a time, against std::sort.
#include "../../founderblockgraphs_b200/csrc/tilesort.cuh"
#include <algorithm>
#include <cstdio>
#include <random>
#include <vector>

struct HostPass {
	uint64_t *key; uint32_t cap;
	void first() { for (uint32_t g = 0; g < cap / 8; g++) fbg::ts_first_phases_group(key, g); }
	template <int NST> void run(uint32_t k, uint32_t jtop) { for (uint32_t g = 0; g < (cap >> NST); g++) fbg::ts_pass_group<NST>(key, k, jtop, g); }
};

int main() {
	std::mt19937_64 rng(12345);
	for (uint32_t cap : {128u, 256u, 2048u, 8192u, 16384u}) {
		for (int rep = 0; rep < 4; rep++) {
			std::vector<uint64_t> a(cap), b;
			for (auto &x : a) x = rep == 1 ? (rng() & 7) : rep == 2 ? ~0ull : rng();
			if (rep == 3) for (uint32_t i = 0; i < cap; i++) a[i] = cap - i;
			b = a;
			std::sort(b.begin(), b.end());
			std::vector<uint64_t> sw(cap);
			for (uint32_t i = 0; i < cap; i++) sw[fbg::ts_sw(i)] = a[i];
			HostPass p{sw.data(), cap};
			fbg::ts_schedule(cap, p);
			for (uint32_t i = 0; i < cap; i++) a[i] = sw[fbg::ts_sw(i)];
			if (a != b) { std::printf("FAIL cap=%u rep=%d\n", cap, rep); return 1; }
		}
	}
	std::printf("ok\n");
	return 0;
}
