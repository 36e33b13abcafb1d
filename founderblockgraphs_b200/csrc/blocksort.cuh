// blocksort.cuh -- shared-memory bitonic sort used by the per-column (K2) and
// per-block (K4) kernels.  A CTA is split into `groups` thread groups of `tpg`
// threads (a multiple of 32); every group sorts its own `mpad` (power of two)
// 64-bit keys.  All groups execute the same step sequence, so the
// __syncthreads() calls are uniform across the CTA.
#pragma once
#include "common.cuh"
#include "tilesort.cuh"

namespace fbg {

__device__ __forceinline__ void group_bitonic_sort(u64 *key, u32 mpad, u32 tg, u32 tpg) {
	for (u32 k = 2; k <= mpad; k <<= 1) {
		for (u32 j = k >> 1; j > 0; j >>= 1) {
			for (u32 e = tg; e < (mpad >> 1); e += tpg) {
				u32 lo = ((e & ~(j - 1)) << 1) | (e & (j - 1));
				u32 hi = lo | j;
				u64 ka = key[lo], kb = key[hi];
				bool up = (lo & k) == 0;
				if ((ka > kb) == up) { key[lo] = kb; key[hi] = ka; }
			}
			__syncthreads();
		}
	}
}

// Same contract on the three-stages-per-pass network of tilesort.cuh: key i of a group
// lives at key[ts_sw(i)] before and after (mpad is a multiple of 32, so the swizzle stays
// inside the group's array).
struct GroupPass {
	u64 *key; u32 mpad, tg, tpg;
	__device__ __forceinline__ void first() {
		for (u32 g = tg; g < mpad / 8; g += tpg) ts_first_phases_group(key, g);
		__syncthreads();
	}
	template <int NST> __device__ __forceinline__ void run(u32 k, u32 jtop) {
		for (u32 g = tg; g < (mpad >> NST); g += tpg) ts_pass_group<NST>(key, k, jtop, g);
		__syncthreads();
	}
};
__device__ __forceinline__ void group_tile_sort(u64 *key, u32 mpad, u32 tg, u32 tpg) {
	GroupPass p{key, mpad, tg, tpg};
	ts_schedule(mpad, p);
}

struct GroupCfg { u32 mpad, tpg, groups, threads; };

// one thread per element pair, <= 1024 threads per group, >= one warp per group,
// >= 256 threads per CTA
static inline GroupCfg group_cfg(u64 m) {
	GroupCfg c;
	c.mpad = 32;
	while (c.mpad < m) c.mpad <<= 1;
	c.tpg = c.mpad / 2;
	if (c.tpg > 1024) c.tpg = 1024;
	if (c.tpg < 32) c.tpg = 32;
	c.threads = c.tpg < 256 ? 256 : c.tpg;
	c.groups = c.threads / c.tpg;
	return c;
}

} // namespace fbg
