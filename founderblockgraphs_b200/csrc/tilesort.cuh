// tilesort.cuh -- bitonic sorting network over one shared-memory tile of 64-bit
// keys, used by the bucket-local passes of the suffix sorter (suffix.cu).
//
// The network is the textbook one (phase k = 2,4,..,CAP; stage j = k/2,..,1;
// element i is compared with i^j, ascending iff (i & k) == 0), but a thread
// executes up to three consecutive stages of a phase on 2/4/8 keys held in
// registers, so a tile makes ~CAP-independent 1/3 of the shared-memory round
// trips and barriers of the stage-per-pass form.  The first three phases
// (k = 2,4,8) run in a single pass on 8 consecutive keys.
//
// The index arithmetic is plain C++ so that tests/host/test_tilesort.cpp can run
// the same code on the CPU (one "thread" at a time) against std::sort.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define FBG_HD __host__ __device__ __forceinline__
#else
#define FBG_HD inline
#endif

namespace fbg {

// Shared-memory placement of key i: the low four index bits are XORed with bits 3..6.
// Whatever the stride of a pass (1, 2, 4, ... keys between the lanes of a half-warp, eight
// keys per lane), the sixteen 8-byte accesses of a half-warp then fall into sixteen
// different bank pairs.  A GF(2)-linear bijection: ts_sw(a ^ b) == ts_sw(a) ^ ts_sw(b).
FBG_HD uint32_t ts_sw(uint32_t i) { return i ^ ((i >> 3) & 15u); }

FBG_HD void ts_cx(uint64_t &a, uint64_t &b, bool up) {
	// ascending (up): a <= b afterwards
	const bool sw = (a > b) == up;
	const uint64_t t = a;
	a = sw ? b : a;
	b = sw ? t : b;
}

// Stages jtop, jtop/2, .., jtop >> (NST-1) of phase k for group g.  A group is
// the 2^NST keys whose indices agree outside bits log2(low)..log2(jtop), low =
// jtop >> (NST-1).  All of them share the direction bit (k > jtop).
template <int NST> FBG_HD void ts_pass_group(uint64_t *key, uint32_t k, uint32_t jtop, uint32_t g) {
	constexpr int E = 1 << NST;
	const uint32_t low = jtop >> (NST - 1);
	const uint32_t base = ((g & ~(low - 1)) << NST) | (g & (low - 1));
	const bool up = (base & k) == 0;
	uint64_t v[E];
#pragma unroll
	for (int c = 0; c < E; c++) v[c] = key[ts_sw(base + (uint32_t) c * low)];
#pragma unroll
	for (int st = NST - 1; st >= 0; st--) {
#pragma unroll
		for (int c = 0; c < E; c++)
			if ((c & (1 << st)) == 0) ts_cx(v[c], v[c | (1 << st)], up);
	}
#pragma unroll
	for (int c = 0; c < E; c++) key[ts_sw(base + (uint32_t) c * low)] = v[c];
}

// Phases k = 2, 4, 8 on the 8 consecutive keys [8g, 8g+8).
FBG_HD void ts_first_phases_group(uint64_t *key, uint32_t g) {
	const uint32_t base = g << 3;
	uint64_t v[8];
#pragma unroll
	for (int c = 0; c < 8; c++) v[c] = key[ts_sw(base + c)];
#pragma unroll
	for (int k = 2; k <= 8; k <<= 1) {
#pragma unroll
		for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
			for (int c = 0; c < 8; c++)
				if ((c & j) == 0) ts_cx(v[c], v[c | j], (((base + c) & k) == 0));
		}
	}
#pragma unroll
	for (int c = 0; c < 8; c++) key[ts_sw(base + c)] = v[c];
}

// Schedule of one phase k >= 16: stages log2(k)-1 .. 0 in passes of <= 3.  The
// callback runs one pass over every group; the caller puts a barrier after it.
// Host and device use the same driver through the PASS functor.
template <class PASS> FBG_HD void ts_schedule(uint32_t cap, PASS &pass) {
	pass.first();
	for (uint32_t k = 16; k <= cap; k <<= 1) {
		uint32_t j = k >> 1;
		// peel 3 stages at a time from the top, leaving the remainder (1 or 2) for the last pass
		// number of stages of this phase = log2(k)
#if defined(__CUDA_ARCH__)
		int stages = 31 - __clz((int) k);
#else
		int stages = 31 - __builtin_clz(k);
#endif
		while (stages > 0) {
			int nst = stages >= 3 ? 3 : stages;
			if (nst == 3) pass.template run<3>(k, j);
			else if (nst == 2) pass.template run<2>(k, j);
			else pass.template run<1>(k, j);
			j >>= nst;
			stages -= nst;
		}
	}
}

#if defined(__CUDACC__)
template <int CAP, int THREADS> struct TsDevicePass {
	uint64_t *key;
	__device__ __forceinline__ void first() {
		for (uint32_t g = threadIdx.x; g < CAP / 8; g += THREADS) ts_first_phases_group(key, g);
		__syncthreads();
	}
	template <int NST> __device__ __forceinline__ void run(uint32_t k, uint32_t jtop) {
		for (uint32_t g = threadIdx.x; g < (CAP >> NST); g += THREADS) ts_pass_group<NST>(key, k, jtop, g);
		__syncthreads();
	}
};

// Same network on the first `net` keys only (net = power of two, 32 <= net <= tile capacity,
// uniform over the CTA): a tile that turned out half empty pays for the smaller network.
template <int THREADS> struct TsDevicePassRt {
	uint64_t *key; uint32_t net;
	__device__ __forceinline__ void first() {
		for (uint32_t g = threadIdx.x; g < net / 8; g += THREADS) ts_first_phases_group(key, g);
		__syncthreads();
	}
	template <int NST> __device__ __forceinline__ void run(uint32_t k, uint32_t jtop) {
		for (uint32_t g = threadIdx.x; g < (net >> NST); g += THREADS) ts_pass_group<NST>(key, k, jtop, g);
		__syncthreads();
	}
};
template <int THREADS> __device__ __forceinline__ void tile_bitonic_sort_rt(uint64_t *key, uint32_t net) {
	TsDevicePassRt<THREADS> p{key, net};
	ts_schedule(net, p);
}

// Sorts the tile ascending: key i lives at key[ts_sw(i)] before and after.  CAP is a power
// of two >= 128; all THREADS threads of the CTA call it; the keys must be visible (barrier)
// before the call and are visible after.
template <int CAP, int THREADS> __device__ __forceinline__ void tile_bitonic_sort(uint64_t *key) {
	TsDevicePass<CAP, THREADS> p{key};
	ts_schedule(CAP, p);
}
#endif

} // namespace fbg
