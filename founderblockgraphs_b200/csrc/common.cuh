// common.cuh -- shared types, error handling and warp/block helpers for the
// sm_100a kernels of the founder-block-graph hot path.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;
typedef uint64_t u64;
typedef int64_t i64;

namespace fbg {

struct cuda_error : std::runtime_error {
	explicit cuda_error(const std::string &s) : std::runtime_error(s) {}
};

// a size or shape this build does not handle (reported as FBG_ERR_UNSUPPORTED, not as a CUDA failure)
struct unsupported_error : std::runtime_error {
	explicit unsupported_error(const std::string &s) : std::runtime_error(s) {}
};

#define FBG_CUDA(expr)                                                                       \
	do {                                                                                     \
		cudaError_t e__ = (expr);                                                            \
		if (e__ != cudaSuccess)                                                              \
			throw fbg::cuda_error(std::string(#expr) + " failed at " + __FILE__ + ":" +     \
			                      std::to_string(__LINE__) + ": " + cudaGetErrorString(e__)); \
	} while (0)

#define FBG_LAUNCH_CHECK() FBG_CUDA(cudaGetLastError())

static inline u32 div_up(u64 a, u64 b) { return (u32) ((a + b - 1) / b); }
static inline u64 round_up(u64 a, u64 b) { return (a + b - 1) / b * b; }

constexpr u32 U32_MAX = 0xFFFFFFFFu;
constexpr int NUM_SMS = 148; // B200

// ---- per-kernel accounting (roofline.achieved = algorithmic bytes / event time)
enum KernelId {
	K_PACK = 0,      // K0: masks, ranks, text scatter
	K_KMER,          // K1: initial k-mer keys
	K_RADIX_HIST,    // K1/K4: radix digit histograms
	K_RADIX_SCATTER, // K1/K4: radix scatter passes
	K_SCAN,          // all prefix scans
	K_DOUBLING_AUX,  // K1: key build / head flags / rank update / compaction
	K_LCP,           // K1: irreducible LCP + fill + permute
	K_FCOL,          // K2: per-column f(x)
	K_DP,            // K3: segmentation DP + backtrack
	K_GRAPH,         // K4: label hash / dedup / edges
	K_TILE_SORT,     // K1: bucket-local shared-memory sort of tied suffixes
	K_COMM,          // NCCL collectives of a multi-GPU build (rank replicas, f)
	K_EMIT,          // K5: xGFA formatting (item lengths, offsets, byte writes)
	K_EDGES,         // K4: node ids, per-block edge sort / unique, gather (K_GRAPH keeps hash, dedup, verification)
	K_COUNT
};

struct KernelStat {
	double ms = 0.0;     // accumulated CUDA-event time
	double bytes = 0.0;  // accumulated algorithmic bytes
	u64 launches = 0;
};

// ---- device helpers
__device__ __forceinline__ u32 lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ u32 lanemask_lt() {
	u32 m;
	asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
	return m;
}

// 16-byte streaming load through the read-only path, no L1 allocation.
__device__ __forceinline__ uint4 ld_nc_u4(const void *p) {
	uint4 r;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
	             : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
	return r;
}

template <typename T> __device__ __forceinline__ T warp_incl_sum(T v) {
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		T t = __shfl_up_sync(0xffffffffu, v, o);
		if ((int) lane_id() >= o) v += t;
	}
	return v;
}
template <typename T> __device__ __forceinline__ T warp_incl_max(T v) {
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		T t = __shfl_up_sync(0xffffffffu, v, o);
		if ((int) lane_id() >= o) v = t > v ? t : v;
	}
	return v;
}
template <typename T> __device__ __forceinline__ T warp_incl_min(T v) {
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		T t = __shfl_up_sync(0xffffffffu, v, o);
		if ((int) lane_id() >= o) v = t < v ? t : v;
	}
	return v;
}
template <typename T> __device__ __forceinline__ T warp_reduce_max(T v) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		T t = __shfl_xor_sync(0xffffffffu, v, o);
		v = t > v ? t : v;
	}
	return v;
}
template <typename T> __device__ __forceinline__ T warp_reduce_sum(T v) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

} // namespace fbg
