// dist.cuh -- the multi-GPU side of one build: one process per GPU, every context
// joined to an NCCL communicator over NVLink/NVSwitch.
//
// Sharding of ONE MSA across G GPUs:
//   K1  suffix sorting is sharded by k-mer key range: rank r sorts the suffixes whose
//       leading k-mer falls in its range (whole buckets, so every later refinement
//       stays local).  Ranks are global SA indices.  The rank array R is replicated;
//       after every doubling round the replicas are merged with an all-reduce(max):
//       ranks only ever grow, and a rank array entry is written by its owner alone.
//   K2  f(x) is sharded by contiguous column ranges and merged with the same all-reduce.
//   K3  the chunk guesses and the verify passes of the DP are sharded by chunk / column range
//       and merged the same way; plateau and backtrack run replicated.
//   K0, LCP, K4 run replicated (O(N) streaming passes).
// NCCL is loaded with dlopen so that the single-GPU library has no link-time dependency
// on it; the few entry points used are declared here with NCCL's own ABI.
#pragma once
#include "common.cuh"
#include <dlfcn.h>
#include <cstring>
#include <string>

namespace fbg {

struct NcclUniqueId { char internal[128]; };

struct NcclApi {
	void *lib = nullptr;
	int (*GetUniqueId)(NcclUniqueId *) = nullptr;
	int (*CommInitRank)(void **, int, NcclUniqueId, int) = nullptr;
	int (*CommDestroy)(void *) = nullptr;
	int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
	int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
	const char *(*GetErrorString)(int) = nullptr;

	bool load(std::string &err) {
		if (lib) return true;
		for (const char *nm : {"libnccl.so.2", "libnccl.so"}) {
			lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
			if (lib) break;
		}
		if (!lib) { err = std::string("cannot load NCCL: ") + dlerror(); return false; }
		GetUniqueId = (decltype(GetUniqueId)) dlsym(lib, "ncclGetUniqueId");
		CommInitRank = (decltype(CommInitRank)) dlsym(lib, "ncclCommInitRank");
		CommDestroy = (decltype(CommDestroy)) dlsym(lib, "ncclCommDestroy");
		AllReduce = (decltype(AllReduce)) dlsym(lib, "ncclAllReduce");
		AllGather = (decltype(AllGather)) dlsym(lib, "ncclAllGather");
		GetErrorString = (decltype(GetErrorString)) dlsym(lib, "ncclGetErrorString");
		if (!GetUniqueId || !CommInitRank || !CommDestroy || !AllReduce || !AllGather || !GetErrorString) { err = "NCCL library lacks an expected symbol"; return false; }
		return true;
	}
};

constexpr int NCCL_UINT8 = 1, NCCL_UINT32 = 3, NCCL_UINT64 = 5, NCCL_SUM = 0, NCCL_MAX = 2;
constexpr int MAX_PEERS = 16; // GPUs of one box

// What a kernel of a peer-mode build needs to find the authoritative copy of a rank: rank array of every GPU
// (its own included; the others mapped through CUDA IPC over NVLink) and the column ranges the GPUs own.
struct PeerView {
	u32 *R[MAX_PEERS];
	u32 xb[MAX_PEERS + 1]; // GPU g owns the text positions of columns [xb[g], xb[g+1]); '#' and the sentinel (column >= n) go to the last
	int me, world;
	const u32 *notmine;    // one bit per text position: its column belongs to another GPU (62 MB for 5e8 positions: the look-ups of a round stay in L2)
	__host__ __device__ int owner(u32 c) const {
		int g = 0;
		while (g + 1 < world && c >= xb[g + 1]) g++;
		return g;
	}
};

struct Dist {
	int rank = 0, world = 1;
	void *comm = nullptr;
	NcclApi api;

	void check(int rc, const char *what) const {
		if (rc != 0) throw cuda_error(std::string(what) + " failed: " + (api.GetErrorString ? api.GetErrorString(rc) : "NCCL error"));
	}
	// in-place
	void allreduce_max_u32(u32 *buf, u64 count, cudaStream_t s) const {
		if (world > 1 && count) check(api.AllReduce(buf, buf, count, NCCL_UINT32, NCCL_MAX, comm, s), "ncclAllReduce(max,u32)");
	}
	void allreduce_max_u8(u8 *buf, u64 count, cudaStream_t s) const {
		if (world > 1 && count) check(api.AllReduce(buf, buf, count, NCCL_UINT8, NCCL_MAX, comm, s), "ncclAllReduce(max,u8)");
	}
	void allreduce_sum_u64(u64 *buf, u64 count, cudaStream_t s) const {
		if (world > 1 && count) check(api.AllReduce(buf, buf, count, NCCL_UINT64, NCCL_SUM, comm, s), "ncclAllReduce(sum,u64)");
	}

	// ---- peer memory (one process per GPU: CUDA IPC)
	struct PeerMsg { cudaIpcMemHandle_t handle; u64 offset; u64 pad[7]; }; // 128 bytes
	u8 *xchg = nullptr;                 // device staging of the all-gather
	void *peer_map[MAX_PEERS] = {};     // mapped base of every peer's persistent arena
	cudaIpcMemHandle_t peer_handle[MAX_PEERS];
	// Every rank publishes (arena base, offset of an array inside it); returns the array's address on every GPU.
	void open_peers(void *arena_base, void *array, void **out /* [world] */, cudaStream_t s) {
		if (world > MAX_PEERS) throw cuda_error("more GPUs than one box holds");
		if (!xchg) FBG_CUDA(cudaMalloc(&xchg, sizeof(PeerMsg) * (MAX_PEERS + 1)));
		PeerMsg mine;
		memset(&mine, 0, sizeof mine);
		FBG_CUDA(cudaIpcGetMemHandle(&mine.handle, arena_base));
		mine.offset = (u64) ((char *) array - (char *) arena_base);
		FBG_CUDA(cudaMemcpyAsync(xchg + sizeof(PeerMsg) * MAX_PEERS, &mine, sizeof mine, cudaMemcpyHostToDevice, s));
		check(api.AllGather(xchg + sizeof(PeerMsg) * MAX_PEERS, xchg, sizeof(PeerMsg), NCCL_UINT8, comm, s), "ncclAllGather(ipc handles)");
		PeerMsg all[MAX_PEERS];
		FBG_CUDA(cudaMemcpyAsync(all, xchg, sizeof(PeerMsg) * world, cudaMemcpyDeviceToHost, s));
		FBG_CUDA(cudaStreamSynchronize(s));
		for (int g = 0; g < world; g++) {
			if (g == rank) { out[g] = array; continue; }
			// mapping a peer's arena costs tens of milliseconds: kept across builds, renewed when the peer's arena moved
			if (peer_map[g] && memcmp(&peer_handle[g], &all[g].handle, sizeof(cudaIpcMemHandle_t)) != 0) { cudaIpcCloseMemHandle(peer_map[g]); peer_map[g] = nullptr; }
			if (!peer_map[g]) {
				FBG_CUDA(cudaIpcOpenMemHandle(&peer_map[g], all[g].handle, cudaIpcMemLazyEnablePeerAccess));
				peer_handle[g] = all[g].handle;
			}
			out[g] = (char *) peer_map[g] + all[g].offset;
		}
	}
	// (leaving the communicator, or giving the peer form up for a build)
	void close_peers() {
		for (int g = 0; g < MAX_PEERS; g++)
			if (peer_map[g]) { cudaIpcCloseMemHandle(peer_map[g]); peer_map[g] = nullptr; }
	}
};

// ---- host-side plans (pure functions; exported through the C ABI for the CPU tests)

// Key-range plan: hist[b] = number of suffixes whose k-mer falls in bin b (bins ordered like
// the keys).  bounds[r] .. bounds[r+1] are rank r's bins; cut points are the first bin at
// which the running count reaches r/world of the total, so equal keys never straddle ranks.
static inline void plan_key_ranges(const u32 *hist, u32 nbins, int world, u32 *bounds) {
	u64 total = 0;
	for (u32 b = 0; b < nbins; b++) total += hist[b];
	bounds[0] = 0;
	u64 run = 0;
	u32 b = 0;
	for (int r = 1; r < world; r++) {
		u64 target = total / (u64) world * (u64) r + (total % (u64) world) * (u64) r / (u64) world;
		while (b < nbins && run < target) run += hist[b++];
		bounds[r] = b;
	}
	bounds[world] = nbins;
}

// Column-range plan of K2: rank r owns [x0, x1)
static inline void plan_columns(u64 n, int world, int rank, u64 *x0, u64 *x1) {
	*x0 = n / (u64) world * (u64) rank + (n % (u64) world) * (u64) rank / (u64) world;
	*x1 = rank + 1 == world ? n : n / (u64) world * (u64) (rank + 1) + (n % (u64) world) * (u64) (rank + 1) / (u64) world;
}

} // namespace fbg
