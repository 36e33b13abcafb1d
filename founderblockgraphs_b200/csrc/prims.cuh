// prims.cuh -- device-wide primitives written for this path: workspace arena,
// per-kernel event accounting, three-phase tiled scans and the LSD radix sort
// (warp match/ballot ranking + shared-memory staged bucket histograms).
#pragma once
#include <cstdlib>
#include <cstdio>
#include "common.cuh"
#include <vector>

namespace fbg {

// ---------------------------------------------------------------- device ctx
struct Dev {
	int device = 0;
	cudaStream_t stream = nullptr;
	// bump arena for scratch (one cudaMalloc, re-used by every phase)
	u8 *arena = nullptr;
	u64 arena_cap = 0, arena_top = 0, arena_peak = 0;
	// event accounting
	struct Rec { int id; cudaEvent_t a, b; double bytes; };
	std::vector<Rec> recs;
	std::vector<cudaEvent_t> pool;
	KernelStat stats[K_COUNT];
	bool timing = true;
	int open_id = -1;
	cudaEvent_t open_ev = nullptr;

	void reserve(u64 bytes) {
		if (bytes <= arena_cap) return;
		if (arena) FBG_CUDA(cudaFree(arena));
		arena = nullptr; arena_cap = 0;
		FBG_CUDA(cudaMalloc(&arena, bytes));
		arena_cap = bytes;
	}
	// phase entry: make sure the (currently empty) scratch arena can hold `bytes`
	void ensure(u64 bytes) {
		if (bytes <= arena_cap) return;
		if (arena_top) throw cuda_error("scratch arena must be empty when it grows");
		FBG_CUDA(cudaStreamSynchronize(stream));
		reserve(bytes);
	}
	template <typename T> T *alloc(u64 count) {
		u64 off = round_up(arena_top, 256);
		u64 need = off + count * sizeof(T);
		if (need > arena_cap)
			throw cuda_error("device arena exhausted: need " + std::to_string(need) + " of " + std::to_string(arena_cap));
		arena_top = need;
		if (arena_top > arena_peak) arena_peak = arena_top;
		return reinterpret_cast<T *>(arena + off);
	}
	// second bump arena for the arrays that live as long as one MSA (text, SA, ISA, LCP, f, ...):
	// grown, never shrunk, so a context that builds graph after graph makes no cudaMalloc calls
	u8 *pers = nullptr;
	u64 pers_cap = 0, pers_top = 0;
	void reserve_pers(u64 bytes) {
		if (bytes <= pers_cap) return;
		if (pers) FBG_CUDA(cudaFree(pers));
		pers = nullptr; pers_cap = 0;
		FBG_CUDA(cudaMalloc(&pers, bytes));
		pers_cap = bytes;
	}
	template <typename T> T *palloc(u64 count) {
		u64 off = round_up(pers_top, 256);
		u64 need = off + (count ? count : 1) * sizeof(T);
		if (need > pers_cap)
			throw cuda_error("persistent device arena exhausted: need " + std::to_string(need) + " of " + std::to_string(pers_cap));
		pers_top = need;
		return reinterpret_cast<T *>(pers + off);
	}
	u64 mark() const { return arena_top; }
	void release(u64 m) { arena_top = m; }

	cudaEvent_t get_event() {
		if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
		cudaEvent_t e; FBG_CUDA(cudaEventCreate(&e)); return e;
	}
	void begin(int id) {
		if (!timing) return;
		open_id = id; open_ev = get_event();
		FBG_CUDA(cudaEventRecord(open_ev, stream));
	}
	void end(double bytes) {
		if (!timing) return;
		cudaEvent_t b = get_event();
		FBG_CUDA(cudaEventRecord(b, stream));
		recs.push_back({open_id, open_ev, b, bytes});
		stats[open_id].launches++;
		open_id = -1;
	}
	void note_launches(int id, u64 extra) { if (timing) stats[id].launches += extra; }
	// call after a stream synchronize
	void resolve() {
		static const bool trace = getenv("FBG_TRACE") != nullptr; // every timed launch group, in launch order, on stderr
		for (auto &r : recs) {
			float ms = 0.f;
			FBG_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
			if (trace) fprintf(stderr, "[fbg-trace] dev %d group %2d %9.3f ms %12.0f B\n", device, r.id, ms, r.bytes);
			stats[r.id].ms += ms; stats[r.id].bytes += r.bytes;
			pool.push_back(r.a); pool.push_back(r.b);
		}
		recs.clear();
	}
	void reset_stats() { for (auto &s : stats) s = KernelStat(); }
	void destroy() {
		for (auto &r : recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
		for (auto e : pool) cudaEventDestroy(e);
		recs.clear(); pool.clear();
		if (arena) cudaFree(arena);
		arena = nullptr; arena_cap = 0;
		if (pers) cudaFree(pers);
		pers = nullptr; pers_cap = 0; pers_top = 0;
	}
};

// ------------------------------------------------------------------- scans
struct OpSumU32 { typedef u32 T; __host__ __device__ static T id() { return 0; } __host__ __device__ static T f(T a, T b) { return a + b; } };
struct OpSumU64 { typedef u64 T; __host__ __device__ static T id() { return 0; } __host__ __device__ static T f(T a, T b) { return a + b; } };
struct OpMaxU64 { typedef u64 T; __host__ __device__ static T id() { return 0; } __host__ __device__ static T f(T a, T b) { return a > b ? a : b; } };
struct OpMaxU32 { typedef u32 T; __host__ __device__ static T id() { return 0; } __host__ __device__ static T f(T a, T b) { return a > b ? a : b; } };
// two independent 32-bit max-scans packed in one 64-bit word
struct OpMax2x32 {
	typedef u64 T;
	__host__ __device__ static T id() { return 0; }
	__host__ __device__ static T f(T a, T b) {
		u32 ah = (u32) (a >> 32), al = (u32) a, bh = (u32) (b >> 32), bl = (u32) b;
		return ((u64) (ah > bh ? ah : bh) << 32) | (al > bl ? al : bl);
	}
};

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 32;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

template <class Op> __device__ __forceinline__ typename Op::T warp_incl_scan(typename Op::T v) {
	typedef typename Op::T T;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		T t = __shfl_up_sync(0xffffffffu, v, o);
		if ((int) lane_id() >= o) v = Op::f(t, v);
	}
	return v;
}

// Exclusive scan of one value per thread across the block; returns the exclusive
// prefix for this thread and the block total in *total. smem: 32 entries of T.
template <class Op> __device__ __forceinline__ typename Op::T block_excl_scan(typename Op::T v, typename Op::T *sm, typename Op::T *total) {
	typedef typename Op::T T;
	const u32 w = threadIdx.x >> 5, l = lane_id(), nw = blockDim.x >> 5;
	T inc = warp_incl_scan<Op>(v);
	if (l == 31) sm[w] = inc;
	__syncthreads();
	if (w == 0) {
		T x = l < nw ? sm[l] : Op::id();
		T xi = warp_incl_scan<Op>(x);
		sm[l] = xi; // inclusive over warps
	}
	__syncthreads();
	T wprefix = w ? sm[w - 1] : Op::id();
	T exc = __shfl_up_sync(0xffffffffu, inc, 1);
	if (l == 0) exc = Op::id();
	*total = sm[nw - 1];
	T r = Op::f(wprefix, exc);
	__syncthreads();
	return r;
}

template <class Op> __global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(const typename Op::T *__restrict__ in, u64 n, typename Op::T *__restrict__ partial) {
	typedef typename Op::T T;
	__shared__ T sm[32];
	u64 base = (u64) blockIdx.x * SCAN_TILE + (u64) threadIdx.x * SCAN_ITEMS;
	T acc = Op::id();
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) if (base + i < n) acc = Op::f(acc, in[base + i]);
	T tot;
	block_excl_scan<Op>(acc, sm, &tot);
	if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

// single block: exclusive scan of partial[0..nt) in place; total -> *total_out (may be null)
template <class Op> __global__ void __launch_bounds__(1024) scan_partials_kernel(typename Op::T *__restrict__ partial, u32 nt, typename Op::T *__restrict__ total_out) {
	typedef typename Op::T T;
	__shared__ T sm[32];
	T carry = Op::id();
	for (u32 base = 0; base < nt; base += 1024) {
		u32 i = base + threadIdx.x;
		T v = i < nt ? partial[i] : Op::id();
		T tot;
		T ex = block_excl_scan<Op>(v, sm, &tot);
		if (i < nt) partial[i] = Op::f(carry, ex);
		carry = Op::f(carry, tot);
	}
	if (threadIdx.x == 0 && total_out) *total_out = carry;
}

template <class Op, bool EXCL> __global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const typename Op::T *in, typename Op::T *out, u64 n, const typename Op::T *__restrict__ partial) {
	typedef typename Op::T T;
	__shared__ T sm[32];
	u64 base = (u64) blockIdx.x * SCAN_TILE + (u64) threadIdx.x * SCAN_ITEMS;
	T v[SCAN_ITEMS];
	T acc = Op::id();
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = base + i < n ? in[base + i] : Op::id(); acc = Op::f(acc, v[i]); }
	T tot;
	T pre = block_excl_scan<Op>(acc, sm, &tot);
	pre = Op::f(partial[blockIdx.x], pre);
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) {
		T inc = Op::f(pre, v[i]);
		if (base + i < n) out[base + i] = EXCL ? pre : inc;
		pre = inc;
	}
}

// Single-pass scan with decoupled look-back: tiles take tickets in launch order, publish
// their aggregate, and warp 0 walks back over the predecessors (32 at a time) until it meets
// a tile whose inclusive prefix is already known.  Traffic: one read + one write of the data.
// Every operator used here is commutative, so the walk may combine in any order.
// state per tile: flag (0 nothing, 1 aggregate, 2 inclusive prefix), agg, incl.
template <class Op> __device__ __forceinline__ typename Op::T warp_reduce_op(typename Op::T v) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v = Op::f(v, __shfl_xor_sync(0xffffffffu, v, o));
	return v;
}

// SUBIDX (32-bit sums / maxima only): the element index is subtracted from what is written (PLCP from the running
// maximum of position + irreducible LCP, suffix.cu: saves a pass over the array)
template <class Op, bool EXCL, bool SUBIDX = false> __global__ void __launch_bounds__(SCAN_THREADS)
scan_lookback_kernel(const typename Op::T *in, typename Op::T *out, u64 n, u32 *flag, typename Op::T *agg, typename Op::T *incl, u32 *ticket, typename Op::T *total_out) {
	typedef typename Op::T T;
	__shared__ T sm[32];
	__shared__ T s_prefix;
	__shared__ u32 s_tile;
	if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
	__syncthreads();
	const u32 tile = s_tile;
	const u32 ntiles = gridDim.x;
	// warp-striped: warp w owns 32*SCAN_ITEMS consecutive elements, lane l takes element i*32+l
	// of round i (coalesced); v[i] = scan value inside the warp's chunk (exclusive if EXCL)
	const u32 wi = threadIdx.x >> 5, li = threadIdx.x & 31;
	const u64 base = (u64) tile * SCAN_TILE + (u64) wi * (32 * SCAN_ITEMS) + li;
	T v[SCAN_ITEMS];
	T carry = Op::id();
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) {
		const u64 idx = base + (u64) i * 32;
		T x = idx < n ? in[idx] : Op::id();
		T inc = Op::f(carry, warp_incl_scan<Op>(x));
		if (EXCL) {
			T ex = __shfl_up_sync(0xffffffffu, inc, 1);
			v[i] = li == 0 ? carry : ex;
		} else v[i] = inc;
		carry = __shfl_sync(0xffffffffu, inc, 31);
	}
	// carry = total of the warp's chunk (same on every lane); scan the warp totals
	if (li == 0) sm[wi] = carry;
	__syncthreads();
	T pre = Op::id(), tot = Op::id();
	for (u32 ww = 0; ww < SCAN_THREADS / 32; ww++) { T t = sm[ww]; if (ww < wi) pre = Op::f(pre, t); tot = Op::f(tot, t); }
	if (threadIdx.x < 32) {
		const u32 l = threadIdx.x;
		T run = Op::id();
		if (tile == 0) {
			if (l == 0) { incl[0] = tot; __threadfence(); *reinterpret_cast<volatile u32 *>(&flag[0]) = 2u; }
		} else {
			if (l == 0) { agg[tile] = tot; __threadfence(); *reinterpret_cast<volatile u32 *>(&flag[tile]) = 1u; }
			i64 tb = (i64) tile - 1;
			for (;;) {
				i64 t = tb - (i64) l;
				u32 f = 2u;
				T c = Op::id();
				if (t >= 0) {
					do { f = *reinterpret_cast<volatile u32 *>(&flag[t]); } while (f == 0u);
					__threadfence();
					c = f == 2u ? *reinterpret_cast<volatile T *>(&incl[t]) : *reinterpret_cast<volatile T *>(&agg[t]);
				}
				const u32 m2 = __ballot_sync(0xffffffffu, f == 2u);
				const u32 first = (u32) (__ffs(m2) - 1); // m2 != 0 at the latest when the walk passes tile 0
				if (m2 && l > first) c = Op::id();
				run = Op::f(run, warp_reduce_op<Op>(c));
				if (m2) break;
				tb -= 32;
			}
			if (l == 0) { incl[tile] = Op::f(run, tot); __threadfence(); *reinterpret_cast<volatile u32 *>(&flag[tile]) = 2u; }
		}
		if (l == 0) {
			s_prefix = run;
			if (tile == ntiles - 1 && total_out) *total_out = Op::f(run, tot);
		}
	}
	__syncthreads();
	pre = Op::f(s_prefix, pre);
#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) {
		const u64 idx = base + (u64) i * 32;
		if (idx < n) out[idx] = SUBIDX ? (T) (Op::f(pre, v[i]) - (T) idx) : Op::f(pre, v[i]);
	}
}

// Device-wide scan. in/out may alias. total_out: optional device scalar receiving the grand total.
template <class Op, bool EXCL, bool SUBIDX = false> void device_scan(Dev &d, const typename Op::T *in, typename Op::T *out, u64 n, typename Op::T *total_out = nullptr) {
	typedef typename Op::T T;
	if (n == 0) { if (total_out) FBG_CUDA(cudaMemsetAsync(total_out, 0, sizeof(T), d.stream)); return; }
	u32 nt = div_up(n, SCAN_TILE);
	u64 mk = d.mark();
	u32 *flag = d.alloc<u32>((u64) nt + 1); // [nt] = ticket counter
	T *agg = d.alloc<T>(nt), *incl = d.alloc<T>(nt);
	d.begin(K_SCAN);
	FBG_CUDA(cudaMemsetAsync(flag, 0, ((u64) nt + 1) * sizeof(u32), d.stream));
	scan_lookback_kernel<Op, EXCL, SUBIDX><<<nt, SCAN_THREADS, 0, d.stream>>>(in, out, n, flag, agg, incl, flag + nt, total_out);
	FBG_LAUNCH_CHECK();
	d.end((double) n * sizeof(T) * 2.0);
	d.release(mk);
}

// -------------------------------------------------------------- radix sort
// LSD radix sort of (key, u32 value) pairs, digits of <= 8 bits.
// Per pass: (1) per-tile digit histogram staged in shared memory,
//           (2) one CTA per digit: exclusive scan of that digit's counts over the tiles
//               + the digit total (rs_binscan),
//           (3) stable scatter: every CTA scans the 256 digit totals itself, ranks by
//               warp-level __match_any, per-warp digit counters in shared memory, tile-local
//               reorder in shared memory so the global writes of one digit are contiguous.
constexpr int RS_THREADS = 512;          // scatter
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 8;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS; // 4096 keys per CTA
constexpr int RS_BINS = 256;
constexpr int RS_HTHREADS = 256;         // histogram: one thread per bin
constexpr int RS_HITEMS = RS_TILE / RS_HTHREADS;

template <typename K> __device__ __forceinline__ u32 rs_digit(K k, int shift, u32 mask) { return (u32) (k >> shift) & mask; }

template <typename K> __global__ void __launch_bounds__(RS_HTHREADS) rs_hist_kernel(const K *__restrict__ keys, u64 n, int shift, u32 mask, u32 ntiles, u32 *__restrict__ counts) {
	__shared__ u32 h[RS_BINS];
	h[threadIdx.x] = 0;
	__syncthreads();
	u64 base = (u64) blockIdx.x * RS_TILE;
#pragma unroll 4
	for (int r = 0; r < RS_HITEMS; r++) {
		u64 i = base + (u64) r * RS_HTHREADS + threadIdx.x;
		bool ok = i < n;
		u32 dg = ok ? rs_digit(keys[i], shift, mask) : 0xFFFFFFFFu;
		// warp-aggregated shared-memory atomics
		u32 peers = __match_any_sync(0xffffffffu, dg);
		if (ok && (peers & lanemask_lt()) == 0) atomicAdd(&h[dg], (u32) __popc(peers));
	}
	__syncthreads();
	counts[(u64) threadIdx.x * ntiles + blockIdx.x] = h[threadIdx.x];
}

// CTA d: counts[d][0..ntiles) -> exclusive prefix over the tiles (in place), totals[d] = sum
static __global__ void __launch_bounds__(1024) rs_binscan_kernel(u32 *__restrict__ counts, u32 ntiles, u64 *__restrict__ totals) {
	__shared__ u32 sm[32];
	u32 *row = counts + (u64) blockIdx.x * ntiles;
	u32 carry = 0;
	for (u32 base = 0; base < ntiles; base += 1024) {
		u32 i = base + threadIdx.x;
		u32 v = i < ntiles ? row[i] : 0u;
		u32 tot;
		u32 ex = block_excl_scan<OpSumU32>(v, sm, &tot);
		if (i < ntiles) row[i] = carry + ex;
		carry += tot;
	}
	if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

template <typename K> struct RsSmem {
	K keys[RS_TILE];
	u32 vals[RS_TILE];
	u32 warp_cnt[RS_WARPS][RS_BINS];
	u32 digit_base[RS_BINS];
	u32 gbase_lo[RS_BINS];
	u32 gbase_hi[RS_BINS];
	u64 scan_sm[32];
	u32 scan_sm32[32];
};

template <typename K, bool HAS_VAL> __global__ void __launch_bounds__(RS_THREADS, 2)
rs_scatter_kernel(const K *__restrict__ kin, const u32 *__restrict__ vin, K *__restrict__ kout, u32 *__restrict__ vout, u64 n, int shift, u32 mask, u32 ntiles,
                  const u32 *__restrict__ tile_off, const u64 *__restrict__ totals) {
	extern __shared__ __align__(16) unsigned char rs_raw[];
	RsSmem<K> &S = *reinterpret_cast<RsSmem<K> *>(rs_raw);
	const u32 tid = threadIdx.x, w = tid >> 5, l = tid & 31;
	for (int i = tid; i < RS_WARPS * RS_BINS; i += RS_THREADS) (&S.warp_cnt[0][0])[i] = 0;
	__syncthreads();
	const u64 tbase = (u64) blockIdx.x * RS_TILE;
	const u32 tcount = (u32) (n - tbase < (u64) RS_TILE ? n - tbase : (u64) RS_TILE);
	// warp-striped: warp w owns tile items [w*256, (w+1)*256), round r covers 32 consecutive keys
	K key[RS_ITEMS];
	u32 val[RS_ITEMS];
	u32 rank[RS_ITEMS];
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		u32 li = w * (32 * RS_ITEMS) + r * 32 + l;
		bool ok = li < tcount;
		key[r] = ok ? kin[tbase + li] : (K) ~(K) 0;
		if (HAS_VAL) val[r] = ok ? vin[tbase + li] : 0;
	}
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		u32 dg = rs_digit(key[r], shift, mask);
		u32 peers = __match_any_sync(0xffffffffu, dg);
		u32 prev = S.warp_cnt[w][dg];
		__syncwarp();
		if ((peers & lanemask_lt()) == 0) S.warp_cnt[w][dg] = prev + __popc(peers);
		__syncwarp();
		rank[r] = prev + __popc(peers & lanemask_lt());
	}
	__syncthreads();
	// thread d < 256: exclusive prefix over warps for digit d, digit totals in the tile, digit
	// bases in the tile and in the output (sum of the totals of the smaller digits + the tiles before)
	{
		u32 run = 0;
		u64 gtot = 0;
		if (tid < RS_BINS) {
#pragma unroll
			for (int ww = 0; ww < RS_WARPS; ww++) { u32 c = S.warp_cnt[ww][tid]; S.warp_cnt[ww][tid] = run; run += c; }
			gtot = totals[tid];
		}
		u32 tot;
		u32 ex = block_excl_scan<OpSumU32>(run, S.scan_sm32, &tot);
		u64 gt;
		u64 gex = block_excl_scan<OpSumU64>(gtot, S.scan_sm, &gt);
		if (tid < RS_BINS) {
			S.digit_base[tid] = ex;
			// padded keys (all ones) fall in the last bin after every real key, so the
			// global base of a digit only depends on the scanned real counts
			u64 g = gex + tile_off[(u64) tid * ntiles + blockIdx.x] - ex;
			S.gbase_lo[tid] = (u32) g; S.gbase_hi[tid] = (u32) (g >> 32);
		}
	}
	__syncthreads();
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		u32 dg = rs_digit(key[r], shift, mask);
		u32 pos = S.digit_base[dg] + S.warp_cnt[w][dg] + rank[r];
		S.keys[pos] = key[r];
		if (HAS_VAL) S.vals[pos] = val[r];
	}
	__syncthreads();
	for (u32 j = tid; j < tcount; j += RS_THREADS) {
		K k = S.keys[j];
		u32 dg = rs_digit(k, shift, mask);
		u64 g = (((u64) S.gbase_hi[dg] << 32) | S.gbase_lo[dg]) + j;
		kout[g] = k;
		if (HAS_VAL) vout[g] = S.vals[j];
	}
}

struct RadixBuffers {
	void *k[2]; u32 *v[2]; // ping-pong
	int cur = 0;
};

// Sorts bits [lo_bit, hi_bit) of the keys. Returns index (0/1) of the buffer holding the result.
template <typename K, bool HAS_VAL> int radix_sort(Dev &d, RadixBuffers &rb, u64 n, int lo_bit, int hi_bit) {
	if (n == 0 || hi_bit <= lo_bit) return rb.cur;
	int total = hi_bit - lo_bit;
	int passes = (total + 7) / 8;
	u32 ntiles = div_up(n, RS_TILE);
	u64 mk = d.mark();
	u32 *counts = d.alloc<u32>((u64) RS_BINS * ntiles);
	u64 *totals = d.alloc<u64>(RS_BINS);
	size_t smem = sizeof(RsSmem<K>);
	FBG_CUDA(cudaFuncSetAttribute(rs_scatter_kernel<K, HAS_VAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	int bit = lo_bit;
	for (int p = 0; p < passes; p++) {
		int bits = (hi_bit - bit + (passes - p) - 1) / (passes - p); // even split
		u32 mask = (1u << bits) - 1;
		const K *kin = (const K *) rb.k[rb.cur];
		K *kout = (K *) rb.k[rb.cur ^ 1];
		d.begin(K_RADIX_HIST);
		rs_hist_kernel<K><<<ntiles, RS_HTHREADS, 0, d.stream>>>(kin, n, bit, mask, ntiles, counts);
		rs_binscan_kernel<<<RS_BINS, 1024, 0, d.stream>>>(counts, ntiles, totals);
		FBG_LAUNCH_CHECK();
		d.end((double) n * sizeof(K) + (double) RS_BINS * ntiles * 12.0);
		d.note_launches(K_RADIX_HIST, 1);
		d.begin(K_RADIX_SCATTER);
		rs_scatter_kernel<K, HAS_VAL><<<ntiles, RS_THREADS, smem, d.stream>>>(kin, rb.v[rb.cur], kout, rb.v[rb.cur ^ 1], n, bit, mask, ntiles, counts, totals);
		FBG_LAUNCH_CHECK();
		d.end((double) n * (sizeof(K) + (HAS_VAL ? 4 : 0)) * 2.0);
		rb.cur ^= 1;
		bit += bits;
	}
	d.release(mk);
	return rb.cur;
}

} // namespace fbg
