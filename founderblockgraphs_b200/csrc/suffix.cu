// suffix.cu -- K1: suffix array, inverse suffix array and LCP of the gap-stripped
// text, by prefix doubling on radix passes.
//
// Replaces sdsl::construct(cst_sct3) as called at fbg.cpp:428 (divsufsort SA,
// construct_sa.hpp:183-190; PHI LCP, construct_lcp.hpp:216; BPS tree,
// cst_sct3.hpp:346).  The hot path only ever consumes ISA (csa.isa[], select_leaf),
// LCP-interval depths (depth/parent) and SA (sn); here they are flat arrays.
//
// Round 0 sorts all suffixes by a packed k-mer of K = floor(64/b) symbols
// (b = bits of the dense alphabet code).  Every later round sorts only the
// suffixes still tied ("active"), keyed by (rank[p], rank[p+h]); ranks are bucket
// start indices, so the new rank of a suffix is its old rank plus the offset of
// its sub-bucket.  LCP is computed at irreducible positions only (BWT[k] !=
// BWT[k-1]) by direct comparison and filled with PLCP[p] = max_{q<=p,irr}(PLCP[q]+q) - p.
//
// On a gapless MSA the later stages need less than that: prefix doubling stops as soon as every
// tied bucket lies in one column ("sufficient index"), and when the k-mer round itself is
// column-pure no per-position array is written at all ("virtual index"; see the block above
// k1_round0_dict and fcol.cu k2_from_dict).  fbg_build_index and the read-backs build the full arrays.
#include "pipeline.cuh"
#include "tilesort.cuh"
#include "dist.cuh"
#include <cstdlib>
#include <vector>

namespace fbg {

// ------------------------------------------------------------ alphabet / k-mers
__global__ void k1_presence_kernel(const u8 *__restrict__ T, u64 N, u32 *__restrict__ present) {
	__shared__ u32 s[256];
	s[threadIdx.x] = 0;
	__syncthreads();
	// 16 bytes per load, four loads in flight per thread (one byte per load left the kernel at 0.3 TB/s: latency-bound)
	const u64 nv = N / 16, stride = (u64) gridDim.x * blockDim.x;
	const uint4 *T4 = reinterpret_cast<const uint4 *>(T); // T is 256-byte aligned
	for (u64 i0 = (u64) blockIdx.x * blockDim.x + threadIdx.x; i0 < nv; i0 += 4 * stride) {
		uint4 v[4];
#pragma unroll
		for (int k = 0; k < 4; k++) { const u64 i = i0 + (u64) k * stride; v[k] = i < nv ? T4[i] : make_uint4(0, 0, 0, 0); }
#pragma unroll
		for (int k = 0; k < 4; k++) {
			if (i0 + (u64) k * stride < nv) {
				const u32 w[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
				for (int q = 0; q < 4; q++) { s[w[q] & 0xFFu] = 1; s[(w[q] >> 8) & 0xFFu] = 1; s[(w[q] >> 16) & 0xFFu] = 1; s[w[q] >> 24] = 1; }
			}
		}
	}
	for (u64 i = nv * 16 + (u64) blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) s[T[i]] = 1;
	__syncthreads();
	if (s[threadIdx.x]) present[threadIdx.x] = 1;
}

struct CodeTable { u8 code[256]; };

constexpr int KM_THREADS = 256;
constexpr int KM_ITEMS = 8;
constexpr int KM_TILE = KM_THREADS * KM_ITEMS;

// MODE 0: (key, position) of every suffix.  Modes 1-3 serve the key-range sharding of a
// multi-GPU build (dist.cuh): 1 = histogram of the top KM_HBITS key bits, 2 = per-CTA count
// of the keys whose bin is in [bin_lo, bin_hi), 3 = write those pairs, compacted in text order.
constexpr int KM_HBITS = 16;
struct KmerShard { u32 bin_lo, bin_hi; int hshift; const u64 *block_off; u32 *hist; u64 *block_cnt; };

// MODE 4: round 0 without a sort of all suffixes (k1_round0_dict below).  The rows of an MSA
// repeat each other, so the text holds few distinct k-mers (2-8 % of N on the bench
// inputs): every position enters its k-mer into an open-addressing table and leaves with
// (slot, arrival number inside the k-mer's bucket).
constexpr u64 KD_EMPTY = ~0ull;
constexpr u32 KD_MAX_PROBES = 4096;
struct KmerDict { u64 *key; u32 *cnt, *first; u32 mask; u32 *slot_of, *ord_of; u32 *overflow; u32 *slotcol; const u32 *col; u32 *impure; u32 ncol1; };
// MODE 5: the light form of MODE 4 for the "virtual" index of a gapless MSA (below): a position
// only makes sure its k-mer is in the table and that the bucket lies in ONE column (the first
// arrival leaves its column in slotcol, everybody else compares) -- in the steady state two L2
// reads per position and no atomic.  No arrival numbers, no smallest member, nothing per position.
constexpr u32 KD_NOCOL = 0xFFFFFFFFu;
__device__ __forceinline__ u32 kd_hash(u64 k) {
	k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
	return (u32) k;
}

template <int MODE> __global__ void __launch_bounds__(KM_THREADS) k1_kmer_kernel(const u8 *__restrict__ T, u64 N, CodeTable tab, int K, int b, KmerShard sh,
                                                                                 u64 *__restrict__ keys, u32 *__restrict__ vals, KmerDict dk) {
	__shared__ u8 sc[KM_TILE + 64];
	__shared__ u8 stab[256];
	__shared__ u32 ssum[32];
	stab[threadIdx.x] = tab.code[threadIdx.x];
	if ((MODE == 4 || MODE == 5) && threadIdx.x == 0) ssum[0] = *reinterpret_cast<volatile u32 *>(dk.overflow) | (MODE == 5 ? *reinterpret_cast<volatile u32 *>(dk.impure) : 0u);
	__syncthreads();
	if ((MODE == 4 || MODE == 5) && ssum[0]) return; // the table overflowed (or, MODE 5, a bucket spans two columns): the caller falls back
	const u64 base = (u64) blockIdx.x * KM_TILE;
	for (u32 i = threadIdx.x; i < KM_TILE + 64; i += KM_THREADS) {
		u64 g = base + i;
		sc[i] = g < N ? stab[T[g]] : (u8) 0;
	}
	__syncthreads();
	const u64 kmask = (K * b >= 64) ? ~0ull : ((1ull << (K * b)) - 1ull);
	u32 o = threadIdx.x * KM_ITEMS;
	u64 key = 0;
	for (int c = 0; c < K; c++) key = (key << b) | sc[o + c];
	u64 kk[KM_ITEMS];
	u32 mine = 0;
#pragma unroll
	for (int it = 0; it < KM_ITEMS; it++) {
		u64 g = base + o + it;
		kk[it] = key;
		if (MODE == 0) { if (g < N) { keys[g] = key; vals[g] = (u32) g; } }
		else if (MODE == 4 || MODE == 5) {
			// entered below, all items of the thread together
		} else {
			u32 bin = (u32) (key >> sh.hshift);
			if (MODE == 1) { if (g < N) atomicAdd(&sh.hist[bin], 1u); }
			else if (g < N && bin >= sh.bin_lo && bin < sh.bin_hi) mine |= 1u << it;
		}
		key = ((key << b) | sc[o + it + K]) & kmask;
	}
	if (MODE == 4 || MODE == 5) {
		// The eight table accesses of a thread are independent: first probes, arrival counters and
		// results each go out as a batch, so a thread waits for L2 three times, not 8 x 3 times.
		const u64 g0 = base + o;
		u32 sl[KM_ITEMS], od[KM_ITEMS];
		u64 cur[KM_ITEMS];
#pragma unroll
		for (int it = 0; it < KM_ITEMS; it++) {
			sl[it] = kd_hash(kk[it]) & dk.mask;
			cur[it] = g0 + it < N ? __ldcg(dk.key + sl[it]) : kk[it];
		}
#pragma unroll
		for (int it = 0; it < KM_ITEMS; it++) {
			if (g0 + it < N && cur[it] != kk[it]) {
				const u64 key = kk[it];
				u32 s = sl[it], probes = 0;
				u64 c = cur[it];
				for (;;) {
					// a stale EMPTY only costs the CAS, and a slot that holds a key never changes
					if (c == key) break;
					if (c == KD_EMPTY) {
						u64 prev = atomicCAS(reinterpret_cast<unsigned long long *>(dk.key + s), (unsigned long long) KD_EMPTY, (unsigned long long) key);
						if (prev == KD_EMPTY || prev == key) break;
					}
					s = (s + 1) & dk.mask;
					if (++probes > KD_MAX_PROBES) { *dk.overflow = 1u; s = U32_MAX; break; }
					c = __ldcg(dk.key + s);
				}
				sl[it] = s;
			}
		}
		if (MODE == 5) {
			// gapless rows: text position p = i (n + 1) + x, so the column needs no look-up
			// (ncol1 = n + 1; '#' -> n like the col array; the final sentinel -> n + 1)
			u32 cc[KM_ITEMS], sv[KM_ITEMS];
			u32 xc = (u32) (g0 % dk.ncol1);
#pragma unroll
			for (int it = 0; it < KM_ITEMS; it++) {
				const bool ok = g0 + it < N && sl[it] != U32_MAX;
				cc[it] = g0 + it == N - 1 ? dk.ncol1 : xc;
				xc = xc + 1 == dk.ncol1 ? 0u : xc + 1;
				sv[it] = ok ? __ldcg(dk.slotcol + sl[it]) : cc[it];
			}
#pragma unroll
			for (int it = 0; it < KM_ITEMS; it++) {
				if (g0 + it < N && sl[it] != U32_MAX) {
					u32 v = sv[it];
					if (v == KD_NOCOL) { v = atomicCAS(dk.slotcol + sl[it], KD_NOCOL, cc[it]); if (v == KD_NOCOL) v = cc[it]; }
					if (v != cc[it]) *dk.impure = 1u;
				}
			}
			return;
		}
		bool full = g0 + KM_ITEMS <= N;
#pragma unroll
		for (int it = 0; it < KM_ITEMS; it++) {
			od[it] = 0;
			if (g0 + it < N && sl[it] != U32_MAX) od[it] = atomicAdd(dk.cnt + sl[it], 1u);
			else full = false;
		}
		if (full) {
			// 32 contiguous bytes per thread and array
			uint4 *so = reinterpret_cast<uint4 *>(dk.slot_of + g0), *oo = reinterpret_cast<uint4 *>(dk.ord_of + g0);
			__stcs(so, make_uint4(sl[0], sl[1], sl[2], sl[3])); __stcs(so + 1, make_uint4(sl[4], sl[5], sl[6], sl[7]));
			__stcs(oo, make_uint4(od[0], od[1], od[2], od[3])); __stcs(oo + 1, make_uint4(od[4], od[5], od[6], od[7]));
		}
#pragma unroll
		for (int it = 0; it < KM_ITEMS; it++) {
			if (g0 + it < N && sl[it] != U32_MAX) {
				if (!full) { dk.slot_of[g0 + it] = sl[it]; dk.ord_of[g0 + it] = od[it]; }
				// one small member of the bucket orders the S / M lists by text position and stands for the bucket's
				// column(s) where any member will do: the minimum over the first 16 arrivals (CTAs run roughly in text
				// order, so that is the true minimum or close to it; an atomicMin for EVERY position cost 5 ms on cfg3,
				// the first arrival alone gave a worse layout and lost the 5 ms again in the sorters)
				if (od[it] < 16) atomicMin(dk.first + sl[it], (u32) (g0 + it));
			}
		}
	}
	if (MODE == 2 || MODE == 3) {
		u32 tot;
		u32 ex = block_excl_scan<OpSumU32>((u32) __popc(mine), ssum, &tot);
		if (MODE == 2) { if (threadIdx.x == 0) sh.block_cnt[blockIdx.x] = tot; }
		else {
			u64 q = sh.block_off[blockIdx.x] + ex;
#pragma unroll
			for (int it = 0; it < KM_ITEMS; it++)
				if (mine >> it & 1u) { keys[q] = kk[it]; vals[q] = (u32) (base + o + it); q++; }
		}
	}
}

// ------------------------------------------------------------- doubling rounds
// Tied suffixes live in two lists, each with every bucket contiguous:
//   S  buckets of <= BSMALL suffixes: (lp, lr, r2) = text position, bucket rank,
//      rank of the suffix h further on.  One CTA sorts a tile of whole buckets in
//      shared memory (k1_tile_sort_kernel).
//   B  larger buckets: 64-bit keys (lr << nbits | r2) + position, sorted by the
//      device-wide radix passes.
// A rank is the SA index of the bucket head, so a sub-bucket's new rank is the old
// rank plus its offset inside the (sorted) bucket.  All rank look-ups R[p+h] of a
// round are done by the compaction kernel of the round before, i.e. before any
// rank of the round changes.
// Where a suffix leaves with its final rank: the SA entry, and -- single-GPU full index -- the BWT character
// T[p - 1] next to it (0 for p = 0: no other position has that predecessor, the sentinel is the last byte).  The
// lists are in text order, so the T reads here are cheap, while the LCP stage would otherwise fetch
// T[SA[k] - 1] for every k at random: ncu showed 55 GB of DRAM reads for that on cfg3 (k1_irr_flag_kernel).
struct SaOut {
	u32 *SA; u8 *BWT; const u8 *T;
	__device__ __forceinline__ void put(u32 rank, u32 p) const {
		SA[rank] = p;
		if (BWT) BWT[rank] = p ? __ldg(T + p - 1) : (u8) 0;
	}
};

constexpr u32 BSMALL = 8192;     // largest bucket the shared-memory paths take (CAP_max / 2)
// ... and, when the M list is in use, the S list only keeps buckets below MID_MIN: buckets of
// MID_MIN .. BSMALL suffixes go to the M list (same record format), where one CTA streams one
// bucket (k1_bucket_sort_kernel) instead of sorting a window of whole buckets in shared memory
constexpr u32 MID_MIN = 1024;
constexpr u32 CLS_DROP = 0, CLS_SMALL = 1, CLS_BIG = 2, CLS_MID = 3, CLS_NONE = 0xFF;

// HH[t] = (bucket head index << 32) | sub-bucket head index, 0 where not a head;
// HR[na-1-t] = na - t at sub-bucket heads (a max-scan of HR gives the next head)
__global__ void k1_heads_kernel(const u64 *__restrict__ keys, u64 na, int hi_shift, u64 *__restrict__ HH, u32 *__restrict__ HR) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na) return;
	u64 k = keys[t];
	u64 kp = t ? keys[t - 1] : 0;
	bool hs = (t == 0) || (k != kp);
	bool hb = (t == 0) || (hi_shift < 64 && (k >> hi_shift) != (kp >> hi_shift));
	HH[t] = ((hb ? t : 0) << 32) | (hs ? t : 0);
	HR[na - 1 - t] = hs ? (u32) (na - t) : 0u;
}

// new rank = bucket rank + (sub-head index - bucket-head index); class by sub-bucket size
__global__ void __launch_bounds__(256) k1_update_kernel(const u64 *__restrict__ keys, const u32 *__restrict__ vals, const u64 *__restrict__ HH, const u32 *__restrict__ HR,
                                                        u64 na, int hi_shift, u32 rank_off, u32 mid_min, u32 *__restrict__ R, SaOut so, u32 *__restrict__ newr, u8 *__restrict__ cls, u32 *__restrict__ maxb) {
	__shared__ u32 smax[8];
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	u32 small_size = 0;
	if (t < na) {
		u64 k = keys[t];
		u64 hh = HH[t];
		u32 hb = (u32) (hh >> 32), hs = (u32) hh;
		u32 base = hi_shift < 64 ? (u32) (k >> hi_shift) : rank_off; // round 0: first global rank of this GPU's key range
		u32 nr = base + (hs - hb);
		u32 nh = (t + 1 < na) ? (u32) na - HR[na - 2 - t] : (u32) na; // HR scanned: 0 = no head after t
		u32 size = nh - hs;
		u32 c = size == 1 ? CLS_DROP : (size < mid_min ? CLS_SMALL : (size <= BSMALL ? CLS_MID : CLS_BIG)); // mid_min = BSMALL + 1: no M list
		const u32 pv = vals[t];
		R[pv] = nr;
		if (c == CLS_DROP) so.put(nr, pv); // a suffix alone in its bucket has its final rank: the SA entry is written as it leaves
		newr[t] = nr;
		cls[t] = (u8) c;
		if (c == CLS_SMALL && hs == (u32) t) small_size = size;
	}
	small_size = warp_reduce_max(small_size);
	if (lane_id() == 0) smax[threadIdx.x >> 5] = small_size;
	__syncthreads();
	if (threadIdx.x == 0) {
		u32 v = 0;
		for (int w = 0; w < 8; w++) v = smax[w] > v ? smax[w] : v;
		if (v) atomicMax(maxb, v);
	}
}

// ---- compaction of (lp, lr, cls) into the next S and B lists, with the rank look-up fused
constexpr int CP_THREADS = 256;
constexpr int CP_ITEMS = 16;
constexpr int CP_TILE = CP_THREADS * CP_ITEMS;
constexpr int CP_CHUNK = 4;      // items whose loads / look-ups / stores go out together in the write kernel
constexpr int CP_MINB = 6;       // CTAs per SM the write kernel's registers are held to (it is bound by memory latency)

// Round 0 lays the S list out by text position instead of by rank: buckets are ordered by
// their smallest member, so neighbouring tiles hold neighbouring MSA columns and the rank
// look-ups R[p+h] / updates R[p] of consecutive tiles share 32-byte sectors (the S list only
// needs every bucket contiguous, not any global order).
//   A[p0] = size of the small bucket whose smallest member is p0 (stable sort: its head)
__global__ void k1_bucket_size_kernel(const u64 *__restrict__ HH, const u32 *__restrict__ HR, const u32 *__restrict__ vals, const u8 *__restrict__ cls, u64 na, u32 *__restrict__ A) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na || cls[t] != CLS_SMALL || (u32) HH[t] != (u32) t) return;
	u32 nh = (t + 1 < na) ? (u32) na - HR[na - 2 - t] : (u32) na;
	A[vals[t]] = nh - (u32) t;
}
//   A exclusive-scanned: where each bucket starts in the S list
__global__ void k1_ordered_write_kernel(const u64 *__restrict__ HH, const u32 *__restrict__ vals, const u32 *__restrict__ newr, const u8 *__restrict__ cls, u64 na,
                                        const u32 *__restrict__ A, const u32 *__restrict__ R, u64 hnext, u64 N,
                                        u32 *__restrict__ Slp, u32 *__restrict__ Slr, u32 *__restrict__ Sr2) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na || cls[t] != CLS_SMALL) return;
	u32 hs = (u32) HH[t];
	u32 q = A[vals[hs]] + ((u32) t - hs);
	u32 p = vals[t];
	u64 pn = (u64) p + hnext;
	Slp[q] = p; Slr[q] = newr[t]; Sr2[q] = pn < N ? R[pn] : 0u;
}

// number of bytes of w equal to v (v < 128 or CLS_NONE)
__device__ __forceinline__ u32 bytes_equal(u32 w, u32 v) {
	const u32 x = w ^ (v * 0x01010101u);
	return (u32) __popc(~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u);
}
__device__ __forceinline__ u64 cls_count16(uint4 v, u32 ca, u32 cb) {
	// bytes equal to class ca -> low word, equal to class cb -> high word
	const u32 a = bytes_equal(v.x, ca) + bytes_equal(v.y, ca) + bytes_equal(v.z, ca) + bytes_equal(v.w, ca);
	const u32 b = bytes_equal(v.x, cb) + bytes_equal(v.y, cb) + bytes_equal(v.z, cb) + bytes_equal(v.w, cb);
	return ((u64) b << 32) | a;
}

// One compaction moves the survivors of two classes (ca, cb; CLS_NONE = unused) of one sorted list into
// the next round's lists.
__global__ void __launch_bounds__(CP_THREADS) k1_cls_count_kernel(const u8 *__restrict__ cls, u64 na, u32 ca, u32 cb, u64 *__restrict__ partial) {
	__shared__ u64 sm[32];
	u64 base = (u64) blockIdx.x * CP_TILE + (u64) threadIdx.x * CP_ITEMS;
	u64 c = 0;
	if (base + CP_ITEMS <= na) c = cls_count16(*reinterpret_cast<const uint4 *>(cls + base), ca, cb);
	else for (u64 i = base; i < na; i++) { u8 v = cls[i]; c += v == ca ? 1ull : (v == cb ? (1ull << 32) : 0ull); }
	u64 tot;
	block_excl_scan<OpSumU64>(c, sm, &tot);
	if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

constexpr int TGT_S = 0, TGT_M = 1, TGT_B = 2;
struct CompactOut {
	u32 *Slp, *Slr, *Sr2;  // next S list
	u32 *Mlp, *Mlr, *Mr2;  // next M list (same records)
	u64 *Bkey; u32 *Bval;  // next B list
};

// partial: exclusive-scanned per-tile (class a | class b << 32) counts; base_a / base_b: device words with
// the offsets this call appends at in its two target lists (null = 0).  A warp owns 512
// consecutive entries and walks them 32 at a time: coalesced reads, ballot/popc ranks,
// compacted (hence nearly coalesced) writes.
// PEER (a multi-GPU build in peer mode, dist.cuh): the rank of position q is read from the GPU that owns q's column
template <bool PEER>
__global__ void __launch_bounds__(CP_THREADS, CP_MINB) k1_cls_write_kernel(const u8 *__restrict__ cls, const u32 *__restrict__ lp, const u32 *__restrict__ lr, u64 na,
                                                                  const u64 *__restrict__ partial, const u64 *__restrict__ base_a, const u64 *__restrict__ base_b,
                                                                  u32 ca, u32 cb, int tgt_a, int tgt_b,
                                                                  const u32 *__restrict__ R, u64 hnext, u64 N, int nbits, CompactOut o, const u32 *__restrict__ col, const PeerView *__restrict__ pvp) {
	__shared__ u64 wtot[CP_THREADS / 32];
	__shared__ PeerView pv; // (dynamic indexing of a by-value kernel parameter would put a copy into every thread's local memory)
	if (PEER) {
		for (u32 q = threadIdx.x; q < sizeof(PeerView) / 4; q += CP_THREADS) reinterpret_cast<u32 *>(&pv)[q] = reinterpret_cast<const u32 *>(pvp)[q];
		// wtot's barrier below also publishes pv
	}
	const u32 w = threadIdx.x >> 5, l = threadIdx.x & 31;
	const u64 w0 = (u64) blockIdx.x * CP_TILE + (u64) w * (32 * CP_ITEMS);
	// pass 1: this warp's counts
	u32 cs = 0, cbn = 0;
	u8 c[CP_ITEMS];
#pragma unroll
	for (int i = 0; i < CP_ITEMS; i++) {
		u64 t = w0 + (u64) i * 32 + l;
		u8 v = t < na ? __ldcs(cls + t) : (u8) CLS_DROP; // (the lists are read once and written once: streaming hints keep R in L2)
		v = v == ca ? (u8) 1 : (v == cb ? (u8) 2 : (u8) 0); // 1 -> target a, 2 -> target b, 0 -> not moved by this call
		c[i] = v;
		cs += __popc(__ballot_sync(0xffffffffu, v == 1));
		cbn += __popc(__ballot_sync(0xffffffffu, v == 2));
	}
	if (l == 0) wtot[w] = ((u64) cbn << 32) | cs;
	__syncthreads();
	u64 ex = partial[blockIdx.x];
	for (u32 ww = 0; ww < w; ww++) ex += wtot[ww];
	u32 qa = (u32) ex + (base_a ? (u32) *base_a : 0u), qb = (u32) (ex >> 32) + (base_b ? (u32) *base_b : 0u);
	u32 *alp = tgt_a == TGT_S ? o.Slp : o.Mlp, *alr = tgt_a == TGT_S ? o.Slr : o.Mlr, *ar2 = tgt_a == TGT_S ? o.Sr2 : o.Mr2;
	u32 *blp = tgt_b == TGT_S ? o.Slp : o.Mlp, *blr = tgt_b == TGT_S ? o.Slr : o.Mlr, *br2 = tgt_b == TGT_S ? o.Sr2 : o.Mr2;
	// pass 2, CP_CHUNK items at a time: their loads, then their rank look-ups, then the writes, each as
	// one batch (the output arrays may alias the inputs as far as the compiler knows, so it would
	// not reorder an item's loads over the stores of the item before)
#pragma unroll
	for (int i0 = 0; i0 < CP_ITEMS; i0 += CP_CHUNK) {
		u32 pp[CP_CHUNK], rr[CP_CHUNK], r2[CP_CHUNK];
#pragma unroll
		for (int k = 0; k < CP_CHUNK; k++) {
			const u64 t = w0 + (u64) (i0 + k) * 32 + l;
			pp[k] = 0; rr[k] = 0;
			if (c[i0 + k]) { pp[k] = __ldcs(lp + t); rr[k] = __ldcs(lr + t); }
		}
		if (PEER) {
			// nearly every look-up stays on this GPU (a bucket's members lie in its columns, p + h leaves them only near
			// the range's end): one bit per position says so; the column -- hence the owner -- is looked up for the rest
			u32 nm[CP_CHUNK];
#pragma unroll
			for (int k = 0; k < CP_CHUNK; k++) {
				const u64 q = (u64) pp[k] + hnext;
				nm[k] = (c[i0 + k] && q < N) ? (__ldg(pv.notmine + (q >> 5)) >> (q & 31)) & 1u : 0u;
			}
#pragma unroll
			for (int k = 0; k < CP_CHUNK; k++) {
				const u64 q = (u64) pp[k] + hnext;
				r2[k] = 0u;
				if (c[i0 + k] && q < N) r2[k] = nm[k] ? __ldcg(pv.R[pv.owner(col[q])] + q) // never through L1: the owner rewrites it every round
				                                      : R[q];
			}
		} else {
#pragma unroll
			for (int k = 0; k < CP_CHUNK; k++) {
				const u64 q = (u64) pp[k] + hnext;
				r2[k] = (c[i0 + k] && q < N) ? R[q] : 0u;
			}
		}
#pragma unroll
		for (int k = 0; k < CP_CHUNK; k++) {
			const int i = i0 + k;
			const u32 ma = __ballot_sync(0xffffffffu, c[i] == 1), mb = __ballot_sync(0xffffffffu, c[i] == 2);
			if (c[i] == 1) { const u32 d = qa + __popc(ma & lanemask_lt()); __stcs(alp + d, pp[k]); __stcs(alr + d, rr[k]); __stcs(ar2 + d, r2[k]); }
			else if (c[i] == 2) {
				const u32 d = qb + __popc(mb & lanemask_lt());
				if (tgt_b == TGT_B) { o.Bkey[d] = ((u64) rr[k] << nbits) | r2[k]; o.Bval[d] = pp[k]; }
				else { __stcs(blp + d, pp[k]); __stcs(blr + d, rr[k]); __stcs(br2 + d, r2[k]); }
			}
			qa += __popc(ma); qb += __popc(mb);
		}
	}
}

// the totals of one compaction call are added to the fill counts of its two target lists
__global__ void k1_add_totals_kernel(const u64 *tot, u64 *cnt_a, u64 *cnt_b) {
	const u64 t = *tot;
	if (cnt_a) *cnt_a += t & 0xFFFFFFFFull;
	if (cnt_b) *cnt_b += t >> 32;
}

// ---- bucket-local sort of the S list
// CTA b owns the buckets whose head lies in [b*tile, (b+1)*tile) and loads the window
// [b*tile, b*tile + CAP); tile <= CAP - (largest bucket), so every owned bucket fits.
//
// The rows of an MSA are near copies of each other, so most of a bucket (the rows that agree
// at a column over h characters) lands in ONE bucket again h characters further on: nearly
// all of its r2 values are equal.  Every bucket therefore takes the r2 of its middle element
// as a pivot; elements equal to the pivot keep their relative order and leave at once
// (new rank = bucket rank + number of smaller elements), and only the others -- the
// "minority" -- go through the shared-memory sorting network, whose size follows their
// number.  Any pivot is correct; a bad one only sends more keys through the network.
//   key = bucket head | greater-than-pivot | r2 | window index of the element
// Shared memory: key u64[CAP] (first an r2 copy for the pivot look-ups), PXI u32[CAP+1]
// (inclusive prefix counts, minority | smaller << 16), bend u16[CAP] (bucket end by head).
template <int CAP> struct TileCfg {
	static constexpr int THREADS = CAP >= 8192 ? 1024 : CAP / 8;
	static constexpr int MINB = CAP >= 8192 ? 1 : CAP >= 4096 ? 2 : 5; // CTAs per SM the register budget is held to
	// windows up to 8192 slots also stage lr and lp in shared memory (one burst of asynchronous
	// copies at the start, every later access at shared-memory latency)
	static constexpr bool STAGE = CAP <= 8192;
	static constexpr size_t SMEM = (size_t) CAP * 8 + ((size_t) CAP + 4) * 4 + (size_t) CAP * 2 + (STAGE ? ((size_t) CAP + 8) * 4 + (size_t) CAP * 4 : 0);
};

__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gsrc) {
	const u32 d = (u32) __cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
	const u32 d = (u32) __cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
// 1-D bulk copy global -> shared through the TMA unit (cp.async.bulk, SASS UBLKCP): issued by ONE thread, no per-thread
// address arithmetic; completion is counted in bytes on an mbarrier.  Addresses and size are multiples of 16.
__device__ __forceinline__ void mbar_init(u64 *bar, u32 count) {
	const u32 a = (u32) __cvta_generic_to_shared(bar);
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count) : "memory");
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, u32 bytes) {
	const u32 a = (u32) __cvta_generic_to_shared(bar);
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, u32 bytes, u64 *bar) {
	const u32 d = (u32) __cvta_generic_to_shared(smem_dst), a = (u32) __cvta_generic_to_shared(bar);
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d), "l"(gsrc), "r"(bytes), "r"(a) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64 *bar, u32 parity) {
	const u32 a = (u32) __cvta_generic_to_shared(bar);
	u32 ok;
	do {
		asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(a), "r"(parity) : "memory");
	} while (!ok);
}
__device__ __forceinline__ void cp_async_wait_all() {
	asm volatile("cp.async.commit_group;" ::: "memory");
	asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// inclusive "index + 1 of the last flagged element at or before me" over a 32-element chunk
// whose lane 0 holds element `base`; carry = the value flowing in from the chunks before
__device__ __forceinline__ u32 ts_last_flag(u32 fm, u32 base, u32 &carry) {
	const u32 m = fm & (lanemask_lt() | (1u << lane_id()));
	const u32 inc = m ? base + (32u - (u32) __clz(m)) : carry;
	if (fm) carry = base + (32u - (u32) __clz(fm));
	return inc;
}

template <int CAP> __global__ void __launch_bounds__(TileCfg<CAP>::THREADS, TileCfg<CAP>::MINB)
k1_tile_sort_kernel(const u32 *__restrict__ lp, const u32 *__restrict__ lr, const u32 *__restrict__ r2, u64 na, u32 tile,
                    u32 *__restrict__ R, SaOut so, u32 *__restrict__ lp_out, u32 *__restrict__ lr_out, u8 *__restrict__ cls_out, u32 *__restrict__ scal) {
	constexpr int THREADS = TileCfg<CAP>::THREADS;
	constexpr int ITEMS = CAP / THREADS;
	constexpr int NW = THREADS / 32;
	constexpr int WCH = 32 * ITEMS;
	constexpr int LOGC = CAP == 16384 ? 14 : CAP == 8192 ? 13 : CAP == 4096 ? 12 : CAP == 2048 ? 11 : CAP == 1024 ? 10 : -1;
	static_assert(LOGC > 0, "unsupported tile capacity");
	extern __shared__ __align__(16) unsigned char k1_raw[];
	u64 *key = reinterpret_cast<u64 *>(k1_raw);
	u32 *r2s = reinterpret_cast<u32 *>(k1_raw);                                              // until the keys are written
	unsigned short *hbs = reinterpret_cast<unsigned short *>(k1_raw + (size_t) CAP * 4);     // until step 3 has read it
	u32 *PXI = reinterpret_cast<u32 *>(k1_raw + (size_t) CAP * 8);
	unsigned short *bend = reinterpret_cast<unsigned short *>(k1_raw + (size_t) CAP * 8 + ((size_t) CAP + 4) * 4);
	constexpr bool STAGE = TileCfg<CAP>::STAGE;
	u32 *lrs = reinterpret_cast<u32 *>(k1_raw + (size_t) CAP * 8 + ((size_t) CAP + 4) * 4 + (size_t) CAP * 2); // lrs[idx + 4] = lr of window element idx
	u32 *lps = lrs + CAP + 8;
	__shared__ u32 s_first, s_end, s_wtot[32];
	const u32 tid = threadIdx.x, w = tid >> 5, l = tid & 31;
	const u64 w0 = (u64) blockIdx.x * tile;
	const u32 wlen = (u32) (na - w0 < (u64) CAP ? na - w0 : (u64) CAP);
	const u32 tlen = tile < wlen ? tile : wlen;
	if (tid == 0) { s_first = U32_MAX; s_end = U32_MAX; }
	// The kernel alternates between short bursts of loads and long stretches of shared-memory
	// work, which leaves too few bytes in flight to cover DRAM latency: every CTA therefore
	// pulls the window of the CTA that will run about one wave later into L2.
	{
		constexpr u32 LINES = CAP / 32; // 128-byte lines per array and window
		const u64 pb = ((u64) blockIdx.x + (u64) NUM_SMS * (TileCfg<CAP>::MINB + 1)) * tile;
		for (u32 q = tid; q < 3 * LINES; q += THREADS) {
			const u32 arr = q / LINES;
			const u64 e = pb + (u64) (q % LINES) * 32;
			if (e < na) {
				const u32 *a = (arr == 0 ? lr : arr == 1 ? r2 : lp) + e;
				asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
			}
		}
	}
	if (STAGE) {
		// The three windows come in by bulk copies (TMA unit): one thread names source, destination and size, the
		// other 255 issue nothing (the per-thread cp.async form spent 13 % of the kernel's instructions here).
		// The tile length is a multiple of four elements, so window and lists are 16-byte aligned; the last
		// window of a list may end off a multiple of 16 bytes: those (at most three) elements are copied by hand.
		__shared__ __align__(8) u64 s_bar;
		const u32 nq = wlen >> 2;
		if (tid == 0) mbar_init(&s_bar, 1);
		__syncthreads();
		if (tid == 0 && nq) {
			mbar_expect_tx(&s_bar, 3 * 16 * nq);
			bulk_g2s(lrs + 4, lr + w0, 16 * nq, &s_bar);
			bulk_g2s(r2s, r2 + w0, 16 * nq, &s_bar);
			bulk_g2s(lps, lp + w0, 16 * nq, &s_bar);
		}
		const u32 idx = 4 * nq + tid;
		if (idx < wlen) { lrs[4 + idx] = lr[w0 + idx]; r2s[idx] = r2[w0 + idx]; lps[idx] = lp[w0 + idx]; }
		if (tid == 32 && w0 > 0) lrs[3] = lr[w0 - 1];
		if (nq) mbar_wait(&s_bar, 0);
	}
	__syncthreads();

	// 1. bucket heads in the window (warp w scans [w*WCH, (w+1)*WCH)); st = index + 1 of the
	//    last head at or before the element, as far as this warp can see.  All loads first:
	//    the scan below is one dependent chain.
	u32 st[ITEMS];
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const u32 idx = w * WCH + i * 32 + l;
		st[i] = idx < wlen ? (STAGE ? lrs[idx + 4] : lr[w0 + idx]) : 0u;
	}
	if (!STAGE) {
#pragma unroll
		for (int i = 0; i < ITEMS; i++) {
			const u32 idx = w * WCH + i * 32 + l;
			if (idx < wlen) r2s[idx] = r2[w0 + idx];
		}
	}
	u32 hmask = 0;
	u32 carry = 0, fmin = U32_MAX, emin = U32_MAX;
	u32 pl = 0;
	if (!STAGE && l == 0 && w * WCH < wlen && w0 + w * WCH > 0) pl = lr[w0 + w * WCH - 1];
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const u32 base = w * WCH + i * 32;
		const u32 idx = base + l;
		const u32 v = st[i];
		u32 pv;
		if (STAGE) pv = lrs[idx + 3]; // the element before (lrs[3] = the one before the window)
		else {
			pv = __shfl_up_sync(0xffffffffu, v, 1);
			if (l == 0) pv = pl;
			pl = __shfl_sync(0xffffffffu, v, 31);
		}
		const bool head = idx < wlen && (w0 + idx == 0 || pv != v);
		const u32 hm = __ballot_sync(0xffffffffu, head);
		hmask |= (head ? 1u : 0u) << i;
		st[i] = ts_last_flag(hm, base, carry);
		// first head inside the tile, first head behind it (warp-uniform)
		const u32 inl = tlen > base ? (tlen - base >= 32 ? 0xffffffffu : (1u << (tlen - base)) - 1u) : 0u;
		if ((hm & inl) && fmin == U32_MAX) fmin = base + (u32) __ffs(hm & inl) - 1;
		if ((hm & ~inl) && emin == U32_MAX) emin = base + (u32) __ffs(hm & ~inl) - 1;
	}
	if (l == 0) {
		s_wtot[w] = carry;
		if (fmin != U32_MAX) atomicMin(&s_first, fmin);
		if (emin != U32_MAX) atomicMin(&s_end, emin);
	}
	__syncthreads();
	const u32 first = s_first;
	if (first == U32_MAX) return; // no bucket starts in this tile
	u32 endidx = s_end;
	if (endidx == U32_MAX) {
		if (w0 + wlen == na) endidx = wlen;
		else { if (tid == 0) scal[1] = 1u; return; } // a bucket larger than CAP - tile: host contract broken
	}
	const u32 cnt = endidx - first;
	u32 wpre = 0;
	for (u32 ww = 0; ww < w; ww++) { u32 x = s_wtot[ww]; wpre = x > wpre ? x : wpre; }

	// 2. every owned element posts its bucket head; the last element of a bucket posts the end
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const u32 idx = w * WCH + i * 32 + l;
		const u32 hm = __ballot_sync(0xffffffffu, (hmask >> i) & 1u);
		const u32 hn = i + 1 < ITEMS ? __ballot_sync(0xffffffffu, (hmask >> (i + 1 < ITEMS ? i + 1 : i)) & 1u) : 0u;
		if (idx >= first && idx < endidx) {
			const u32 hb = (st[i] > wpre ? st[i] : wpre) - 1;
			hbs[idx] = (unsigned short) hb;
			bool last;
			if (idx + 1 == endidx) last = true;
			else if (l < 31) last = (hm >> (l + 1)) & 1u;
			else if (i + 1 < ITEMS) last = hn & 1u;
			else last = STAGE ? lrs[idx + 5] != lrs[idx + 4] : lr[w0 + idx + 1] != lr[w0 + idx];
			if (last) bend[hb] = (unsigned short) (idx + 1);
		}
	}
	__syncthreads();

	// From here on the cnt owned elements are dealt out evenly: warp w takes
	// [first + w*per, first + (w+1)*per), per a multiple of 32.
	const u32 per = ((cnt + NW * 32 - 1) / (NW * 32)) * 32;
	const u32 wb = first + w * per;

	// 3. pivot of the bucket, minority / smaller flags, their prefix counts over the owned range
	u32 mmask = 0, gmask = 0, run = 0;
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		if ((u32) i * 32 < per) {
			const u32 idx = wb + i * 32 + l;
			bool mn = false, ls = false;
			if (idx < endidx) {
				const u32 hb = hbs[idx];
				st[i] = hb;
				const u32 pv = r2s[hb + (((u32) bend[hb] - hb) >> 1)];
				const u32 rr = r2s[idx];
				mn = rr != pv; ls = rr < pv;
			}
			const u32 bm = __ballot_sync(0xffffffffu, mn), bl = __ballot_sync(0xffffffffu, ls);
			const u32 lt = lanemask_lt();
			if (idx < endidx) PXI[idx + 1] = run + (u32) __popc(bm & lt) + (mn ? 1u : 0u) + (((u32) __popc(bl & lt) + (ls ? 1u : 0u)) << 16); // inside the warp, for now
			run += (u32) __popc(bm) + ((u32) __popc(bl) << 16);
			mmask |= (mn ? 1u : 0u) << i;
			gmask |= ((mn && !ls) ? 1u : 0u) << i;
		}
	}
	if (l == 0) s_wtot[w] = run;
	if (tid == 0) PXI[first] = 0;
	__syncthreads();
	u32 ppre = 0, ptot = 0;
	for (u32 ww = 0; ww < NW; ww++) { u32 x = s_wtot[ww]; if (ww < w) ppre += x; ptot += x; }
	const u32 nmin = ptot & 0xFFFFu;
	if (ppre) {
#pragma unroll
		for (int i = 0; i < ITEMS; i++)
			if ((u32) i * 32 < per && wb + i * 32 + l < endidx) PXI[wb + i * 32 + l + 1] += ppre;
	}
	__syncthreads();

	// 4. elements equal to their pivot leave; the minority writes its keys (r2s and hbs are
	//    dead: their last reads lie before the two barriers above)
	u32 mx = 0;
	constexpr int P4C = ITEMS < 8 ? ITEMS : 8; // global loads of a chunk of items go out together
#pragma unroll
	for (int i0 = 0; i0 < ITEMS; i0 += P4C) {
		u32 pv4[P4C], rv4[P4C];
#pragma unroll
		for (int c = 0; c < P4C; c++) {
			const int i = i0 + c;
			const u32 idx = wb + i * 32 + l;
			pv4[c] = 0; rv4[c] = 0;
			if ((u32) i * 32 < per && idx < endidx) {
				if ((mmask >> i) & 1u) rv4[c] = r2[w0 + idx];
				else if (STAGE) { pv4[c] = lps[idx]; rv4[c] = lrs[idx + 4]; }
				else { pv4[c] = lp[w0 + idx]; rv4[c] = lr[w0 + idx]; }
			}
		}
#pragma unroll
		for (int c = 0; c < P4C; c++) {
			const int i = i0 + c;
			const u32 idx = wb + i * 32 + l;
			if ((u32) i * 32 < per && idx < endidx) {
				const u32 hb = st[i];
				const u32 x = PXI[idx];
				if ((mmask >> i) & 1u) {
					key[ts_sw(x & 0xFFFFu)] = ((u64) hb << (33 + LOGC)) | ((u64) ((gmask >> i) & 1u) << (32 + LOGC)) | ((u64) rv4[c] << LOGC) | (u64) idx;
				} else {
					const u32 end = bend[hb];
					const u32 xh = PXI[hb], xe = PXI[end];
					const u32 nless = (xe >> 16) - (xh >> 16);
					const u32 neq = (end - hb) - ((xe & 0xFFFFu) - (xh & 0xFFFFu));
					const u32 out = hb + nless + (idx - hb) - ((x & 0xFFFFu) - (xh & 0xFFFFu));
					const u32 p = pv4[c];
					const u32 nr = rv4[c] + nless;
					const bool single = neq == 1;
					if (nless) R[p] = nr;           // otherwise the rank is the one the suffix already has
					if (single) so.put(nr, p);
					lp_out[w0 + out] = p;
					lr_out[w0 + out] = nr;
					cls_out[w0 + out] = (u8) (single ? CLS_DROP : CLS_SMALL);
					if (!single) mx = neq > mx ? neq : mx;
				}
			}
		}
	}
	if (nmin == 0) {
		mx = warp_reduce_max(mx);
		if (l == 0 && mx > *reinterpret_cast<volatile u32 *>(scal)) atomicMax(&scal[0], mx);
		return;
	}
	u32 net = 32;
	while (net < nmin) net <<= 1;
	for (u32 slot = nmin + tid; slot < net; slot += THREADS) key[ts_sw(slot)] = ~0ull;
	__syncthreads();

	// 5. sort the minority
	tile_bitonic_sort_rt<THREADS>(key, net);

	// 6. sub-bucket heads among them (dealt out like the owned range): st = index + 1 of the
	//    last sub-head at or before the element
	const u32 per2 = ((nmin + NW * 32 - 1) / (NW * 32)) * 32;
	const u32 sb = w * per2;
	carry = 0;
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		if ((u32) i * 32 < per2) {
			const u32 s = sb + i * 32 + l;
			bool sh = false;
			if (s < nmin) sh = s == 0 || (key[ts_sw(s)] >> LOGC) != (key[ts_sw(s - 1)] >> LOGC);
			st[i] = ts_last_flag(__ballot_sync(0xffffffffu, sh), sb + i * 32, carry);
		}
	}
	if (l == 0) s_wtot[w] = carry;
	__syncthreads();
	wpre = 0;
	for (u32 ww = 0; ww < w; ww++) { u32 x = s_wtot[ww]; wpre = x > wpre ? x : wpre; }

	// 7. ranks, classes, outputs of the minority: smaller ones in front of the bucket's equal
	//    elements, greater ones behind them
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const u32 s = sb + i * 32 + l;
		if ((u32) i * 32 < per2 && s < nmin) {
			const u64 k = key[ts_sw(s)];
			const u32 hs = (st[i] > wpre ? st[i] : wpre) - 1;
			const u32 hb = (u32) (k >> (33 + LOGC));
			const u32 e = (u32) k & (u32) (CAP - 1);
			const bool last = (s + 1 >= nmin) || ((key[ts_sw(s + 1)] >> LOGC) != (k >> LOGC));
			const u32 end = bend[hb];
			const u32 xh = PXI[hb], xe = PXI[end];
			const u32 pmh = xh & 0xFFFFu;
			const u32 neq = (end - hb) - ((xe & 0xFFFFu) - pmh);
			const u32 off = ((k >> (32 + LOGC)) & 1ull) ? neq : 0u;
			const u32 out = hb + (s - pmh) + off;
			const u32 hrel = (hs - pmh) + off;     // sub-bucket head, relative to the bucket head
			const u32 nr = (STAGE ? lrs[hb + 4] : lr[w0 + hb]) + hrel;
			const u32 p = STAGE ? lps[e] : lp[w0 + e];
			const bool single = hs == s && last;
			if (hrel) R[p] = nr;
			if (single) so.put(nr, p);
			lp_out[w0 + out] = p;
			lr_out[w0 + out] = nr;
			cls_out[w0 + out] = (u8) (single ? CLS_DROP : CLS_SMALL);
			if (last && s != hs) { u32 sz = s - hs + 1; mx = sz > mx ? sz : mx; }
		}
	}
	mx = warp_reduce_max(mx);
	if (l == 0 && mx > *reinterpret_cast<volatile u32 *>(scal)) atomicMax(&scal[0], mx);
}

template <int CAP> static void launch_tile_sort(Dev &d, const u32 *lp, const u32 *lr, const u32 *r2, u64 na, u32 maxb,
                                                u32 *R, SaOut SA, u32 *lp_out, u32 *lr_out, u8 *cls_out, u32 *scal) {
	u32 tile = ((u32) CAP - maxb) & ~3u; // a multiple of four: 16-byte aligned windows for the staging copies
	const size_t smem = TileCfg<CAP>::SMEM;
	FBG_CUDA(cudaFuncSetAttribute(k1_tile_sort_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	k1_tile_sort_kernel<CAP><<<div_up(na, tile), TileCfg<CAP>::THREADS, smem, d.stream>>>(lp, lr, r2, na, tile, R, SA, lp_out, lr_out, cls_out, scal);
	FBG_LAUNCH_CHECK();
}

// Appends the survivors of classes ca / cb of one sorted list to the next round's lists tgt_a / tgt_b.
// cnt_a / cnt_b: device words with the fill counts of those lists: read as the append offsets, then
// raised by this call's totals (null for an unused class).  tot: a scratch device word.
static void compact_lists(Dev &d, const u8 *cls, const u32 *lp, const u32 *lr, u64 na, u32 ca, int tgt_a, u64 *cnt_a, u32 cb, int tgt_b, u64 *cnt_b, u64 *tot,
                          const u32 *R, u64 hnext, u64 N, int nbits, CompactOut o, const u32 *col = nullptr, const PeerView *pv = nullptr) {
	if (na == 0) return;
	u32 nt = div_up(na, CP_TILE);
	u64 mk = d.mark();
	u64 *partial = d.alloc<u64>(nt);
	d.begin(K_DOUBLING_AUX);
	k1_cls_count_kernel<<<nt, CP_THREADS, 0, d.stream>>>(cls, na, ca, cb, partial);
	scan_partials_kernel<OpSumU64><<<1, 1024, 0, d.stream>>>(partial, nt, tot);
	if (pv) k1_cls_write_kernel<true><<<nt, CP_THREADS, 0, d.stream>>>(cls, lp, lr, na, partial, cnt_a, cnt_b, ca, cb, tgt_a, tgt_b, R, hnext, N, nbits, o, col, pv);
	else k1_cls_write_kernel<false><<<nt, CP_THREADS, 0, d.stream>>>(cls, lp, lr, na, partial, cnt_a, cnt_b, ca, cb, tgt_a, tgt_b, R, hnext, N, nbits, o, nullptr, nullptr);
	k1_add_totals_kernel<<<1, 1, 0, d.stream>>>(tot, cnt_a, cnt_b);
	FBG_LAUNCH_CHECK();
	d.end((double) na * (2.0 + 8.0 + 4.0 + 12.0));
	d.note_launches(K_DOUBLING_AUX, 3);
	d.release(mk);
}

__global__ void k1_active_total_kernel(const u64 *cnt, u64 *tot) {
	*tot = cnt[0] + cnt[1] + cnt[2]; // next S, M, B lists
}

// ---- bucket-streaming sort of the M list (buckets of MID_MIN .. BSMALL suffixes)
// Same job as k1_tile_sort_kernel -- order every bucket by r2, new rank = old rank + offset of the
// sub-bucket, class of every element, rank scatter, SA entries of single suffixes -- for buckets too
// large to sort several of per shared-memory window.  CTA b owns the buckets whose head lies in list slots
// [b * BS_TILE, (b + 1) * BS_TILE) and STREAMS each of them twice instead of holding it in shared memory:
//   pass A  finds the end of the bucket (lr changes) and counts the elements below / above the pivot
//           (the majority value of three probes; any value is correct, a bad one only sends more
//           keys through the network);
//   pass B  writes the elements equal to the pivot straight to their final slots (block-wide ballot
//           scan: they keep their order, which is text order) and collects the others -- the
//           minority -- as 64-bit keys (above-pivot | r2 | index in bucket) in shared memory;
//   then    the swizzled bitonic network (tilesort.cuh), sized to the minority, sub-bucket heads by
//           ballots, ranks / classes / outputs of the minority.
// Per element: 8 B (lr, r2) + 8 B (r2 again, mostly from L2, lp) read, 9 B written, 4 B rank scatter, and
// about a tenth of the instructions of the window kernel.
constexpr int BS_THREADS = 512;
constexpr int BS_TILE = 2048;
constexpr int BS_CH = BS_THREADS * 4;   // elements per streaming step
constexpr int BS_MAXHEADS = 32;         // bucket heads per BS_TILE slots (2 when every bucket has >= MID_MIN elements)
constexpr size_t BS_SMEM = (size_t) BSMALL * 8 + ((size_t) BSMALL / 32 + 1) * 4;

__global__ void __launch_bounds__(BS_THREADS, 3)
k1_bucket_sort_kernel(const u32 *__restrict__ lp, const u32 *__restrict__ lr, const u32 *__restrict__ r2, u64 na, u32 mid_min,
                      u32 *__restrict__ R, SaOut so, u32 *__restrict__ lp_out, u32 *__restrict__ lr_out, u8 *__restrict__ cls_out, u32 *__restrict__ scal) {
	extern __shared__ __align__(16) unsigned char bs_raw[];
	u64 *key = reinterpret_cast<u64 *>(bs_raw);
	u32 *bal = reinterpret_cast<u32 *>(bs_raw + (size_t) BSMALL * 8);
	__shared__ u32 s_heads[BS_MAXHEADS], s_nheads, s_red[3][BS_THREADS / 32], s_cnt[4 * (BS_THREADS / 32)], s_pre[4 * (BS_THREADS / 32)], s_tot;
	constexpr u32 NW = BS_THREADS / 32;
	const u32 tid = threadIdx.x, w = tid >> 5, l = tid & 31;
	const u64 w0 = (u64) blockIdx.x * BS_TILE;
	if (tid == 0) s_nheads = 0;
	__syncthreads();
#pragma unroll
	for (int j = 0; j < BS_TILE / BS_THREADS; j++) {
		const u64 t = w0 + (u64) j * BS_THREADS + tid;
		if (t < na) {
			const u32 v = lr[t];
			if (t == 0 || lr[t - 1] != v) { const u32 q = atomicAdd(&s_nheads, 1u); if (q < BS_MAXHEADS) s_heads[q] = (u32) (t - w0); }
		}
	}
	__syncthreads();
	const u32 nh = s_nheads;
	if (nh == 0) return;
	if (nh > BS_MAXHEADS) { if (tid == 0) scal[1] = 1u; return; } // smaller buckets than the host promised
	u32 mx = 0;
	for (u32 hi = 0; hi < nh; hi++) {
		const u64 h = w0 + s_heads[hi];
		// the bucket is walked in steps of BS_CH elements, each step waiting for its loads: pull everything the
		// two passes will read into L2 now (the end of the bucket is not known yet; what lies behind it is
		// the next bucket, which its own CTA is about to ask for)
		for (u32 q = tid; q < 3 * (BSMALL / 32); q += BS_THREADS) {
			const u32 arr = q / (BSMALL / 32);
			const u64 e = h + BS_CH * (arr < 2 ? 1u : 0u) + (u64) (q % (BSMALL / 32)) * 32;
			if (e < na && e < h + BSMALL) {
				const u32 *a = (arr == 0 ? lr : arr == 1 ? r2 : lp) + e;
				asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
			}
		}
		const u32 lrh = lr[h];
		u32 pv;
		{
			const u64 tb = h + mid_min / 3, tc = h + 2 * (mid_min / 3);
			const u32 pa = r2[h], pb = tb < na ? r2[tb] : pa, pc = tc < na ? r2[tc] : pa;
			pv = pb == pc ? pb : pa;
		}
		// pass A
		u32 cnt = 0, nl = 0, ng = 0;
		bool broken = false;
		for (u32 c0 = 0;; c0 += BS_CH) {
			u32 a[4], b[4];
#pragma unroll
			for (int j = 0; j < 4; j++) {
				const u64 t = h + c0 + (u32) j * BS_THREADS + tid;
				a[j] = ~lrh; b[j] = pv;
				if (t < na) { a[j] = lr[t]; b[j] = r2[t]; }
			}
			bool all_in = true;
#pragma unroll
			for (int j = 0; j < 4; j++) {
				const u64 t = h + c0 + (u32) j * BS_THREADS + tid;
				const bool in = t < na && a[j] == lrh;
				cnt += in ? 1u : 0u; nl += (in && b[j] < pv) ? 1u : 0u; ng += (in && b[j] > pv) ? 1u : 0u;
				all_in = all_in && in;
			}
			if (!__syncthreads_and(all_in ? 1 : 0)) break;
			if (c0 + BS_CH > BSMALL) { broken = true; break; } // a bucket above BSMALL: host contract broken
		}
		if (broken) { if (tid == 0) scal[1] = 1u; return; }
		cnt = warp_reduce_sum(cnt); nl = warp_reduce_sum(nl); ng = warp_reduce_sum(ng);
		if (l == 0) { s_red[0][w] = cnt; s_red[1][w] = nl; s_red[2][w] = ng; }
		__syncthreads();
		u32 size = 0, nless = 0, ngr = 0;
#pragma unroll
		for (u32 ww = 0; ww < NW; ww++) { size += s_red[0][ww]; nless += s_red[1][ww]; ngr += s_red[2][ww]; }
		if (size > BSMALL) { if (tid == 0) scal[1] = 1u; return; }
		const u32 nmin = nless + ngr, neq = size - nmin;
		const u32 cls_eq = neq == 1 ? CLS_DROP : (neq < mid_min ? CLS_SMALL : CLS_MID);
		const u32 req = lrh + nless; // rank of the elements equal to the pivot
		if (cls_eq == CLS_SMALL) mx = neq > mx ? neq : mx;
		// pass B
		u32 eq_run = 0, mn_run = 0;
		for (u32 c0 = 0; c0 < size; c0 += BS_CH) {
			u32 rv[4], pp[4];
#pragma unroll
			for (int j = 0; j < 4; j++) {
				const u32 q = c0 + (u32) j * BS_THREADS + tid;
				rv[j] = pv; pp[j] = 0;
				if (q < size) { rv[j] = r2[h + q]; pp[j] = lp[h + q]; }
			}
			u32 be[4], bm[4];
#pragma unroll
			for (int j = 0; j < 4; j++) {
				const u32 q = c0 + (u32) j * BS_THREADS + tid;
				const bool in = q < size;
				be[j] = __ballot_sync(0xffffffffu, in && rv[j] == pv);
				bm[j] = __ballot_sync(0xffffffffu, in && rv[j] != pv);
				if (l == 0) s_cnt[j * NW + w] = (u32) __popc(be[j]) | ((u32) __popc(bm[j]) << 16);
			}
			__syncthreads();
			if (w == 0) { // exclusive scan of the 4 * NW slab counts (two per lane)
				const u32 x = s_cnt[2 * l], y = s_cnt[2 * l + 1];
				const u32 inc = warp_incl_sum(x + y);
				s_pre[2 * l] = inc - x - y; s_pre[2 * l + 1] = inc - y;
				if (l == 31) s_tot = inc;
			}
			__syncthreads();
#pragma unroll
			for (int j = 0; j < 4; j++) {
				const u32 q = c0 + (u32) j * BS_THREADS + tid;
				if (q < size) {
					const u32 pre = s_pre[j * NW + w];
					if (rv[j] == pv) {
						const u64 out = h + nless + eq_run + (pre & 0xFFFFu) + (u32) __popc(be[j] & lanemask_lt());
						__stcs(lp_out + out, pp[j]); __stcs(lr_out + out, req); cls_out[out] = (u8) cls_eq;
						if (nless) R[pp[j]] = req;
						if (neq == 1) so.put(req, pp[j]);
					} else {
						const u32 mp = mn_run + (pre >> 16) + (u32) __popc(bm[j] & lanemask_lt());
						key[ts_sw(mp)] = ((u64) (rv[j] > pv ? 1u : 0u) << 45) | ((u64) rv[j] << 13) | (u64) q;
					}
				}
			}
			const u32 tot = s_tot;
			eq_run += tot & 0xFFFFu; mn_run += tot >> 16;
		}
		if (nmin) {
			u32 net = 32;
			while (net < nmin) net <<= 1;
			for (u32 slot = nmin + tid; slot < net; slot += BS_THREADS) key[ts_sw(slot)] = ~0ull;
			__syncthreads();
			tile_bitonic_sort_rt<BS_THREADS>(key, net);
			// sub-bucket heads of the sorted minority, one ballot word per 32 slots
			for (u32 s0 = 0; s0 < nmin; s0 += BS_THREADS) {
				const u32 sx = s0 + tid;
				const bool hd = sx < nmin && (sx == 0 || (key[ts_sw(sx)] >> 13) != (key[ts_sw(sx - 1)] >> 13));
				const u32 bw = __ballot_sync(0xffffffffu, hd);
				if (l == 0 && sx < nmin) bal[sx >> 5] = bw;
			}
			__syncthreads();
			const u32 ngrp = (nmin + 31) >> 5;
			for (u32 s0 = 0; s0 < nmin; s0 += BS_THREADS) {
				const u32 sx = s0 + tid;
				if (sx < nmin) {
					const u64 k = key[ts_sw(sx)];
					const u32 g = sx >> 5, bi = sx & 31;
					u32 m1 = bal[g] & (0xFFFFFFFFu >> (31 - bi)), g1 = g;
					while (m1 == 0) m1 = bal[--g1];                 // slot 0 is a head
					const u32 hs = g1 * 32 + 31 - (u32) __clz(m1);  // head of my sub-bucket
					u32 m2 = bi == 31 ? 0u : (bal[g] & (0xFFFFFFFFu << (bi + 1))), g2 = g;
					while (m2 == 0 && ++g2 < ngrp) m2 = bal[g2];
					const u32 he = m2 ? g2 * 32 + (u32) __ffs(m2) - 1 : nmin; // head of the next one
					const u32 sz = he - hs;
					const bool gt = (k >> 45) & 1ull;
					const u32 p = lp[h + ((u32) k & (BSMALL - 1))];
					const u64 out = h + (gt ? neq : 0u) + sx;
					const u32 nr = lrh + (gt ? neq : 0u) + hs;
					const u32 c = sz == 1 ? CLS_DROP : (sz < mid_min ? CLS_SMALL : CLS_MID);
					if (nr != lrh) R[p] = nr;
					if (sz == 1) so.put(nr, p);
					lp_out[out] = p; lr_out[out] = nr; cls_out[out] = (u8) c;
					if (c == CLS_SMALL && sx == hs) mx = sz > mx ? sz : mx;
				}
			}
		}
		__syncthreads(); // key, bal and the counters are used again by the next bucket
	}
	mx = warp_reduce_max(mx);
	if (l == 0 && mx > *reinterpret_cast<volatile u32 *>(scal)) atomicMax(&scal[0], mx);
}

static void launch_bucket_sort(Dev &d, const u32 *lp, const u32 *lr, const u32 *r2, u64 na, u32 mid_min,
                               u32 *R, SaOut SA, u32 *lp_out, u32 *lr_out, u8 *cls_out, u32 *scal) {
	FBG_CUDA(cudaFuncSetAttribute(k1_bucket_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) BS_SMEM));
	k1_bucket_sort_kernel<<<div_up(na, BS_TILE), BS_THREADS, BS_SMEM, d.stream>>>(lp, lr, r2, na, mid_min, R, SA, lp_out, lr_out, cls_out, scal);
	FBG_LAUNCH_CHECK();
}

// ---- sufficient index for gapless MSAs: stop as soon as every tied bucket is COLUMN-PURE
// What K2 takes from the index is, per column x, the runs of consecutive ranks inside
// C_x = {rank of p_i(x)} and minima of LCP over them (fcol.cu).  In a gapless MSA C_x is the set
// of ALL text positions of column x, so a bucket (= all suffixes sharing a prefix of length D)
// whose members lie in one column is a subset of C_x: it sits inside one run whatever the order
// of its members, the LCP between two of its members is >= D, and the LCP between a member and
// anything outside is < D and the same for every member.  Every minimum K2 takes ranges over at
// least one bucket boundary, so it never sees an LCP from inside a bucket: the order inside
// column-pure buckets and the exact LCP values inside them cannot change f.  Prefix doubling
// may therefore stop at the first depth D at which all buckets still tied are column-pure
// (cfg2: D = K = 21, straight after the k-mer round; MSAs with repeats across columns go on
// until the repeats are resolved).  The members of the remaining buckets get consecutive
// ranks in list order, PLCP = D for all but the first, and LCP at bucket heads and single
// suffixes is found by direct comparison (< D characters).  SA, ISA and PLCP then differ from
// the true arrays only inside those buckets; fbg_build_index and the read-back calls build the
// full index.  tests/test_gpu_parity.py::test_sufficient_index_* compare both on gapless inputs.
constexpr u64 LAZY_MAX_DEPTH = 96;   // direct LCP comparisons stay below this many characters
constexpr u32 PLCP_TODO = 0xFFFFFFFFu; // mark: bucket head or single suffix, LCP still to be found by comparison

// *impure = 1 when two neighbours of the list share the bucket rank but not the column
template <bool BLIST> __global__ void k1_pure_check_kernel(const u32 *__restrict__ lp, const u32 *__restrict__ lr, const u64 *__restrict__ bkey, int nbits, u64 na,
                                                           const u32 *__restrict__ col, u32 *__restrict__ impure) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t + 1 >= na) return;
	const u32 ra = BLIST ? (u32) (bkey[t] >> nbits) : lr[t], rb = BLIST ? (u32) (bkey[t + 1] >> nbits) : lr[t + 1];
	if (ra == rb && col[lp[t]] != col[lp[t + 1]]) *impure = 1u;
}
// hd[t] = t + 1 at the first entry of a bucket, else 0 (inclusive max-scan: list index + 1 of the bucket's first entry)
template <bool BLIST> __global__ void k1_pure_heads_kernel(const u32 *__restrict__ lr, const u64 *__restrict__ bkey, int nbits, u64 na, u32 *__restrict__ hd) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na) return;
	const u32 ra = BLIST ? (u32) (bkey[t] >> nbits) : lr[t];
	const u32 rp = t ? (BLIST ? (u32) (bkey[t - 1] >> nbits) : lr[t - 1]) : 0u;
	hd[t] = (t == 0 || ra != rp) ? (u32) t + 1u : 0u;
}
// consecutive ranks in list order; all but the first member are marked "inside"
template <bool BLIST> __global__ void k1_pure_final_kernel(const u32 *__restrict__ lp, const u32 *__restrict__ lr, const u64 *__restrict__ bkey, int nbits, u64 na,
                                                           const u32 *__restrict__ hd, u32 depth, u32 *__restrict__ R, u32 *__restrict__ SA, u32 *__restrict__ PLCP) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na) return;
	const u32 ra = BLIST ? (u32) (bkey[t] >> nbits) : lr[t];
	const u32 j = (u32) t - (hd[t] - 1u);
	const u32 p = lp[t];
	R[p] = ra + j;
	SA[ra + j] = p;
	if (j) PLCP[p] = depth;
}
// LCP of the sufficient index, in text order: members inside a bucket already hold the depth; a
// bucket head or single suffix (still marked) is compared with its predecessor in SA -- fewer
// than `depth` characters agree
__global__ void k1_lazy_lcp_kernel(const u8 *__restrict__ T, const u32 *__restrict__ SA, const u32 *__restrict__ ISA, u64 N, u32 *__restrict__ PLCP) {
	u64 p = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= N || PLCP[p] != PLCP_TODO) return;
	const u32 k = ISA[p];
	u32 len = 0;
	if (k > 0) {
		const u32 q = SA[k - 1];
		while (p + len < N && (u64) q + len < N && T[p + len] == T[q + len]) len++;
	}
	PLCP[p] = len;
}

// The "virtual" index: the distinct k-mers in rank order with the column of their bucket and the
// LCP (in characters) to the k-mer before -- the keys hold K order-preserving b-bit codes, first
// character in the top bits, so the LCP of two buckets is read off the XOR of their keys.
__global__ void k1_dict_virtual_kernel(const u64 *__restrict__ skey, const u32 *__restrict__ lv, const u32 *__restrict__ slotcol,
                                       u32 nd, int K, int b, u32 *__restrict__ dcol, u32 *__restrict__ dlcp) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t > nd) return;
	if (t == nd) { dlcp[t] = 0; return; } // nothing follows the last bucket
	dcol[t] = slotcol[lv[t]];
	u32 l = 0;
	if (t > 0) {
		const u64 x = skey[t] ^ skey[t - 1];
		l = (u32) ((__clzll((long long) x) - (64 - K * b)) / b);
	}
	dlcp[t] = l;
}
// Round 0 through the dictionary, speculating that every k-mer bucket is column-pure (the usual
// case for a gapless MSA): final ranks (bucket rank + arrival number), SA entries and PLCP marks
// straight from the table, in text order, and no S / B lists.  *impure is raised when a bucket
// holds two columns; the caller then emits the lists as usual (R, SA are rewritten).
__global__ void __launch_bounds__(256) k1_dict_emit_pure_kernel(const u32 *__restrict__ slot_of, const u32 *__restrict__ ord_of, const uint4 *__restrict__ info,
                                                                const u32 *__restrict__ col, u64 N, u32 depth, u32 *__restrict__ R, u32 *__restrict__ SA,
                                                                u32 *__restrict__ PLCP, u32 *__restrict__ impure) {
	u64 p = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= N) return;
	const uint4 v = __ldg(info + slot_of[p]);
	const u32 j = v.z == 1 ? 0u : ord_of[p];
	R[p] = v.x + j;
	SA[v.x + j] = (u32) p;
	PLCP[p] = j ? depth : PLCP_TODO;
	if (v.z > 1 && col[p] != __ldg(col + v.w)) *impure = 1u; // v.w = the bucket's smallest member
}

// ---- round 0 through the k-mer dictionary (single GPU)
__global__ void k1_dict_flag_kernel(const u64 *__restrict__ key, u32 cap, u32 *__restrict__ flag) {
	u32 s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s < cap) flag[s] = key[s] != KD_EMPTY ? 1u : 0u;
}
__global__ void k1_dict_list_kernel(const u64 *__restrict__ key, const u32 *__restrict__ pos, u32 cap, u64 *__restrict__ lk, u32 *__restrict__ lv) {
	u32 s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= cap) return;
	u64 k = key[s];
	if (k != KD_EMPTY) { u32 q = pos[s]; lk[q] = k; lv[q] = s; }
}
// Peer mode (one MSA across several GPUs, dist.cuh): the dictionary is built by every GPU, the tied buckets are
// dealt out by column range.  A bucket whose members all lie in the columns of one GPU belongs to that GPU
// (the usual case: one column, or a few neighbouring ones -- the character before a gap run often shares
// its k-mer with the column the run ends in).  A bucket that spans the ranges of several GPUs (a gap run
// across a range boundary, a repeat) is kept by EVERY GPU: each sorts it in full, identically, so whoever
// owns a member's column holds that member's up-to-date rank without any rank crossing the wire.
__global__ void k1_notmine_kernel(const u32 *__restrict__ col, u64 N, u32 x0, u32 x1, int last, u32 *__restrict__ bits) {
	const u64 p = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	bool nm = false;
	if (p < N) { const u32 c = __ldcs(col + p); nm = !(c >= x0 && (c < x1 || last)); }
	const u32 w = __ballot_sync(0xffffffffu, nm);
	if ((threadIdx.x & 31) == 0 && p < N) bits[p >> 5] = w;
}
// the text positions of row i whose column lies in [x0, x1): own_a[i] <= p < own_b[i] (the last GPU: through the row's '#',
// and the final sentinel after the last row)
__global__ void k1_own_ranges_kernel(const u32 *__restrict__ RK, const u32 *__restrict__ NG, const u32 *__restrict__ start, const u32 *__restrict__ rowlen,
                                     u64 m, u64 n, u32 x0, u32 x1, int last, u32 *__restrict__ own_a, u32 *__restrict__ own_b) {
	const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= m) return;
	auto rank_at = [&](u32 x) -> u32 {
		if (x >= n) return rowlen[i];
		const u64 t = x >> 5;
		return RK[t * m + i] + (u32) __popc(NG[t * m + i] & ((1u << (x & 31)) - 1u));
	};
	own_a[i] = start[i] + rank_at(x0);
	own_b[i] = last ? start[i] + rowlen[i] + 1u + (i == m - 1 ? 1u : 0u) : start[i] + rank_at(x1);
}
// (the N-wide pass compares two bytes -- the owner of the position's column, from a table of n + 2 entries, and the owner of
// the bucket's smallest member, found once per distinct k-mer -- instead of looking up first[s] and col[first[s]] per position)
__global__ void k1_owner_table_kernel(u32 ncols, PeerView pv, u8 *__restrict__ ownc) {
	u32 c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c < ncols) ownc[c] = (u8) pv.owner(c);
}
__global__ void k1_dict_own1_kernel(const u32 *__restrict__ lv, const u32 *__restrict__ first, const u32 *__restrict__ col, const u8 *__restrict__ ownc, u32 nd, u32 *__restrict__ own1) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nd) return;
	const u32 s = lv[t];
	own1[s] = ownc[col[first[s]]];
}
__global__ void k1_dict_spans_kernel(const u32 *__restrict__ slot_of, const u32 *__restrict__ col, const u8 *__restrict__ ownc, const u32 *__restrict__ own1, u64 N, u8 *__restrict__ spans) {
	u64 p = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= N) return;
	const u32 s = __ldcs(slot_of + p);
	if (__ldg(ownc + __ldcs(col + p)) != __ldg(own1 + s)) spans[s] = 1;
}
__global__ void k1_dict_mine_kernel(const u32 *__restrict__ lv, const u32 *__restrict__ own1, const u8 *__restrict__ spans, u32 nd, int me, u8 *__restrict__ mine, u32 *__restrict__ nspan) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nd) return;
	const u32 s = lv[t];
	mine[t] = (spans[s] || own1[s] == (u32) me) ? 1 : 0;
	if (spans[s]) atomicAdd(nspan, 1u);
}
// bucket sizes in k-mer order: all of them (their exclusive scan = the ranks) and the big ones (B list offsets);
// mine: null, or which buckets this GPU sorts
__global__ void k1_dict_counts_kernel(const u32 *__restrict__ lv, const u32 *__restrict__ cnt, const u8 *__restrict__ mine, u32 nd, u32 *__restrict__ call, u32 *__restrict__ cbig) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nd) return;
	u32 c = cnt[lv[t]];
	call[t] = c; cbig[t] = (c > BSMALL && (!mine || mine[t])) ? c : 0u;
}
// info[slot] = (rank, start of the bucket in its list, size, smallest member); the start of a B bucket is set here
// (B list in k-mer order), that of an S / M bucket by k1_dict_place_kernel
__global__ void k1_dict_assign_kernel(const u32 *__restrict__ lv, const u32 *__restrict__ cnt, const u32 *__restrict__ first, const u32 *__restrict__ rk, const u32 *__restrict__ bs,
                                      const u8 *__restrict__ mine, u32 nd, u32 rank_off, u32 mid_min, uint4 *__restrict__ info, u32 *__restrict__ rank_of_slot, u32 *__restrict__ maxb) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	u32 small = 0;
	if (t < nd) {
		u32 s = lv[t], c = cnt[s], f = first[s], start = 0;
		if (mine && !mine[t]) start = U32_MAX; // sorted by another GPU: no list entry here
		else if (c > BSMALL) start = bs[t];
		else if (c > 1 && c < mid_min) small = c;
		info[s] = make_uint4(rank_off + rk[t], start, c, f);
		rank_of_slot[s] = rank_off + rk[t];
	}
	small = warp_reduce_max(small);
	if (lane_id() == 0 && small) atomicMax(maxb, small);
}
// The S and M lists are laid out by text position: buckets ordered by their smallest member, so
// neighbouring tiles hold neighbouring MSA columns (see k1_bucket_size_kernel).  The distinct k-mers
// are sorted by that position (a few passes over nd keys, not a directory of N words): ok = key, ov = t
__global__ void k1_dict_firstkeys_kernel(const u32 *__restrict__ lv, const u32 *__restrict__ first, u32 nd, u64 *__restrict__ ok, u32 *__restrict__ ov) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nd) return;
	ok[t] = first[lv[t]]; ov[t] = t;
}
// in that order: sizes of the S-class and of the M-class buckets (their exclusive scans = the list offsets)
__global__ void k1_dict_sizes_kernel(const u32 *__restrict__ ov, const u32 *__restrict__ lv, const u32 *__restrict__ cnt, const u8 *__restrict__ mine, u32 nd, u32 mid_min, u32 *__restrict__ zs, u32 *__restrict__ zm) {
	u32 j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= nd) return;
	u32 c = cnt[lv[ov[j]]];
	if (mine && !mine[ov[j]]) c = 0;
	zs[j] = (c > 1 && c < mid_min) ? c : 0u;
	zm[j] = (c >= mid_min && c <= BSMALL) ? c : 0u;
}
__global__ void k1_dict_place_kernel(const u32 *__restrict__ ov, const u32 *__restrict__ lv, u32 nd, u32 mid_min, const u32 *__restrict__ zs, const u32 *__restrict__ zm, uint4 *__restrict__ info) {
	u32 j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= nd) return;
	const u32 s = lv[ov[j]];
	uint4 v = info[s];
	if (v.z > 1 && v.z <= BSMALL && v.y != U32_MAX) { v.y = v.z < mid_min ? zs[j] : zm[j]; info[s] = v; }
}
// what the sort-based round 0 leaves behind: ranks, SA entries of the suffixes alone in their
// bucket, the S, M and B lists of the next round with R[p+h] looked up (through the table: no
// rank has to be visible yet)
// PEER: this GPU only needs the ranks of its own columns [x0, x1) (the last GPU also takes '#' and the sentinel) and the
// members of the buckets it sorts; a position outside its columns matters only when its bucket spans GPUs
// MODE 0: one GPU, every position.  Peer mode: MODE 1 = this GPU's own positions (grid.y = row, its text range from
// k1_own_ranges_kernel: no thread for the (world - 1) / world of the text that belongs to the others), MODE 2 = the
// members of the buckets every GPU sorts that lie in other GPUs' columns (N-wide; returns at once when no bucket spans).
template <int MODE>
__global__ void __launch_bounds__(256) k1_dict_emit_kernel(const u32 *__restrict__ slot_of, const u32 *__restrict__ ord_of, const uint4 *__restrict__ info, const u32 *__restrict__ rank_of_slot,
                                                           u64 N, u64 hnext, int nbits, u32 mid_min, u32 *__restrict__ R, SaOut so, CompactOut o,
                                                           const u32 *__restrict__ notmine, const u8 *__restrict__ spans, const u32 *__restrict__ own_a, const u32 *__restrict__ own_b, const u32 *__restrict__ nspan) {
	u64 p = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (MODE == 1) {
		p += own_a[blockIdx.y];
		if (p >= own_b[blockIdx.y]) return;
	}
	if (MODE == 2 && *nspan == 0) return;
	if (p >= N) return;
	if (MODE == 2) {
		{
			if (!((notmine[p >> 5] >> (p & 31)) & 1u)) return; // mine: written by the MODE 1 pass
			if (!spans[slot_of[p]]) return;
			// a member of a bucket every GPU sorts: listed below; its rank and SA entry are its owner's business
			const uint4 v = __ldg(info + slot_of[p]);
			if (v.z == 1 || v.y == U32_MAX) return;
			const u64 pn = p + hnext;
			const u32 r2 = pn < N ? __ldg(rank_of_slot + slot_of[pn]) : 0u;
			const u32 q = v.y + ord_of[p];
			if (v.z < mid_min) { o.Slp[q] = (u32) p; o.Slr[q] = v.x; o.Sr2[q] = r2; }
			else if (v.z <= BSMALL) { o.Mlp[q] = (u32) p; o.Mlr[q] = v.x; o.Mr2[q] = r2; }
			else { o.Bkey[q] = ((u64) v.x << nbits) | r2; o.Bval[q] = (u32) p; }
			return;
		}
	}
	const uint4 v = __ldg(info + __ldcs(slot_of + p)); // (slot_of, ord_of, R stream through once: hints keep the table in L2)
	__stcs(R + p, v.x);
	if (v.z == 1) { so.put(v.x, (u32) p); return; }
	if (v.y == U32_MAX) return; // peer mode: another GPU sorts this bucket
	const u64 pn = p + hnext;
	const u32 r2 = pn < N ? __ldg(rank_of_slot + slot_of[pn]) : 0u; // 4-byte entries: a quarter of the L2 traffic of info[]
	const u32 q = v.y + __ldcs(ord_of + p);
	if (v.z < mid_min) { o.Slp[q] = (u32) p; o.Slr[q] = v.x; o.Sr2[q] = r2; }
	else if (v.z <= BSMALL) { o.Mlp[q] = (u32) p; o.Mlr[q] = v.x; o.Mr2[q] = r2; }
	else { o.Bkey[q] = ((u64) v.x << nbits) | r2; o.Bval[q] = (u32) p; }
}

// ---- the insert of MODE 4 in TILE order (texts with gaps: the RK / NG tables give the tiles).  In text order the 32
// positions of a warp hold 32 different k-mers: 32 probes, 32 counter atomics, 32 minimum atomics into a 1 GB
// table.  A CTA here takes 32 rows x the 32 columns of one bitmap word: the keys are built along the rows (text
// staged per row, coalesced), then a warp takes a COLUMN -- its 32 rows mostly hold ONE k-mer: the lanes with equal
// keys elect a leader (__match_any_sync), the leader alone probes the table, adds the size of its group to the
// bucket's counter and folds its position (the smallest of the group: the lowest row) into the bucket's minimum; the
// members take consecutive arrival numbers.  Slots and arrival numbers go back through shared memory and are
// written along the rows.  '#' and the final sentinel: k1_kmer_ends_kernel.
__device__ __forceinline__ u32 kd_insert(const KmerDict &dk, u64 key) {
	u32 s = kd_hash(key) & dk.mask, probes = 0;
	u64 c = __ldcg(dk.key + s);
	for (;;) {
		if (c == key) break;
		if (c == KD_EMPTY) {
			u64 prev = atomicCAS(reinterpret_cast<unsigned long long *>(dk.key + s), (unsigned long long) KD_EMPTY, (unsigned long long) key);
			if (prev == KD_EMPTY || prev == key) break;
		}
		s = (s + 1) & dk.mask;
		if (++probes > KD_MAX_PROBES) { *dk.overflow = 1u; return U32_MAX; }
		c = __ldcg(dk.key + s);
	}
	return s;
}
__global__ void __launch_bounds__(256) k1_kmer_tile_kernel(const u8 *__restrict__ T, u64 N, CodeTable tab, int K, int b, KmerDict dk,
                                                           const u32 *__restrict__ RK, const u32 *__restrict__ NG, const u32 *__restrict__ start, u64 m, u64 W,
                                                           const u8 *__restrict__ ownc, u32 *__restrict__ own1, u8 *__restrict__ spans) {
	// peer mode (ownc != null): the leader also leaves the GPU that owns its column with the bucket (first come) and
	// raises spans[slot] when the bucket already belongs to another GPU -- the N-wide span test for free
	__shared__ u64 sK[32][33];
	__shared__ u32 sS[32][33], sO[32][33];
	__shared__ u8 sT[32][64]; // the codes of a row's 32 + K - 1 <= 63 characters from its first position in this word
	__shared__ u8 stab[256];
	__shared__ u32 sNG[32], sP0[32], s_stop;
	const u32 w = threadIdx.x >> 5, lane = threadIdx.x & 31;
	// Blocks run word-fastest: a row group is (nearly) done before the next one starts, so the arrival numbers of a bucket
	// grow with the row as they do with the text-order insert.  (Row-group-fastest made the order of the 32-row groups
	// inside a bucket random; compaction and bucket sort of every later round then ran 10-20 % slower: neighbouring buckets
	// no longer ask for the ranks of a row at about the same time.)
	const u32 nwords = (u32) W;
	const u64 i0 = (u64) (blockIdx.x / nwords) * 32, t = blockIdx.x % nwords;
	stab[threadIdx.x] = tab.code[threadIdx.x];
	if (threadIdx.x == 0) s_stop = *reinterpret_cast<volatile u32 *>(dk.overflow);
	if (threadIdx.x < 32) {
		const u64 i = i0 + threadIdx.x;
		sNG[threadIdx.x] = i < m ? NG[t * m + i] : 0u;
		sP0[threadIdx.x] = i < m ? start[i] + RK[t * m + i] : 0u;
	}
	__syncthreads();
	if (s_stop) return; // the table overflowed: the caller falls back
	const u64 kmask = (K * b >= 64) ? ~0ull : ((1ull << (K * b)) - 1ull);
	// phase 1: rows -- text codes, then one key per position
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 rr = w + 8 * k;
		const u32 cnt = (u32) __popc(sNG[rr]);
		if (cnt) { // (uniform over the warp)
			const u64 g0 = (u64) sP0[rr] + lane, g1 = g0 + 32;
			sT[rr][lane] = g0 < N ? stab[T[g0]] : (u8) 0;
			sT[rr][lane + 32] = g1 < N ? stab[T[g1]] : (u8) 0;
			__syncwarp();
			if (lane < cnt) {
				u64 key = 0;
				for (int c = 0; c < K; c++) key = (key << b) | sT[rr][lane + c];
				sK[rr][lane] = key & kmask;
			}
		}
	}
	__syncthreads();
	// phase 2: columns (lane = row)
	const u32 ng = sNG[lane], p0 = sP0[lane];
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 c = w + 8 * k;
		const bool act = (ng >> c) & 1u;
		const u32 am = __ballot_sync(0xffffffffu, act);
		if (act) {
			const u32 j = (u32) __popc(ng & ((1u << c) - 1u));
			const u64 key = sK[lane][j];
			const u32 grp = __match_any_sync(am, key);
			const int leader = __ffs(grp) - 1;
			u32 sl = 0, base = 0;
			if ((int) lane == leader) {
				sl = kd_insert(dk, key);
				if (sl != U32_MAX) {
					base = atomicAdd(dk.cnt + sl, (u32) __popc(grp)); atomicMin(dk.first + sl, p0 + j);
					if (ownc) {
						const u32 ow = ownc[(u32) t * 32u + c];
						const u32 prev = atomicCAS(own1 + sl, U32_MAX, ow);
						if (prev != U32_MAX && prev != ow) spans[sl] = 1;
					}
				}
			}
			sl = __shfl_sync(grp, sl, leader); base = __shfl_sync(grp, base, leader);
			sS[lane][j] = sl; sO[lane][j] = base + (u32) __popc(grp & lanemask_lt());
		}
	}
	__syncthreads();
	// phase 3: rows
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 rr = w + 8 * k;
		if (lane < (u32) __popc(sNG[rr])) {
			const u64 p = (u64) sP0[rr] + lane;
			__stcs(dk.slot_of + p, sS[rr][lane]); __stcs(dk.ord_of + p, sO[rr][lane]);
		}
	}
}
__global__ void k1_kmer_ends_kernel(const u8 *__restrict__ T, u64 N, CodeTable tab, int K, int b, KmerDict dk, const u32 *__restrict__ start, const u32 *__restrict__ rowlen, u64 m,
                                    u32 ncols, const u8 *__restrict__ ownc, u32 *__restrict__ own1, u8 *__restrict__ spans) {
	const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (i > m) return;
	const u64 p = i < m ? (u64) start[i] + rowlen[i] : N - 1;
	const u64 kmask = (K * b >= 64) ? ~0ull : ((1ull << (K * b)) - 1ull);
	u64 key = 0;
	for (int c = 0; c < K; c++) key = (key << b) | (p + c < N ? tab.code[T[p + c]] : (u8) 0);
	key &= kmask;
	const u32 sl = kd_insert(dk, key);
	if (sl == U32_MAX) return;
	dk.slot_of[p] = sl; dk.ord_of[p] = atomicAdd(dk.cnt + sl, 1u);
	atomicMin(dk.first + sl, (u32) p);
	if (ownc) { // '#' has column n, the sentinel n + 1
		const u32 ow = ownc[i < m ? ncols : ncols + 1u];
		const u32 prev = atomicCAS(own1 + sl, U32_MAX, ow);
		if (prev != U32_MAX && prev != ow) spans[sl] = 1;
	}
}

// ---- the same emission in TILE order (one GPU): a CTA takes 32 rows x the 32 columns of one bitmap word.
// In text order a warp's 32 positions belong to 32 different buckets: 32 sectors of info[], 32 of rank_of_slot[] and
// 3 x 32 of the lists per warp instruction -- ncu: 38 GB of DRAM traffic for 14 GB of algorithmic bytes, L2-bound.
// Here phase 1 reads the per-position inputs along the ROWS (a warp per row, consecutive text positions: coalesced)
// into shared memory; phase 2 turns the tile: a warp takes a COLUMN, its lanes the 32 rows -- nearly always one
// bucket, so the table look-ups of a warp hit one or two sectors and its list writes go to (nearly) consecutive
// slots, the arrival numbers of consecutive rows being (nearly) consecutive; phase 3 writes the ranks along the
// rows again.  Same bytes in the same places as the text-order kernel; '#' and the sentinel by k1_dict_emit_ends.
__global__ void __launch_bounds__(256) k1_dict_emit_tile_kernel(const u32 *__restrict__ slot_of, const u32 *__restrict__ ord_of, const uint4 *__restrict__ info, const u32 *__restrict__ rank_of_slot,
                                                                u64 N, u64 hnext, int nbits, u32 mid_min, u32 *__restrict__ R, SaOut so, CompactOut o,
                                                                const u32 *__restrict__ RK, const u32 *__restrict__ NG, const u32 *__restrict__ start, u64 m,
                                                                u32 tw0, u32 x0, u32 x1) {
	// [x0, x1): the columns this GPU owns (peer mode; one GPU: all of them), tw0 = x0 / 32
	__shared__ u32 sS[32][33], sO[32][33], sN[32][33]; // slot of p, arrival number of p, slot of p + hnext -> (phase 2) rank of p
	__shared__ u32 sNG[32], sP0[32];
	const u32 w = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const u32 ngroups = (u32) ((m + 31) / 32); // (one-dimensional grid, row group fastest: the blocks in flight work on few buckets)
	const u64 i0 = (u64) (blockIdx.x % ngroups) * 32, t = (u64) tw0 + blockIdx.x / ngroups;
	if (threadIdx.x < 32) {
		const u64 i = i0 + threadIdx.x;
		sNG[threadIdx.x] = i < m ? NG[t * m + i] : 0u;
		sP0[threadIdx.x] = i < m ? start[i] + RK[t * m + i] : 0u;
	}
	__syncthreads();
	// phase 1: rows
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 rr = w + 8 * k;
		const u32 cnt = (u32) __popc(sNG[rr]);
		if (lane < cnt) {
			const u64 p = (u64) sP0[rr] + lane, pn = p + hnext;
			sS[rr][lane] = __ldcs(slot_of + p);
			sO[rr][lane] = __ldcs(ord_of + p);
			sN[rr][lane] = pn < N ? slot_of[pn] : U32_MAX;
		}
	}
	__syncthreads();
	// phase 2: columns (lane = row)
	const u32 ng = sNG[lane], p0 = sP0[lane];
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 c = w + 8 * k;
		const u32 x = (u32) t * 32u + c;
		if (((ng >> c) & 1u) && x >= x0 && x < x1) {
			const u32 j = (u32) __popc(ng & ((1u << c) - 1u));
			const u32 p = p0 + j;
			const uint4 v = __ldg(info + sS[lane][j]);
			const u32 sn = sN[lane][j];
			sN[lane][j] = v.x; // the rank, for phase 3
			if (v.z == 1) so.put(v.x, p);
			else if (v.y != U32_MAX) { // (peer mode: U32_MAX = another GPU sorts this bucket)
				const u32 r2 = sn != U32_MAX ? __ldg(rank_of_slot + sn) : 0u;
				const u32 q = v.y + sO[lane][j];
				if (v.z < mid_min) { o.Slp[q] = p; o.Slr[q] = v.x; o.Sr2[q] = r2; }
				else if (v.z <= BSMALL) { o.Mlp[q] = p; o.Mlr[q] = v.x; o.Mr2[q] = r2; }
				else { o.Bkey[q] = ((u64) v.x << nbits) | r2; o.Bval[q] = p; }
			}
		}
	}
	__syncthreads();
	// phase 3: rows
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 rr = w + 8 * k;
		if (lane < (u32) __popc(sNG[rr])) {
			const u32 x = (u32) t * 32u + __fns(sNG[rr], 0, (int) lane + 1); // the column of the row's lane-th character in this word
			if (x >= x0 && x < x1) __stcs(R + (u64) sP0[rr] + lane, sN[rr][lane]);
		}
	}
}
// the '#' behind every row and the final sentinel (text-order code, m + 1 positions)
__global__ void k1_dict_emit_ends_kernel(const u32 *__restrict__ slot_of, const u32 *__restrict__ ord_of, const uint4 *__restrict__ info, const u32 *__restrict__ rank_of_slot,
                                         u64 N, u64 hnext, int nbits, u32 mid_min, u32 *__restrict__ R, SaOut so, CompactOut o,
                                         const u32 *__restrict__ start, const u32 *__restrict__ rowlen, u64 m) {
	const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (i > m) return;
	const u64 p = i < m ? (u64) start[i] + rowlen[i] : N - 1;
	const uint4 v = __ldg(info + slot_of[p]);
	R[p] = v.x;
	if (v.z == 1) { so.put(v.x, (u32) p); return; }
	if (v.y == U32_MAX) return;
	const u64 pn = p + hnext;
	const u32 r2 = pn < N ? __ldg(rank_of_slot + slot_of[pn]) : 0u;
	const u32 q = v.y + ord_of[p];
	if (v.z < mid_min) { o.Slp[q] = (u32) p; o.Slr[q] = v.x; o.Sr2[q] = r2; }
	else if (v.z <= BSMALL) { o.Mlp[q] = (u32) p; o.Mlr[q] = v.x; o.Mr2[q] = r2; }
	else { o.Bkey[q] = ((u64) v.x << nbits) | r2; o.Bval[q] = (u32) p; }
}

static u32 dict_capacity(u64 N) {
	u32 cap = 1u << 16;
	while ((u64) cap * 8 < N && cap < (1u << 30)) cap <<= 1;
	return cap;
}
static u64 dict_scratch_bytes(u64 N) {
	u64 cap = dict_capacity(N);
	return cap * (8 + 4 + 4 + 16 + 4 + 1 + 4) + (cap / 2) * (16 + 8 + 8 + 16 + 8 + 1) + (cap / SCAN_TILE + 2) * 24 + (cap / 2 / RS_TILE + 2) * RS_BINS * 12 + (1u << 16) + N + (1u << 12);
}

// Returns false (nothing written that matters) when the text has too many distinct k-mers for
// the table: the caller then sorts all suffixes as before.
// pv: null, or the column ranges of a peer-mode build (this GPU lists only the buckets it sorts; list_cap = room in its lists)
static bool k1_round0_dict(Pipeline &p, const CodeTable &tab, int K, int b, int key_bits0, u32 km_grid, u32 *slot_of, u32 *ord_of, u32 mid_min,
                           u64 hnext, int nbits, CompactOut o, u64 *nS, u64 *nM, u64 *nB, u32 *maxb, bool try_pure, const PeerView *pv = nullptr, u64 list_cap = 0) {
	Dev &d = p.dev;
	const u64 N = p.N;
	const u32 cap = dict_capacity(N), ndmax = cap / 2;
	u64 mk = d.mark();
	KmerDict dk;
	dk.key = d.alloc<u64>(cap); dk.cnt = d.alloc<u32>(cap); dk.first = d.alloc<u32>(cap);
	dk.mask = cap - 1; dk.slot_of = slot_of; dk.ord_of = ord_of;
	u32 *sc = reinterpret_cast<u32 *>(p.scalars + 48); // [0] distinct, [1] overflow, [2] nS, [3] nB, [4] maxb, [5] nM
	u32 *h_sc = reinterpret_cast<u32 *>(p.h_scalars + 48);
	dk.overflow = sc + 1;
	uint4 *info = d.alloc<uint4>(cap);
	u32 *pos = d.alloc<u32>(cap);
	u32 *rank_of_slot = pos; // the occupancy scan is dead once the distinct k-mers are listed
	RadixBuffers lb;
	lb.k[0] = d.alloc<u64>(ndmax); lb.k[1] = d.alloc<u64>(ndmax);
	lb.v[0] = d.alloc<u32>(ndmax); lb.v[1] = d.alloc<u32>(ndmax);
	lb.cur = 0;
	u32 *call = d.alloc<u32>(ndmax), *cbig = d.alloc<u32>(ndmax);
	RadixBuffers ob; // the distinct k-mers by smallest member (layout of the S and M lists)
	ob.k[0] = d.alloc<u64>(ndmax); ob.k[1] = d.alloc<u64>(ndmax);
	ob.v[0] = d.alloc<u32>(ndmax); ob.v[1] = d.alloc<u32>(ndmax);
	ob.cur = 0;
	const u32 gc = div_up(cap, 256);
	bool known_impure = false;
	if (try_pure && !p.force_arrays && p.opt.n_ignore == 0) {
		// "virtual" index: a light pass enters the k-mers and checks that every bucket lies in one
		// column; if so, the sorted list of distinct k-mers is all K2 needs (fcol.cu k2_from_dict)
		u32 *impure = reinterpret_cast<u32 *>(p.scalars + 12);
		u32 *h_impure = reinterpret_cast<u32 *>(p.h_scalars + 12);
		dk.slotcol = dk.cnt; dk.col = p.col; dk.impure = impure; dk.ncol1 = (u32) p.n + 1u;
		FBG_CUDA(cudaMemsetAsync(p.scalars + 48, 0, 4 * sizeof(u64), d.stream));
		FBG_CUDA(cudaMemsetAsync(impure, 0, sizeof(u64), d.stream));
		d.begin(K_KMER);
		FBG_CUDA(cudaMemsetAsync(dk.key, 0xFF, (u64) cap * sizeof(u64), d.stream));
		FBG_CUDA(cudaMemsetAsync(dk.slotcol, 0xFF, (u64) cap * sizeof(u32), d.stream));
		KmerShard sh0;
		sh0.bin_lo = 0; sh0.bin_hi = 0; sh0.hshift = 0; sh0.block_off = nullptr; sh0.hist = nullptr; sh0.block_cnt = nullptr;
		k1_kmer_kernel<5><<<km_grid, KM_THREADS, 0, d.stream>>>(p.T, N, tab, K, b, sh0, nullptr, nullptr, dk);
		FBG_LAUNCH_CHECK();
		d.end((double) N * (1.0 + 8.0 + 4.0));
		d.begin(K_DOUBLING_AUX);
		k1_dict_flag_kernel<<<gc, 256, 0, d.stream>>>(dk.key, cap, pos);
		FBG_LAUNCH_CHECK();
		d.end((double) cap * 12.0);
		device_scan<OpSumU32, true>(d, pos, pos, cap, sc);
		FBG_CUDA(cudaMemcpyAsync(h_sc, sc, 2 * sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaMemcpyAsync(h_impure, impure, sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		const u32 ndv = h_sc[0];
		if (h_sc[1] || ndv > ndmax) { d.release(mk); return false; }
		if (*h_impure == 0) {
			if (!p.dict_col) { p.dict_col = p.dev.palloc<u32>(ndmax); p.dict_lcp = p.dev.palloc<u32>((u64) ndmax + 1); }
			d.begin(K_DOUBLING_AUX);
			k1_dict_list_kernel<<<gc, 256, 0, d.stream>>>(dk.key, pos, cap, (u64 *) lb.k[0], lb.v[0]);
			FBG_LAUNCH_CHECK();
			d.end((double) cap * 12.0 + (double) ndv * 12.0);
			const int curv = radix_sort<u64, true>(d, lb, ndv, 0, key_bits0);
			d.begin(K_DOUBLING_AUX);
			k1_dict_virtual_kernel<<<div_up((u64) ndv + 1, 256), 256, 0, d.stream>>>((const u64 *) lb.k[curv], lb.v[curv], dk.slotcol, ndv, K, b, p.dict_col, p.dict_lcp);
			FBG_LAUNCH_CHECK();
			d.end((double) ndv * 28.0);
			FBG_CUDA(cudaStreamSynchronize(d.stream)); // the scratch is released below
			p.index_lazy = true; p.lazy_depth = (u64) K;
			p.index_virtual = true; p.dict_nd = ndv;
			*nS = 0; *nM = 0; *nB = 0; *maxb = 0;
			d.release(mk);
			return true;
		}
		known_impure = true; // the full insert below starts from an empty table
		lb.cur = 0;
	}
	FBG_CUDA(cudaMemsetAsync(p.scalars + 48, 0, 4 * sizeof(u64), d.stream));
	d.begin(K_KMER);
	FBG_CUDA(cudaMemsetAsync(dk.key, 0xFF, (u64) cap * sizeof(u64), d.stream));
	FBG_CUDA(cudaMemsetAsync(dk.cnt, 0, (u64) cap * sizeof(u32), d.stream));
	FBG_CUDA(cudaMemsetAsync(dk.first, 0xFF, (u64) cap * sizeof(u32), d.stream));
	KmerShard sh;
	sh.bin_lo = 0; sh.bin_hi = 0; sh.hshift = 0; sh.block_off = nullptr; sh.hist = nullptr; sh.block_cnt = nullptr;
	u8 *spans = nullptr, *ownc = nullptr;
	u32 *own1 = nullptr;
	bool spans_done = false;
	if (pv) { // peer mode: which GPU a bucket belongs to, which buckets span GPUs (k1_dict_mine_kernel below)
		spans = d.alloc<u8>(cap); own1 = d.alloc<u32>(cap); ownc = d.alloc<u8>(p.n + 2);
		FBG_CUDA(cudaMemsetAsync(spans, 0, cap, d.stream));
		FBG_CUDA(cudaMemsetAsync(own1, 0xFF, (u64) cap * sizeof(u32), d.stream));
		k1_owner_table_kernel<<<div_up(p.n + 2, 256), 256, 0, d.stream>>>((u32) p.n + 2u, *pv, ownc);
	}
	if (p.gaps != 0 && p.RK && p.NG && div_up(p.m, 32) * p.W <= 0x7FFFFFFFull && !getenv("FBG_KMER_TEXT_ORDER")) { // FBG_KMER_TEXT_ORDER=1: the text-order insert, for comparison
		k1_kmer_tile_kernel<<<(u32) (div_up(p.m, 32) * p.W), 256, 0, d.stream>>>(p.T, N, tab, K, b, dk, p.RK, p.NG, p.start, p.m, p.W, ownc, own1, spans);
		k1_kmer_ends_kernel<<<div_up(p.m + 1, 256), 256, 0, d.stream>>>(p.T, N, tab, K, b, dk, p.start, p.rowlen, p.m, (u32) p.n, ownc, own1, spans);
		spans_done = pv != nullptr;
	} else k1_kmer_kernel<4><<<km_grid, KM_THREADS, 0, d.stream>>>(p.T, N, tab, K, b, sh, nullptr, nullptr, dk);
	FBG_LAUNCH_CHECK();
	d.end((double) N * (1.0 + 8.0 + 8.0));
	d.begin(K_DOUBLING_AUX);
	k1_dict_flag_kernel<<<gc, 256, 0, d.stream>>>(dk.key, cap, pos);
	FBG_LAUNCH_CHECK();
	d.end((double) cap * 12.0);
	device_scan<OpSumU32, true>(d, pos, pos, cap, sc);
	FBG_CUDA(cudaMemcpyAsync(h_sc, sc, 2 * sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	const u32 nd = h_sc[0];
	if (h_sc[1] || nd > ndmax) { d.release(mk); return false; }
	d.begin(K_DOUBLING_AUX);
	k1_dict_list_kernel<<<gc, 256, 0, d.stream>>>(dk.key, pos, cap, (u64 *) lb.k[0], lb.v[0]);
	FBG_LAUNCH_CHECK();
	d.end((double) cap * 12.0 + (double) nd * 12.0);
	const int cur = radix_sort<u64, true>(d, lb, nd, 0, key_bits0);
	const u32 *lv = lb.v[cur];
	const u32 gn = div_up(nd, 256);
	u8 *mine = nullptr;
	if (pv) {
		mine = d.alloc<u8>(ndmax);
		d.begin(K_DOUBLING_AUX);
		if (!spans_done) { // (the text-order insert does not look at owners)
			k1_dict_own1_kernel<<<gn, 256, 0, d.stream>>>(lv, dk.first, p.col, ownc, nd, own1);
			k1_dict_spans_kernel<<<div_up(N, 256), 256, 0, d.stream>>>(slot_of, p.col, ownc, own1, N, spans);
		}
		k1_dict_mine_kernel<<<gn, 256, 0, d.stream>>>(lv, own1, spans, nd, pv->me, mine, sc + 6);
		FBG_LAUNCH_CHECK();
		d.end((double) N * 8.0 + (double) nd * 13.0);
		d.note_launches(K_DOUBLING_AUX, 1);
	}
	d.begin(K_DOUBLING_AUX);
	k1_dict_counts_kernel<<<gn, 256, 0, d.stream>>>(lv, dk.cnt, mine, nd, call, cbig);
	FBG_LAUNCH_CHECK();
	d.end((double) nd * 16.0);
	device_scan<OpSumU32, true>(d, call, call, nd);
	device_scan<OpSumU32, true>(d, cbig, cbig, nd, sc + 3);
	d.begin(K_DOUBLING_AUX);
	k1_dict_assign_kernel<<<gn, 256, 0, d.stream>>>(lv, dk.cnt, dk.first, call, cbig, mine, nd, 0u, mid_min, info, rank_of_slot, sc + 4);
	FBG_LAUNCH_CHECK();
	d.end((double) nd * 40.0);
	if (try_pure && !known_impure) {
		// sufficient index (above k1_suffix_array): if no bucket spans two columns this is all of K1
		u32 *impure = reinterpret_cast<u32 *>(p.scalars + 12);
		u32 *h_impure = reinterpret_cast<u32 *>(p.h_scalars + 12);
		FBG_CUDA(cudaMemsetAsync(impure, 0, sizeof(u64), d.stream));
		d.begin(K_DOUBLING_AUX);
		k1_dict_emit_pure_kernel<<<div_up(N, 256), 256, 0, d.stream>>>(slot_of, ord_of, info, p.col, N, (u32) K, p.ISA, p.SA, p.PLCP, impure);
		FBG_LAUNCH_CHECK();
		d.end((double) N * (8.0 + 4.0 + 4.0 + 4.0 + 4.0) + (double) nd * 16.0);
		FBG_CUDA(cudaMemcpyAsync(h_impure, impure, sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		if (*h_impure == 0) {
			p.index_lazy = true; p.lazy_depth = (u64) K;
			*nS = 0; *nM = 0; *nB = 0; *maxb = 0;
			d.release(mk);
			return true;
		}
	}
	// offsets of the S and M buckets in their lists, in the order of their smallest members
	d.begin(K_DOUBLING_AUX);
	k1_dict_firstkeys_kernel<<<gn, 256, 0, d.stream>>>(lv, dk.first, nd, (u64 *) ob.k[0], ob.v[0]);
	FBG_LAUNCH_CHECK();
	d.end((double) nd * 20.0);
	const int oc = radix_sort<u64, true>(d, ob, nd, 0, nbits);
	const u32 *ov = ob.v[oc];
	u32 *zs = call, *zm = cbig; // both scans were consumed by k1_dict_assign_kernel
	d.begin(K_DOUBLING_AUX);
	k1_dict_sizes_kernel<<<gn, 256, 0, d.stream>>>(ov, lv, dk.cnt, mine, nd, mid_min, zs, zm);
	FBG_LAUNCH_CHECK();
	d.end((double) nd * 20.0);
	device_scan<OpSumU32, true>(d, zs, zs, nd, sc + 2);
	device_scan<OpSumU32, true>(d, zm, zm, nd, sc + 5);
	if (pv) {
		// the lists were sized for this GPU's share of the columns: check before anything is written into them
		FBG_CUDA(cudaMemcpyAsync(h_sc, sc, 7 * sizeof(u32), cudaMemcpyDeviceToHost, d.stream)); // [6]: buckets that span GPUs
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		// (all GPUs decide alike: the build changes form if any of them is short of room)
		p.h_scalars[63] = ((u64) h_sc[2] + h_sc[5] + h_sc[3] > list_cap) ? 1 : 0;
		FBG_CUDA(cudaMemcpyAsync(p.scalars + 63, p.h_scalars + 63, sizeof(u64), cudaMemcpyHostToDevice, d.stream));
		p.dist.allreduce_sum_u64(p.scalars + 63, 1, d.stream);
		FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 63, p.scalars + 63, sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		if (p.h_scalars[63]) { d.release(mk); return false; }
	}
	d.begin(K_DOUBLING_AUX);
	k1_dict_place_kernel<<<gn, 256, 0, d.stream>>>(ov, lv, nd, mid_min, zs, zm, info);
	if (pv) {
		FBG_CUDA(cudaMemsetAsync(p.ISA, 0, N * sizeof(u32), d.stream)); // ranks outside this GPU's columns stay 0: the inverse is assembled by max at the end
		const u32 x0 = pv->xb[pv->me], x1 = pv->xb[pv->me + 1];
		const int last = pv->me + 1 == pv->world ? 1 : 0;
		u32 *own_a = d.alloc<u32>(p.m), *own_b = d.alloc<u32>(p.m);
		k1_own_ranges_kernel<<<div_up(p.m, 256), 256, 0, d.stream>>>(p.RK, p.NG, p.start, p.rowlen, p.m, p.n, x0, x1, last, own_a, own_b);
		const SaOut sao{p.SA, p.have_bwt ? p.BWT : nullptr, p.T};
		if (p.gaps != 0 && p.RK && p.NG && !getenv("FBG_EMIT_TEXT_ORDER")) { // tile order over this GPU's columns (a gapless text has no RK / NG tables)
			const u32 xe = x1 < (u32) p.n ? x1 : (u32) p.n;
			if (xe > x0) {
				const u64 gt = div_up(p.m, 32) * (((u64) xe + 31) / 32 - x0 / 32);
				if (gt > 0x7FFFFFFFull) throw unsupported_error("round 0: more than 2^31 tiles of 32 rows x 32 columns");
				k1_dict_emit_tile_kernel<<<(u32) gt, 256, 0, d.stream>>>(slot_of, ord_of, info, rank_of_slot, N, hnext, nbits, mid_min, p.ISA, sao, o, p.RK, p.NG, p.start, p.m, x0 / 32, x0, xe);
			}
			if (last) k1_dict_emit_ends_kernel<<<div_up(p.m + 1, 256), 256, 0, d.stream>>>(slot_of, ord_of, info, rank_of_slot, N, hnext, nbits, mid_min, p.ISA, sao, o, p.start, p.rowlen, p.m);
		} else {
			const dim3 g1((u32) div_up((u64) (x1 - x0) + 2, 256), (u32) p.m); // a row has at most x1 - x0 characters in these columns (+ '#', + the sentinel)
			k1_dict_emit_kernel<1><<<g1, 256, 0, d.stream>>>(slot_of, ord_of, info, rank_of_slot, N, hnext, nbits, mid_min, p.ISA, sao, o, pv->notmine, spans, own_a, own_b, sc + 6);
		}
		// (nothing to do -- the usual case -- costs 2 ms of empty blocks at 5e8 positions: not launched then)
		if (h_sc[6]) k1_dict_emit_kernel<2><<<div_up(N, 256), 256, 0, d.stream>>>(slot_of, ord_of, info, rank_of_slot, N, hnext, nbits, mid_min, p.ISA, sao, o, pv->notmine, spans, own_a, own_b, sc + 6);
		d.note_launches(K_DOUBLING_AUX, 2);
	} else if (p.gaps != 0 && p.RK && p.NG && div_up(p.m, 32) * p.W <= 0x7FFFFFFFull && !getenv("FBG_EMIT_TEXT_ORDER")) {
		// (a gapless text has no RK / NG tables; FBG_EMIT_TEXT_ORDER=1: the text-order kernel, for comparison)
		const SaOut sao{p.SA, p.have_bwt ? p.BWT : nullptr, p.T};
		const u32 gt = (u32) (div_up(p.m, 32) * p.W);
		k1_dict_emit_tile_kernel<<<gt, 256, 0, d.stream>>>(slot_of, ord_of, info, rank_of_slot, N, hnext, nbits, mid_min, p.ISA, sao, o, p.RK, p.NG, p.start, p.m, 0u, 0u, (u32) p.n);
		k1_dict_emit_ends_kernel<<<div_up(p.m + 1, 256), 256, 0, d.stream>>>(slot_of, ord_of, info, rank_of_slot, N, hnext, nbits, mid_min, p.ISA, sao, o, p.start, p.rowlen, p.m);
		d.note_launches(K_DOUBLING_AUX, 1);
	} else k1_dict_emit_kernel<0><<<div_up(N, 256), 256, 0, d.stream>>>(slot_of, ord_of, info, rank_of_slot, N, hnext, nbits, mid_min, p.ISA, SaOut{p.SA, p.have_bwt ? p.BWT : nullptr, p.T}, o, nullptr, nullptr, nullptr, nullptr, nullptr);
	FBG_LAUNCH_CHECK();
	d.end((double) N * (8.0 + 4.0 + 4.0 + 12.0) + (double) nd * 40.0);
	d.note_launches(K_DOUBLING_AUX, 1);
	FBG_CUDA(cudaMemcpyAsync(h_sc, sc, 6 * sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	*nS = h_sc[2]; *nB = h_sc[3]; *maxb = h_sc[4]; *nM = h_sc[5];
	d.release(mk);
	return true;
}

void k1_suffix_array(Pipeline &p) {
	Dev &d = p.dev;
	Dist &dist = p.dist;
	const bool sharded = dist.world > 1;
	const u64 N = p.N;
	if (!p.ISA) p.ISA = p.dev.palloc<u32>(N + 1); // (a full index built after a sufficient one re-uses the arrays)
	if (!p.SA) p.SA = p.dev.palloc<u32>(N + 1);
	if (!p.PLCP) p.PLCP = p.dev.palloc<u32>(N + 1);
	// BWT next to the SA (SaOut) for the LCP stage of a full index (the sufficient index finds its LCPs differently);
	// a multi-GPU build merges the bytes like the SA entries: every entry is written by the GPU(s) that sorted its bucket
	p.have_bwt = !getenv("FBG_NO_BWT");
	if (p.have_bwt && !p.BWT) p.BWT = p.dev.palloc<u8>(N + 64);
	if (p.have_bwt && sharded) FBG_CUDA(cudaMemsetAsync(p.BWT, 0, N, d.stream));
	p.index_lazy = false; p.lazy_depth = 0; p.index_virtual = false; p.dict_nd = 0;
	// the sufficient index (above): gapless rows, one GPU, unless the full index was asked for
	const bool lazy_ok = !sharded && p.gaps == 0 && !p.want_full_index && !getenv("FBG_FULL_INDEX");
	d.ensure((sharded ? 1u : 0u) * (u64) (4u << KM_HBITS) + (1u << 20));
	u64 mk = d.mark();

	// alphabet: dense order-preserving codes (sentinel 0 is always present)
	u32 *present = d.alloc<u32>(256);
	FBG_CUDA(cudaMemsetAsync(present, 0, 256 * sizeof(u32), d.stream));
	d.begin(K_PACK); // alphabet scan: accounted with the text preparation, K_KMER is the k-mer pass alone
	k1_presence_kernel<<<NUM_SMS * 4, 256, 0, d.stream>>>(p.T, N, present);
	FBG_LAUNCH_CHECK();
	d.end((double) N);
	u32 hpres[256];
	FBG_CUDA(cudaMemcpyAsync(hpres, present, sizeof hpres, cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	CodeTable tab;
	int sigma = 0;
	for (int c = 0; c < 256; c++) { tab.code[c] = (u8) sigma; if (hpres[c]) sigma++; }
	// tab.code[c] = number of present bytes < c: order preserving, sentinel -> 0
	int b = 1;
	while ((1 << b) < sigma) b++;
	int K = 64 / b;
	if (K > 32) K = 32;
	if (K * b == 64) K--; // keep the all-ones key free: it marks the empty slots of the k-mer dictionary
	const int key_bits0 = K * b;
	const u32 km_grid = div_up(N, KM_TILE);

	// ---- which suffixes this GPU sorts: all of them, or one k-mer key range (dist.cuh)
	KmerShard sh;
	sh.bin_lo = 0; sh.bin_hi = 1u << KM_HBITS; sh.hshift = key_bits0 > KM_HBITS ? key_bits0 - KM_HBITS : 0;
	sh.block_off = nullptr; sh.hist = nullptr; sh.block_cnt = nullptr;
	u64 nloc = N;      // suffixes sorted here
	u32 rank_off = 0;  // global rank of the first of them
	// Peer mode (several GPUs, the preferred form): every GPU builds the k-mer dictionary of round 0, the tied buckets
	// are dealt out by COLUMN range, every GPU keeps the ranks of its own columns up to date and reads the ranks of
	// other columns straight from their owner's memory over NVLink (CUDA IPC).  No rank array is ever all-reduced.
	// The older form (k-mer key ranges, rank replicas merged by all-reduce(max) every round) remains for texts
	// the dictionary does not take (p.no_peer: set for one build when the peer attempt had to be given up).
	bool peer = sharded && key_bits0 < 64 && !p.no_peer && !getenv("FBG_NO_KMER_DICT") && !getenv("FBG_NO_PEER");
	p.no_peer = false;
	if (sharded && !peer) {
		const u32 nbins = 1u << KM_HBITS;
		u64 mk2 = d.mark();
		u32 *hist = d.alloc<u32>(nbins);
		FBG_CUDA(cudaMemsetAsync(hist, 0, nbins * sizeof(u32), d.stream));
		sh.hist = hist;
		d.begin(K_KMER);
		k1_kmer_kernel<1><<<km_grid, KM_THREADS, 0, d.stream>>>(p.T, N, tab, K, b, sh, nullptr, nullptr, KmerDict());
		FBG_LAUNCH_CHECK();
		d.end((double) N);
		std::vector<u32> hh(nbins), bounds(dist.world + 1);
		FBG_CUDA(cudaMemcpyAsync(hh.data(), hist, nbins * sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		// the text is replicated, so every rank derives the same plan without talking
		plan_key_ranges(hh.data(), nbins, dist.world, bounds.data());
		sh.bin_lo = bounds[dist.rank]; sh.bin_hi = bounds[dist.rank + 1];
		u64 before = 0, mine = 0;
		for (u32 q = 0; q < sh.bin_hi; q++) { if (q < sh.bin_lo) before += hh[q]; else mine += hh[q]; }
		nloc = mine; rank_off = (u32) before;
		d.release(mk2);
	}
	u64 cap = nloc ? nloc : 1;
	if (peer) { // this GPU's share of the columns, with room for uneven columns and for the buckets every GPU keeps
		cap = N / (u64) dist.world + N / (2 * (u64) dist.world) + (8u << 20);
		if (cap > N) cap = N;
	}
	const u64 capN = peer ? N : cap; // the dictionary's per-position arrays cover the whole text
	// scratch of this phase: 98 B per suffix sorted here (sort ping-pong 24, heads 16, classes 2,
	// S lists 24, M lists 24, sorter outputs 8) + radix count matrices + (key-range form) the N-word bucket directory
	d.release(mk);
	d.ensure(cap * 100 + (capN - cap) * 8 + (cap / RS_TILE + 2) * RS_BINS * 12 + (u64) km_grid * 8 + (sharded && !peer ? N * 4 + (N / SCAN_TILE + 2) * 24 : dict_scratch_bytes(N)) + (peer ? N / 8 + 4096 : 0) + (16u << 20));
	mk = d.mark();

	RadixBuffers rb;
	rb.k[0] = d.alloc<u64>(cap); rb.k[1] = d.alloc<u64>(cap);
	rb.v[0] = d.alloc<u32>(cap); rb.v[1] = d.alloc<u32>(cap);
	rb.cur = 0;
	u64 *HH = d.alloc<u64>(cap);
	u32 *HR = d.alloc<u32>(capN), *newr = d.alloc<u32>(capN);
	u8 *clsB = d.alloc<u8>(cap + 16), *clsS = d.alloc<u8>(cap + 16);
	u32 *Slp[2], *Slr[2], *Sr2[2], *Mlp[2], *Mlr[2], *Mr2[2];
	for (int q = 0; q < 2; q++) { Slp[q] = d.alloc<u32>(cap); Slr[q] = d.alloc<u32>(cap); Sr2[q] = d.alloc<u32>(cap); }
	for (int q = 0; q < 2; q++) { Mlp[q] = d.alloc<u32>(cap); Mlr[q] = d.alloc<u32>(cap); Mr2[q] = d.alloc<u32>(cap); }
	// outputs of the two bucket-local sorters: S list in [0, nS), M list in [nS, nS + nM)
	u32 *lp_out = d.alloc<u32>(cap), *lr_out = d.alloc<u32>(cap);
	u32 *sc32 = reinterpret_cast<u32 *>(p.scalars + 56);  // [0] maxb, [1] err
	u64 *cnt = p.scalars + 57;                           // fill counts of the next S, M, B lists
	u64 *ctot = p.scalars + 60;                          // scratch word of the compactions
	u64 *gtot = p.scalars + 61;                          // suffixes still tied, over all GPUs
	u64 *h_sc = p.h_scalars + 56;                        // mirror of scalars[56..61]

	int nbits = 1;
	while ((1ull << nbits) < N) nbits++;
	u64 nS = 0, nM = 0, nB = nloc; // round 0: every suffix goes through the device-wide sort ...
	u32 maxb = 0;
	int sc = 0;           // current S / M lists
	// ... unless the k-mer dictionary takes it: no sort, ranks written in text order
	bool by_dict = false;
	PeerView pv, *pv_dev = nullptr;
	memset(&pv, 0, sizeof pv);
	if (peer) {
		pv.me = dist.rank; pv.world = dist.world;
		for (int g = 0; g < dist.world; g++) { u64 a, bq; plan_columns(p.n, dist.world, g, &a, &bq); pv.xb[g] = (u32) a; pv.xb[g + 1] = (u32) bq; }
		void *rp[MAX_PEERS];
		dist.open_peers(d.pers, p.ISA, rp, d.stream);
		for (int g = 0; g < dist.world; g++) pv.R[g] = static_cast<u32 *>(rp[g]);
		u32 *nmb = d.alloc<u32>(N / 32 + 2);
		k1_notmine_kernel<<<div_up(N, 256), 256, 0, d.stream>>>(p.col, N, pv.xb[pv.me], pv.xb[pv.me + 1], pv.me + 1 == pv.world ? 1 : 0, nmb);
		FBG_LAUNCH_CHECK();
		pv.notmine = nmb;
		pv_dev = d.alloc<PeerView>(1);
		FBG_CUDA(cudaMemcpyAsync(pv_dev, &pv, sizeof pv, cudaMemcpyHostToDevice, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream)); // pv lives on this function's stack
		FBG_CUDA(cudaMemsetAsync(p.SA, 0, N * sizeof(u32), d.stream)); // every GPU writes the SA entries of the buckets it sorts; merged by max at the end
	}
	u32 mid_want = MID_MIN; // FBG_MID_MIN: test hook (64 ..: small inputs through the M path; above BSMALL: no M list)
	if (const char *e = getenv("FBG_MID_MIN")) { long v = atol(e); mid_want = v > (long) BSMALL ? BSMALL + 1 : (v < 64 ? 64u : (u32) v); }
	if ((!sharded || peer) && key_bits0 < 64 && !getenv("FBG_NO_KMER_DICT")) {
		CompactOut o;
		o.Slp = Slp[1]; o.Slr = Slr[1]; o.Sr2 = Sr2[1];
		o.Mlp = Mlp[1]; o.Mlr = Mlr[1]; o.Mr2 = Mr2[1];
		o.Bkey = (u64 *) rb.k[1]; o.Bval = rb.v[1];
		by_dict = k1_round0_dict(p, tab, K, b, key_bits0, km_grid, HR, newr, mid_want, (u64) K, nbits, o, &nS, &nM, &nB, &maxb,
		                         lazy_ok && (u64) K <= LAZY_MAX_DEPTH, peer ? &pv : nullptr, cap);
	}
	if (peer && !by_dict) {
		// the dictionary did not take this text (or a GPU's share outgrew its lists) -- on every GPU alike: key-range form
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		d.release(mk);
		p.no_peer = true;
		k1_suffix_array(p);
		return;
	}
	// the M list (buckets of MID_MIN .. BSMALL suffixes streamed one per CTA) needs the text-ordered layout the
	// dictionary's round 0 gives it; builds whose round 0 sorts all suffixes keep everything up to BSMALL in the S list
	const u32 mid_min = by_dict ? mid_want : BSMALL + 1;
	if (!by_dict) d.begin(K_KMER);
	if (by_dict) {
	} else if (!sharded) {
		k1_kmer_kernel<0><<<km_grid, KM_THREADS, 0, d.stream>>>(p.T, N, tab, K, b, sh, (u64 *) rb.k[0], rb.v[0], KmerDict());
	} else {
		u64 mk2 = d.mark();
		u64 *bc = d.alloc<u64>(km_grid);
		sh.block_cnt = bc; sh.block_off = bc;
		k1_kmer_kernel<2><<<km_grid, KM_THREADS, 0, d.stream>>>(p.T, N, tab, K, b, sh, nullptr, nullptr, KmerDict());
		scan_partials_kernel<OpSumU64><<<1, 1024, 0, d.stream>>>(bc, km_grid, nullptr);
		k1_kmer_kernel<3><<<km_grid, KM_THREADS, 0, d.stream>>>(p.T, N, tab, K, b, sh, (u64 *) rb.k[0], rb.v[0], KmerDict());
		FBG_CUDA(cudaMemsetAsync(p.ISA, 0, N * sizeof(u32), d.stream)); // replicas are merged by max
		FBG_CUDA(cudaMemsetAsync(p.SA, 0, N * sizeof(u32), d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream)); // bc is released below
		d.release(mk2);
		d.note_launches(K_KMER, 2);
	}
	FBG_LAUNCH_CHECK();
	if (!by_dict) d.end((double) N * (1.0 + 8.0 + 4.0));

	const SaOut sao{p.SA, p.have_bwt ? p.BWT : nullptr, p.T};
	u64 h = (u64) K;      // depth the current ranks are exact to after the round being run
	int hi_shift = 64;    // round 0: a single level of buckets
	int key_bits = key_bits0;
	p.sa_rounds = 0;
	const bool debug = getenv("FBG_DEBUG") != nullptr;
	if (debug) fprintf(stderr, "[fbg] rank %d K1 round 0 %s\n", dist.rank, by_dict ? "through the k-mer dictionary" : "by sorting all suffixes");
	p.kmer_dict = by_dict ? 1 : 0;
	if (by_dict) {
		// the state a sorted round 0 ends in
		p.sa_rounds = 1;
		if (nB) rb.cur ^= 1;
		sc ^= 1;
		hi_shift = nbits;
		key_bits = 2 * nbits;
		h *= 2;
	}
	for (;;) {
		if (by_dict && nS + nM + nB == 0) break; // every k-mer was unique
		if (lazy_ok && p.sa_rounds >= 1 && h / 2 <= LAZY_MAX_DEPTH && nS + nM + nB) {
			// ranks are exact to depth h / 2: are all buckets still tied column-pure?
			u32 *impure = reinterpret_cast<u32 *>(p.scalars + 12);
			u32 *h_impure = reinterpret_cast<u32 *>(p.h_scalars + 12);
			const u64 *bk = (const u64 *) rb.k[rb.cur];
			const u32 *bv = rb.v[rb.cur];
			FBG_CUDA(cudaMemsetAsync(impure, 0, sizeof(u64), d.stream));
			d.begin(K_DOUBLING_AUX);
			if (nS) k1_pure_check_kernel<false><<<div_up(nS, 256), 256, 0, d.stream>>>(Slp[sc], Slr[sc], nullptr, nbits, nS, p.col, impure);
			if (nM) k1_pure_check_kernel<false><<<div_up(nM, 256), 256, 0, d.stream>>>(Mlp[sc], Mlr[sc], nullptr, nbits, nM, p.col, impure);
			if (nB) k1_pure_check_kernel<true><<<div_up(nB, 256), 256, 0, d.stream>>>(bv, nullptr, bk, nbits, nB, p.col, impure);
			FBG_LAUNCH_CHECK();
			d.end((double) (nS + nM) * 12.0 + (double) nB * 16.0);
			FBG_CUDA(cudaMemcpyAsync(h_impure, impure, sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
			FBG_CUDA(cudaStreamSynchronize(d.stream));
			if (debug) fprintf(stderr, "[fbg] K1 depth %llu: tied buckets %s\n", (unsigned long long) (h / 2), *h_impure ? "span several columns" : "are all column-pure: stop");
			if (*h_impure == 0) {
				const u32 depth = (u32) (h / 2);
				FBG_CUDA(cudaMemsetAsync(p.PLCP, 0xFF, N * sizeof(u32), d.stream)); // PLCP_TODO everywhere
				for (int li = 0; li < 2; li++) { // S list, then M list (same records)
					const u64 nl = li ? nM : nS;
					const u32 *llp = li ? Mlp[sc] : Slp[sc], *llr = li ? Mlr[sc] : Slr[sc];
					if (!nl) continue;
					d.begin(K_DOUBLING_AUX);
					k1_pure_heads_kernel<false><<<div_up(nl, 256), 256, 0, d.stream>>>(llr, nullptr, nbits, nl, lp_out);
					FBG_LAUNCH_CHECK();
					d.end((double) nl * 8.0);
					device_scan<OpMaxU32, false>(d, lp_out, lp_out, nl);
					d.begin(K_DOUBLING_AUX);
					k1_pure_final_kernel<false><<<div_up(nl, 256), 256, 0, d.stream>>>(llp, llr, nullptr, nbits, nl, lp_out, depth, p.ISA, p.SA, p.PLCP);
					FBG_LAUNCH_CHECK();
					d.end((double) nl * 24.0);
				}
				if (nB) {
					d.begin(K_DOUBLING_AUX);
					k1_pure_heads_kernel<true><<<div_up(nB, 256), 256, 0, d.stream>>>(nullptr, bk, nbits, nB, HR);
					FBG_LAUNCH_CHECK();
					d.end((double) nB * 12.0);
					device_scan<OpMaxU32, false>(d, HR, HR, nB);
					d.begin(K_DOUBLING_AUX);
					k1_pure_final_kernel<true><<<div_up(nB, 256), 256, 0, d.stream>>>(bv, nullptr, bk, nbits, nB, HR, depth, p.ISA, p.SA, p.PLCP);
					FBG_LAUNCH_CHECK();
					d.end((double) nB * 28.0);
				}
				p.index_lazy = true; p.lazy_depth = depth;
				break;
			}
		}
		if (debug) fprintf(stderr, "[fbg] rank %d K1 round %d: depth %llu, S list %llu (largest bucket %u), M list %llu, B list %llu\n", dist.rank, p.sa_rounds,
		                   (unsigned long long) (p.sa_rounds ? h / 2 : 0), (unsigned long long) nS, maxb, (unsigned long long) nM, (unsigned long long) nB);
		FBG_CUDA(cudaMemsetAsync(p.scalars + 56, 0, 6 * sizeof(u64), d.stream));
		// ---- S path: windows of whole small buckets sorted in shared memory
		if (nS) {
			d.begin(K_TILE_SORT);
			if (2 * maxb <= 2048) launch_tile_sort<2048>(d, Slp[sc], Slr[sc], Sr2[sc], nS, maxb, p.ISA, sao, lp_out, lr_out, clsS, sc32);
			else if (2 * maxb <= 4096) launch_tile_sort<4096>(d, Slp[sc], Slr[sc], Sr2[sc], nS, maxb, p.ISA, sao, lp_out, lr_out, clsS, sc32);
			else if (2 * maxb <= 8192) launch_tile_sort<8192>(d, Slp[sc], Slr[sc], Sr2[sc], nS, maxb, p.ISA, sao, lp_out, lr_out, clsS, sc32);
			else launch_tile_sort<16384>(d, Slp[sc], Slr[sc], Sr2[sc], nS, maxb, p.ISA, sao, lp_out, lr_out, clsS, sc32);
			d.end((double) nS * (12.0 + 4.0 + 9.0));
		}
		// ---- M path: one CTA streams one bucket
		if (nM) {
			d.begin(K_TILE_SORT);
			launch_bucket_sort(d, Mlp[sc], Mlr[sc], Mr2[sc], nM, mid_min, p.ISA, sao, lp_out + nS, lr_out + nS, clsS + nS, sc32);
			d.end((double) nM * (12.0 + 4.0 + 9.0));
		}
		// ---- B path
		const u32 *sv = nullptr;
		if (nB) {
			int cur = radix_sort<u64, true>(d, rb, nB, 0, key_bits);
			const u64 *sk = (const u64 *) rb.k[cur];
			sv = rb.v[cur];
			u32 g = div_up(nB, 256);
			d.begin(K_DOUBLING_AUX);
			k1_heads_kernel<<<g, 256, 0, d.stream>>>(sk, nB, hi_shift, HH, HR);
			FBG_LAUNCH_CHECK();
			d.end((double) nB * 20.0);
			device_scan<OpMax2x32, false>(d, HH, HH, nB);
			device_scan<OpMaxU32, false>(d, HR, HR, nB);
			d.begin(K_DOUBLING_AUX);
			k1_update_kernel<<<g, 256, 0, d.stream>>>(sk, sv, HH, HR, nB, hi_shift, rank_off, mid_min, p.ISA, sao, newr, clsB, sc32);
			FBG_LAUNCH_CHECK();
			d.end((double) nB * (8.0 + 4.0 + 8.0 + 4.0 + 4.0 + 4.0 + 1.0));
		}
		// ---- key-range form: merge the rank replicas (every GPU raised only the ranks of its own suffixes);
		//      peer mode: only a barrier -- every GPU's rank updates of this round are done before anyone looks ranks up
		if (sharded) {
			d.begin(K_COMM);
			if (peer) dist.allreduce_sum_u64(p.scalars + 62, 1, d.stream);
			else dist.allreduce_max_u32(p.ISA, N, d.stream);
			d.end(peer ? 16.0 : (double) N * 8.0);
		}
		// ---- next lists (all rank updates of this round are enqueued before these look-ups)
		u64 hnext = h;
		CompactOut o;
		o.Slp = Slp[sc ^ 1]; o.Slr = Slr[sc ^ 1]; o.Sr2 = Sr2[sc ^ 1];
		o.Mlp = Mlp[sc ^ 1]; o.Mlr = Mlr[sc ^ 1]; o.Mr2 = Mr2[sc ^ 1];
		o.Bkey = (u64 *) rb.k[rb.cur ^ 1]; o.Bval = rb.v[rb.cur ^ 1];
		const bool ordered = p.sa_rounds == 0; // round 0 by sort: S list in text order (k1_bucket_size_kernel); no M list in such a build
		if (ordered) {
			u64 mk2 = d.mark();
			u32 *A = sharded ? d.alloc<u32>(N) : lp_out; // lp_out is free in round 0 (no S path yet) and holds N words on one GPU
			FBG_CUDA(cudaMemsetAsync(A, 0, N * sizeof(u32), d.stream));
			if (nB) {
				u32 g = div_up(nB, 256);
				d.begin(K_DOUBLING_AUX);
				k1_bucket_size_kernel<<<g, 256, 0, d.stream>>>(HH, HR, sv, clsB, nB, A);
				FBG_LAUNCH_CHECK();
				d.end((double) nB * 9.0 + (double) N * 4.0);
			}
			device_scan<OpSumU32, true>(d, A, A, N, reinterpret_cast<u32 *>(cnt)); // cnt[0] = next nS (the high half of the word is zero)
			if (nB) {
				u32 g = div_up(nB, 256);
				d.begin(K_DOUBLING_AUX);
				k1_ordered_write_kernel<<<g, 256, 0, d.stream>>>(HH, sv, newr, clsB, nB, A, p.ISA, hnext, N, o.Slp, o.Slr, o.Sr2);
				FBG_LAUNCH_CHECK();
				d.end((double) nB * (8.0 + 4.0 + 4.0 + 1.0 + 4.0 + 4.0 + 12.0));
			}
			if (sharded) FBG_CUDA(cudaStreamSynchronize(d.stream)); // A is released below
			d.release(mk2);
			compact_lists(d, clsB, sv, newr, nB, CLS_NONE, TGT_S, nullptr, CLS_BIG, TGT_B, cnt + 2, ctot, p.ISA, hnext, N, nbits, o);
		} else {
			// the sorted S and M lists lie back to back in the sorter outputs: one pass moves both
			const PeerView *pvp = peer && !getenv("FBG_PEER_LOCALREAD") ? pv_dev : nullptr; // (the variable: timing experiments only, wrong ranks)
			compact_lists(d, clsS, lp_out, lr_out, nS + nM, CLS_SMALL, TGT_S, cnt, CLS_MID, TGT_M, cnt + 1, ctot, p.ISA, hnext, N, nbits, o, p.col, pvp);
			compact_lists(d, clsB, sv, newr, nB, CLS_SMALL, TGT_S, cnt, CLS_BIG, TGT_B, cnt + 2, ctot, p.ISA, hnext, N, nbits, o, p.col, pvp);
			if (mid_min <= BSMALL) compact_lists(d, clsB, sv, newr, nB, CLS_MID, TGT_M, cnt + 1, CLS_NONE, TGT_B, nullptr, ctot, p.ISA, hnext, N, nbits, o, p.col, pvp);
		}
		k1_active_total_kernel<<<1, 1, 0, d.stream>>>(cnt, gtot);
		if (sharded) dist.allreduce_sum_u64(gtot, 1, d.stream);
		FBG_CUDA(cudaMemcpyAsync(h_sc, p.scalars + 56, 6 * sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		p.sa_rounds++;
		const u32 *h32 = reinterpret_cast<const u32 *>(h_sc);
		if (h32[1]) throw cuda_error("bucket-local suffix sort met a bucket larger than its tile");
		maxb = h32[0];
		nS = h_sc[1]; nM = h_sc[2]; nB = h_sc[3];
		if (nB) rb.cur ^= 1; // the compaction wrote the next B list into the other ping-pong half
		sc ^= 1;
		if (h_sc[5] == 0) break; // nothing tied on any GPU
		if (h >= N) throw cuda_error("prefix doubling did not converge");
		if (getenv("FBG_TEST_FAIL_K1")) throw cuda_error("FBG_TEST_FAIL_K1: failure injected after the first doubling round (test hook)");
		hi_shift = nbits;
		key_bits = 2 * nbits;
		h *= 2;
	}
	if (sharded) {
		// every GPU wrote the SA entries of its own suffixes only (the rest is still 0): merge like the ranks
		d.begin(K_COMM);
		dist.allreduce_max_u32(p.SA, N, d.stream);
		if (p.have_bwt) dist.allreduce_max_u8(p.BWT, N, d.stream);
		d.end((double) N * 10.0);
	}
	if (peer) {
		// Ranks: every GPU holds the final ranks of its own columns, older (hence smaller) values elsewhere.  K2 also
		// reads ranks beyond its own columns (the first character after a gap run, the '#' of an exhausted row), so
		// the inverse is assembled once, here -- the only rank array that crosses the wire in a build.
		d.begin(K_COMM);
		dist.allreduce_max_u32(p.ISA, N, d.stream);
		d.end((double) N * 8.0);
	}
	d.release(mk);
}

// ----------------------------------------------------------------------- LCP
// flag[k] = 1 where LCP[k] is irreducible (or k = 0)
__global__ void k1_irr_flag_kernel(const u32 *__restrict__ SA, const u8 *__restrict__ T, u64 N, u32 *__restrict__ flag) {
	u64 k = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= N) return;
	u32 irr = 1;
	if (k > 0) {
		u32 a = SA[k], bq = SA[k - 1];
		if (a != 0 && bq != 0 && T[a - 1] == T[bq - 1]) irr = 0;
	}
	flag[k] = irr;
}

// the same from the BWT characters written next to the SA entries (SaOut): BWT[k] = T[SA[k] - 1], 0 for SA[k] = 0
__global__ void k1_irr_flag_bwt_kernel(const u8 *__restrict__ BWT, u64 N, u32 *__restrict__ flag) {
	u64 k = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= N) return;
	flag[k] = (k == 0 || BWT[k] != BWT[k - 1]) ? 1u : 0u;
}

__global__ void k1_irr_compact_kernel(const u32 *__restrict__ SA, const u32 *__restrict__ flag, const u32 *__restrict__ pos, u64 N, u32 *__restrict__ ia, u32 *__restrict__ ib) {
	u64 k = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= N || !flag[k]) return;
	u32 q = pos[k];
	ia[q] = SA[k];
	ib[q] = k ? SA[k - 1] : U32_MAX;
}

// The irreducible ranks straight from the BWT bytes, without an N-word flag array and its scan: a thread takes 16
// consecutive ranks (one 16-byte load + the byte before), a CTA 4096; k1_irr_count leaves the number of irreducible
// ranks per CTA, their exclusive scan (one small CTA) is where k1_irr_emit -- the same flags again, a block-wide
// scan of the per-thread counts -- starts writing (SA[k], SA[k-1]).  0.5 GB read twice instead of 10 GB moved.
constexpr int IRR_ITEMS = 16;
constexpr int IRR_TILE = 256 * IRR_ITEMS;
__device__ __forceinline__ u32 irr_mask16(const u8 *__restrict__ BWT, u64 N, u64 k0) {
	// bit j: rank k0 + j is irreducible (BWT differs from the rank before, or rank 0); BWT is 256-byte aligned, k0 % 16 == 0
	if (k0 >= N) return 0u;
	const uint4 v = *reinterpret_cast<const uint4 *>(BWT + k0); // (BWT holds N + 64 bytes)
	const u32 w[4] = {v.x, v.y, v.z, v.w};
	u32 prev = k0 ? (u32) BWT[k0 - 1] : 0x100u, mask = 0;
#pragma unroll
	for (int j = 0; j < 16; j++) {
		const u32 c = (w[j >> 2] >> (8 * (j & 3))) & 0xFFu;
		if (c != prev && k0 + (u64) j < N) mask |= 1u << j;
		prev = c;
	}
	return mask;
}
__global__ void __launch_bounds__(256) k1_irr_count_kernel(const u8 *__restrict__ BWT, u64 N, u32 *__restrict__ partial) {
	__shared__ u32 sm[32];
	const u32 c = (u32) __popc(irr_mask16(BWT, N, ((u64) blockIdx.x * 256 + threadIdx.x) * IRR_ITEMS));
	u32 tot;
	block_excl_scan<OpSumU32>(c, sm, &tot);
	if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(256) k1_irr_emit_kernel(const u8 *__restrict__ BWT, const u32 *__restrict__ SA, u64 N, const u32 *__restrict__ partial,
                                                          u32 *__restrict__ ia, u32 *__restrict__ ib) {
	__shared__ u32 sm[32];
	const u64 k0 = ((u64) blockIdx.x * 256 + threadIdx.x) * IRR_ITEMS;
	u32 mask = irr_mask16(BWT, N, k0);
	u32 tot;
	u32 q = partial[blockIdx.x] + block_excl_scan<OpSumU32>((u32) __popc(mask), sm, &tot);
	while (mask) {
		const u64 k = k0 + (u32) __ffs(mask) - 1u;
		mask &= mask - 1u;
		ia[q] = SA[k];
		ib[q] = k ? SA[k - 1] : U32_MAX;
		q++;
	}
}

// one warp per irreducible pair: 32 bytes per step, ballot for the first mismatch
__global__ void __launch_bounds__(256) k1_irr_lcp_kernel(const u8 *__restrict__ T, u64 N, const u32 *__restrict__ ia, const u32 *__restrict__ ib, u64 ni, u32 *__restrict__ M) {
	u64 wgl = ((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const u32 l = lane_id();
	for (u64 q = wgl; q < ni; q += ((u64) gridDim.x * blockDim.x) >> 5) {
		u32 a = ia[q], b = ib[q];
		u32 len = 0;
		if (b != U32_MAX) {
			for (;;) {
				u64 x = (u64) a + len + l, y = (u64) b + len + l;
				bool diff = (x >= N) || (y >= N) || (T[x] != T[y]);
				u32 mm = __ballot_sync(0xffffffffu, diff);
				if (mm) { len += __ffs(mm) - 1; break; }
				len += 32;
			}
		}
		if (l == 0) M[a] = a + len;
	}
}

void k1_lcp(Pipeline &p) {
	Dev &d = p.dev;
	const u64 N = p.N;
	if (!p.PLCP) p.PLCP = p.dev.palloc<u32>(N + 1);
	if (p.index_virtual) return; // K2 works from the k-mer list
	if (p.index_lazy) {
		d.begin(K_LCP);
		k1_lazy_lcp_kernel<<<div_up(N, 256), 256, 0, d.stream>>>(p.T, p.SA, p.ISA, N, p.PLCP);
		FBG_LAUNCH_CHECK();
		d.end((double) N * 4.0);
		return;
	}
	const bool from_bwt = p.have_bwt && !getenv("FBG_LCP_FLAGS"); // FBG_LCP_FLAGS=1: the flag-array form also with BWT bytes (test hook)
	d.ensure(N * (from_bwt ? 8 : 16) + (N / SCAN_TILE + 2) * 24 + (N / IRR_TILE + 2) * 4 + (4u << 20));
	u64 mk = d.mark();
	u32 *ia = d.alloc<u32>(N), *ib = d.alloc<u32>(N);
	u32 *total = reinterpret_cast<u32 *>(p.scalars + 8);
	u32 *h_total = reinterpret_cast<u32 *>(p.h_scalars + 8);
	u32 g = div_up(N, 256);
	if (from_bwt) {
		const u32 nt = (u32) div_up(N, IRR_TILE);
		u32 *partial = d.alloc<u32>((u64) nt + 1);
		d.begin(K_LCP);
		k1_irr_count_kernel<<<nt, 256, 0, d.stream>>>(p.BWT, N, partial);
		scan_partials_kernel<OpSumU32><<<1, 1024, 0, d.stream>>>(partial, nt, total);
		k1_irr_emit_kernel<<<nt, 256, 0, d.stream>>>(p.BWT, p.SA, N, partial, ia, ib);
		FBG_CUDA(cudaMemsetAsync(p.PLCP, 0, N * sizeof(u32), d.stream));
		FBG_LAUNCH_CHECK();
		d.note_launches(K_LCP, 2);
	} else {
		u32 *flag = d.alloc<u32>(N), *pos = d.alloc<u32>(N);
		d.begin(K_LCP);
		if (p.have_bwt) k1_irr_flag_bwt_kernel<<<g, 256, 0, d.stream>>>(p.BWT, N, flag);
		else k1_irr_flag_kernel<<<g, 256, 0, d.stream>>>(p.SA, p.T, N, flag);
		FBG_LAUNCH_CHECK();
		d.end((double) N * (4.0 + 2.0 + 4.0));
		device_scan<OpSumU32, true>(d, flag, pos, N, total);
		d.begin(K_LCP);
		k1_irr_compact_kernel<<<g, 256, 0, d.stream>>>(p.SA, flag, pos, N, ia, ib);
		FBG_CUDA(cudaMemsetAsync(p.PLCP, 0, N * sizeof(u32), d.stream));
		FBG_LAUNCH_CHECK();
	}
	FBG_CUDA(cudaMemcpyAsync(h_total, total, sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	u64 ni = *h_total;
	u64 warps = ni < (u64) NUM_SMS * 64 ? ni : (u64) NUM_SMS * 64;
	if (warps == 0) warps = 1;
	k1_irr_lcp_kernel<<<div_up(warps * 32, 256), 256, 0, d.stream>>>(p.T, N, ia, ib, ni, p.PLCP);
	FBG_LAUNCH_CHECK();
	d.end((from_bwt ? (double) N * 6.0 : (double) N * 16.0) + (double) ni * 12.0);
	d.note_launches(K_LCP, 1);
	device_scan<OpMaxU32, false, true>(d, p.PLCP, p.PLCP, N); // PLCP[p] = max_{q <= p} (q + irreducible LCP at q) - p
	(void) g;
	d.release(mk);
}

} // namespace fbg
