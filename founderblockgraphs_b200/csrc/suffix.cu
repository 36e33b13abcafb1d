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
#include "pipeline.cuh"

namespace fbg {

// ------------------------------------------------------------ alphabet / k-mers
__global__ void k1_presence_kernel(const u8 *__restrict__ T, u64 N, u32 *__restrict__ present) {
	__shared__ u32 s[256];
	s[threadIdx.x] = 0;
	__syncthreads();
	for (u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (u64) gridDim.x * blockDim.x) s[T[i]] = 1;
	__syncthreads();
	if (s[threadIdx.x]) present[threadIdx.x] = 1;
}

struct CodeTable { u8 code[256]; };

constexpr int KM_THREADS = 256;
constexpr int KM_ITEMS = 8;
constexpr int KM_TILE = KM_THREADS * KM_ITEMS;

__global__ void __launch_bounds__(KM_THREADS) k1_kmer_kernel(const u8 *__restrict__ T, u64 N, CodeTable tab, int K, int b, u64 *__restrict__ keys, u32 *__restrict__ vals) {
	__shared__ u8 sc[KM_TILE + 64];
	__shared__ u8 stab[256];
	stab[threadIdx.x] = tab.code[threadIdx.x];
	__syncthreads();
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
#pragma unroll
	for (int it = 0; it < KM_ITEMS; it++) {
		u64 g = base + o + it;
		if (g < N) { keys[g] = key; vals[g] = (u32) g; }
		key = ((key << b) | sc[o + it + K]) & kmask;
	}
}

// ------------------------------------------------------------- doubling rounds
// HH[t] = (index of bucket head << 32) | index of sub-bucket head, 0 where not a head
__global__ void k1_heads_kernel(const u64 *__restrict__ keys, u64 na, int hi_shift, u64 *__restrict__ HH) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na) return;
	u64 k = keys[t];
	u64 kp = t ? keys[t - 1] : 0;
	bool hs = (t == 0) || (k != kp);
	bool hb = (t == 0) || (hi_shift < 64 && (k >> hi_shift) != (kp >> hi_shift));
	HH[t] = ((hb ? t : 0) << 32) | (hs ? t : 0);
}

// new rank = bucket rank + (sub-head index - bucket-head index); flag = still tied
__global__ void k1_update_kernel(const u64 *__restrict__ keys, const u32 *__restrict__ vals, const u64 *__restrict__ HH, u64 na, int hi_shift,
                                 u32 *__restrict__ R, u32 *__restrict__ newr, u32 *__restrict__ flag) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na) return;
	u64 k = keys[t];
	u64 hh = HH[t];
	u32 hb = (u32) (hh >> 32), hs = (u32) hh;
	u32 base = hi_shift < 64 ? (u32) (k >> hi_shift) : 0u;
	u32 nr = base + (hs - hb);
	bool single = (hs == (u32) t) && (t + 1 == na || keys[t + 1] != k);
	R[vals[t]] = nr;
	newr[t] = nr;
	flag[t] = single ? 0u : 1u;
}

__global__ void k1_compact_kernel(const u32 *__restrict__ vals, const u32 *__restrict__ newr, const u32 *__restrict__ flag, const u32 *__restrict__ pos, u64 na,
                                  u32 *__restrict__ lp, u32 *__restrict__ lr) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na) return;
	if (flag[t]) { u32 q = pos[t]; lp[q] = vals[t]; lr[q] = newr[t]; }
}

// key = (rank[p] << nbits) | rank[p+h]; tied suffixes never contain the sentinel in
// their first h symbols, so p+h <= N-1
__global__ void k1_keys_kernel(const u32 *__restrict__ lp, const u32 *__restrict__ lr, const u32 *__restrict__ R, u64 na, u64 h, u64 N, int nbits,
                               u64 *__restrict__ keys, u32 *__restrict__ vals) {
	u64 t = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= na) return;
	u32 p = lp[t];
	u64 q = (u64) p + h;
	u32 r2 = q < N ? R[q] : 0u;
	keys[t] = ((u64) lr[t] << nbits) | r2;
	vals[t] = p;
}

__global__ void k1_invert_kernel(const u32 *__restrict__ R, u64 N, u32 *__restrict__ SA) {
	u64 p = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (p < N) SA[R[p]] = (u32) p;
}

void k1_suffix_array(Pipeline &p) {
	Dev &d = p.dev;
	const u64 N = p.N;
	p.ISA = dmalloc<u32>(N + 1);
	p.SA = dmalloc<u32>(N + 1);
	u64 mk = d.mark();

	// alphabet: dense order-preserving codes (sentinel 0 is always present)
	u32 *present = d.alloc<u32>(256);
	FBG_CUDA(cudaMemsetAsync(present, 0, 256 * sizeof(u32), d.stream));
	d.begin(K_KMER);
	k1_presence_kernel<<<NUM_SMS * 4, 256, 0, d.stream>>>(p.T, N, present);
	FBG_LAUNCH_CHECK();
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

	RadixBuffers rb;
	rb.k[0] = d.alloc<u64>(N); rb.k[1] = d.alloc<u64>(N);
	rb.v[0] = d.alloc<u32>(N); rb.v[1] = d.alloc<u32>(N);
	rb.cur = 0;
	u64 *HH = d.alloc<u64>(N);
	u32 *newr = d.alloc<u32>(N), *flag = d.alloc<u32>(N), *pos = d.alloc<u32>(N);
	u32 *lp = d.alloc<u32>(N), *lr = d.alloc<u32>(N);
	u32 *total = reinterpret_cast<u32 *>(p.scalars + 8);
	u32 *h_total = reinterpret_cast<u32 *>(p.h_scalars + 8);

	k1_kmer_kernel<<<div_up(N, KM_TILE), KM_THREADS, 0, d.stream>>>(p.T, N, tab, K, b, (u64 *) rb.k[0], rb.v[0]);
	FBG_LAUNCH_CHECK();
	d.end((double) N * (1.0 + 8.0 + 4.0));
	d.note_launches(K_KMER, 1);

	int nbits = 1;
	while ((1ull << nbits) < N) nbits++;
	u64 na = N;
	u64 h = (u64) K;
	int hi_shift = 64; // round 0: a single level of buckets
	int key_bits = K * b;
	p.sa_rounds = 0;
	for (;;) {
		int cur = radix_sort<u64, true>(d, rb, na, 0, key_bits);
		const u64 *sk = (const u64 *) rb.k[cur];
		const u32 *sv = rb.v[cur];
		u32 g = div_up(na, 256);
		d.begin(K_DOUBLING_AUX);
		k1_heads_kernel<<<g, 256, 0, d.stream>>>(sk, na, hi_shift, HH);
		FBG_LAUNCH_CHECK();
		d.end((double) na * 16.0);
		device_scan<OpMax2x32, false>(d, HH, HH, na);
		d.begin(K_DOUBLING_AUX);
		k1_update_kernel<<<g, 256, 0, d.stream>>>(sk, sv, HH, na, hi_shift, p.ISA, newr, flag);
		FBG_LAUNCH_CHECK();
		d.end((double) na * (8.0 + 4.0 + 8.0 + 4.0 + 4.0 + 4.0));
		device_scan<OpSumU32, true>(d, flag, pos, na, total);
		d.begin(K_DOUBLING_AUX);
		k1_compact_kernel<<<g, 256, 0, d.stream>>>(sv, newr, flag, pos, na, lp, lr);
		FBG_LAUNCH_CHECK();
		d.end((double) na * 16.0);
		FBG_CUDA(cudaMemcpyAsync(h_total, total, sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		p.sa_rounds++;
		na = *h_total;
		if (na == 0) break;
		if (h >= N) throw cuda_error("prefix doubling did not converge");
		d.begin(K_DOUBLING_AUX);
		k1_keys_kernel<<<div_up(na, 256), 256, 0, d.stream>>>(lp, lr, p.ISA, na, h, N, nbits, (u64 *) rb.k[rb.cur], rb.v[rb.cur]);
		FBG_LAUNCH_CHECK();
		d.end((double) na * (4.0 + 4.0 + 4.0 + 8.0 + 4.0));
		hi_shift = nbits;
		key_bits = 2 * nbits;
		h *= 2;
	}
	d.begin(K_DOUBLING_AUX);
	k1_invert_kernel<<<div_up(N, 256), 256, 0, d.stream>>>(p.ISA, N, p.SA);
	FBG_LAUNCH_CHECK();
	d.end((double) N * 8.0);
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

__global__ void k1_irr_compact_kernel(const u32 *__restrict__ SA, const u32 *__restrict__ flag, const u32 *__restrict__ pos, u64 N, u32 *__restrict__ ia, u32 *__restrict__ ib) {
	u64 k = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= N || !flag[k]) return;
	u32 q = pos[k];
	ia[q] = SA[k];
	ib[q] = k ? SA[k - 1] : U32_MAX;
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

__global__ void k1_plcp_kernel(u32 *__restrict__ M, u64 N) {
	u64 p = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (p < N) M[p] = M[p] - (u32) p;
}

// NLCP[SA[k]] = PLCP[SA[k+1]]  (0 for the last rank)
__global__ void k1_nlcp_kernel(const u32 *__restrict__ SA, const u32 *__restrict__ PLCP, u64 N, u32 *__restrict__ NLCP) {
	u64 k = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= N) return;
	NLCP[SA[k]] = (k + 1 < N) ? PLCP[SA[k + 1]] : 0u;
}

void k1_lcp(Pipeline &p) {
	Dev &d = p.dev;
	const u64 N = p.N;
	p.PLCP = dmalloc<u32>(N + 1);
	p.NLCP = dmalloc<u32>(N + 1);
	u64 mk = d.mark();
	u32 *flag = d.alloc<u32>(N), *pos = d.alloc<u32>(N), *ia = d.alloc<u32>(N), *ib = d.alloc<u32>(N);
	u32 *total = reinterpret_cast<u32 *>(p.scalars + 8);
	u32 *h_total = reinterpret_cast<u32 *>(p.h_scalars + 8);
	u32 g = div_up(N, 256);
	d.begin(K_LCP);
	k1_irr_flag_kernel<<<g, 256, 0, d.stream>>>(p.SA, p.T, N, flag);
	FBG_LAUNCH_CHECK();
	d.end((double) N * (4.0 + 2.0 + 4.0));
	device_scan<OpSumU32, true>(d, flag, pos, N, total);
	d.begin(K_LCP);
	k1_irr_compact_kernel<<<g, 256, 0, d.stream>>>(p.SA, flag, pos, N, ia, ib);
	FBG_CUDA(cudaMemsetAsync(p.PLCP, 0, N * sizeof(u32), d.stream));
	FBG_LAUNCH_CHECK();
	FBG_CUDA(cudaMemcpyAsync(h_total, total, sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	u64 ni = *h_total;
	u64 warps = ni < (u64) NUM_SMS * 64 ? ni : (u64) NUM_SMS * 64;
	if (warps == 0) warps = 1;
	k1_irr_lcp_kernel<<<div_up(warps * 32, 256), 256, 0, d.stream>>>(p.T, N, ia, ib, ni, p.PLCP);
	FBG_LAUNCH_CHECK();
	d.end((double) N * 16.0 + (double) ni * 12.0);
	d.note_launches(K_LCP, 1);
	device_scan<OpMaxU32, false>(d, p.PLCP, p.PLCP, N);
	d.begin(K_LCP);
	k1_plcp_kernel<<<g, 256, 0, d.stream>>>(p.PLCP, N);
	k1_nlcp_kernel<<<g, 256, 0, d.stream>>>(p.SA, p.PLCP, N, p.NLCP);
	FBG_LAUNCH_CHECK();
	d.end((double) N * (8.0 + 4.0 + 4.0 + 4.0 + 4.0));
	d.note_launches(K_LCP, 1);
	d.release(mk);
}

} // namespace fbg
