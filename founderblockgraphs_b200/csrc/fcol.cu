// fcol.cu -- K2: the minimal right extension f(x) of every MSA column.
//
// Replaces compute_f / compute_f_range (fbg.cpp:1579-1695, 1475-1577).  The
// reference walks, per column, the compressed suffix tree around the m current
// row suffixes; here one thread group owns one column and works on flat arrays:
//
//   p_i(x) = start_i + rank_i(x)                     (RK table + NG bitmap popcount)
//   k_i    = ISA[p_i],  a_i = LCP[k_i] = PLCP[p_i],  b_i = LCP[k_i+1] (the next row's a inside a
//            run of consecutive ranks, PLCP[SA[k_i+1]] at a run's end)
//   sort the active rows by k_i (bitonic network in shared memory)
//   maximal runs of consecutive ranks = the reference's coloured leaf runs
//   (fbg.cpp:1633-1641); segmented prefix-min of a / suffix-min of b over a run
//   give  g = 1 + max(min LCP[a..k], min LCP[k+1..b+1])  = depth(parent(w)) + 1
//   (fbg.cpp:1644-1656); then fi per fbg.cpp:1657-1670 and f[x] = max(x, max fi).
//
// All ISA/PLCP reads of one row advance by one text position per column, so
// consecutive columns of a CTA re-use the same sectors.
#include "pipeline.cuh"
#include "blocksort.cuh"

namespace fbg {

constexpr u64 KEY_INVALID = ~0ull;

struct FcolArgs {
	const u32 *NG, *IG, *RK, *NIA, *start, *rowlen, *lastcol, *col, *ISA, *SA, *PLCP;
	u32 *f;
	u64 m, n, N;
	u32 mpad;      // power of two >= m
	u32 tpg;       // threads per column group
	u32 groups;    // column groups per CTA
	u64 x0, x1;    // column range [x0, x1)
	int tricks_off, has_ignore;
};

// dynamic smem per group: key u64[mpad] | A u32[mpad] | L u32[mpad] | R u32[mpad] | fl u8[mpad] | fr u8[mpad]
__global__ void __launch_bounds__(1024) k2_fcol_kernel(FcolArgs a) {
	extern __shared__ __align__(16) unsigned char k2_raw[];
	const u32 mpad = a.mpad, tpg = a.tpg;
	const u32 grp = threadIdx.x / tpg, tg = threadIdx.x % tpg;
	const u64 per_group = (u64) mpad * (8 + 3 * 4 + 2);
	unsigned char *basep = k2_raw + (u64) grp * per_group;
	u64 *key = reinterpret_cast<u64 *>(basep);
	u32 *A = reinterpret_cast<u32 *>(basep + (u64) mpad * 8);
	u32 *L = A + mpad, *R = L + mpad;
	u8 *fl = reinterpret_cast<u8 *>(R + mpad), *fr = fl + mpad;
	__shared__ u32 red[32];
	const u64 m = a.m;

	for (u64 xb = a.x0 + (u64) blockIdx.x * a.groups; xb < a.x1; xb += (u64) gridDim.x * a.groups) {
		const u64 x = xb + grp;
		const bool live = x < a.x1;
		const u64 t = x >> 5;
		const u32 dx = (u32) (x & 31), below = (1u << dx) - 1u;
		// 1. gather
		for (u32 i = tg; i < mpad; i += tpg) {
			u64 kk = KEY_INVALID;
			if (live && i < m) {
				u32 rk = a.RK[t * m + i] + __popc(a.NG[t * m + i] & below);
				if (a.tricks_off || rk > 0) {
					u64 p = (u64) a.start[i] + rk;
					kk = ((u64) a.ISA[p] << 24) | i;
					A[i] = a.PLCP[p];
				}
			}
			key[ts_sw(i)] = kk;
		}
		__syncthreads();
		// 2. bitonic sort ascending by (rank, row)
		group_tile_sort(key, mpad, tg, tpg);
		// 3. run flags + initial values (sorted order)
		for (u32 j = tg; j < mpad; j += tpg) {
			u64 kj = key[ts_sw(j)];
			u32 lv = U32_MAX, rv = U32_MAX; u8 s = 1, e = 1;
			if (kj != KEY_INVALID) {
				u32 i = (u32) (kj & 0xFFFFFFu);
				u64 r = kj >> 24;
				lv = A[i];
				s = (j == 0) || ((key[ts_sw(j - 1)] >> 24) + 1 != r);
				u64 kn = j + 1 < mpad ? key[ts_sw(j + 1)] : KEY_INVALID;
				e = (kn == KEY_INVALID) || ((kn >> 24) != r + 1);
				// LCP[r+1]: inside a run it is the next element's own LCP value; only a run's
				// last element has to look the foreign neighbour up (SA, then PLCP)
				if (!e) rv = A[(u32) (kn & 0xFFFFFFu)];
				else rv = r + 1 < a.N ? a.PLCP[a.SA[r + 1]] : 0u;
			}
			L[j] = lv; R[j] = rv; fl[j] = s; fr[j] = e;
		}
		__syncthreads();
		// 4. segmented prefix-min (L, run starts) and suffix-min (R, run ends), Hillis-Steele
		for (u32 o = 1; o < mpad; o <<= 1) {
			// read phase into registers (each thread owns the same j's in both phases)
			u32 nl[8], nr[8]; u8 nfl[8], nfr[8];
			u32 c = 0;
			for (u32 j = tg; j < mpad; j += tpg, c++) {
				u32 lv = L[j], rv = R[j]; u8 s = fl[j], e = fr[j];
				if (!s && j >= o) { u32 q = L[j - o]; lv = q < lv ? q : lv; s = fl[j - o]; }
				if (!e && j + o < mpad) { u32 q = R[j + o]; rv = q < rv ? q : rv; e = fr[j + o]; }
				nl[c] = lv; nr[c] = rv; nfl[c] = s; nfr[c] = e;
			}
			__syncthreads();
			c = 0;
			for (u32 j = tg; j < mpad; j += tpg, c++) { L[j] = nl[c]; R[j] = nr[c]; fl[j] = nfl[c]; fr[j] = nfr[c]; }
			__syncthreads();
		}
		// 5. fi per row, max over the column
		u32 fmax = 0;
		for (u32 j = tg; j < mpad; j += tpg) {
			u64 kj = key[ts_sw(j)];
			if (kj == KEY_INVALID) continue;
			u32 i = (u32) (kj & 0xFFFFFFu);
			u32 lv = L[j], rv = R[j];
			u64 g = (u64) (lv > rv ? lv : rv) + 1;                      // fbg.cpp:1656
			u32 ng = a.NG[t * m + i];
			u64 rk = (u64) a.RK[t * m + i] + __popc(ng & below);
			u64 gg = rk + g;                                            // fbg.cpp:1657
			u32 fi;
			if (gg > a.rowlen[i]) fi = a.tricks_off ? (u32) a.n : a.lastcol[i]; // 1659-1665
			else fi = a.col[(u64) a.start[i] + gg - 1];                 // 1666
			if (a.has_ignore) {                                         // 1669-1670
				u32 ig = a.IG[t * m + i] >> dx;
				u32 ni = ig ? (u32) x + (__ffs(ig) - 1) : a.NIA[t * m + i];
				if (ni < fi) fi = ni;
			}
			fmax = fi > fmax ? fi : fmax;
		}
		// group reduction (groups are warp multiples)
		fmax = warp_reduce_max(fmax);
		if (lane_id() == 0) red[threadIdx.x >> 5] = fmax;
		__syncthreads();
		if (tg == 0 && live) {
			u32 v = 0;
			u32 w0 = threadIdx.x >> 5, nw = tpg >> 5;
			for (u32 w = 0; w < nw; w++) v = red[w0 + w] > v ? red[w0 + w] : v;
			a.f[x] = v > (u32) x ? v : (u32) x;                         // 1618,1681
		}
		__syncthreads();
	}
}

void k2_compute_f(Pipeline &p) {
	Dev &d = p.dev;
	if (!p.f) p.f = p.dev.palloc<u32>(p.n + 1);
	GroupCfg gc = group_cfg(p.m);
	u32 mpad = gc.mpad;
	if (mpad > 8192) throw cuda_error("more than 8192 rows per column group is not supported by the shared-memory f(x) kernel yet");
	FcolArgs a;
	a.NG = p.NG; a.IG = p.IG; a.RK = p.RK; a.NIA = p.NIA; a.start = p.start; a.rowlen = p.rowlen; a.lastcol = p.lastcol;
	a.col = p.col; a.ISA = p.ISA; a.SA = p.SA; a.PLCP = p.PLCP; a.f = p.f;
	a.m = p.m; a.n = p.n; a.N = p.N; a.mpad = mpad;
	// the Hillis-Steele phase keeps <= 8 elements per thread in registers
	u32 threads = gc.threads;
	a.tpg = gc.tpg; a.groups = gc.groups;
	a.x0 = 0; a.x1 = p.n;
	const bool sharded = p.dist.world > 1;
	if (sharded) {
		// contiguous column range per GPU; f >= 0 everywhere, so the slices merge by max
		plan_columns(p.n, p.dist.world, p.dist.rank, &a.x0, &a.x1);
		FBG_CUDA(cudaMemsetAsync(p.f, 0, p.n * sizeof(u32), d.stream));
	}
	a.tricks_off = p.opt.disable_tricks ? 1 : 0;
	a.has_ignore = p.opt.n_ignore > 0 ? 1 : 0;
	size_t smem = (size_t) a.groups * mpad * (8 + 12 + 2);
	FBG_CUDA(cudaFuncSetAttribute(k2_fcol_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	u64 col_groups = (a.x1 - a.x0 + a.groups - 1) / a.groups;
	if (col_groups == 0) col_groups = 1;
	u64 max_grid = (u64) NUM_SMS * 8;
	u32 grid = (u32) (col_groups < max_grid ? col_groups : max_grid);
	d.begin(K_FCOL);
	k2_fcol_kernel<<<grid, threads, smem, d.stream>>>(a);
	FBG_LAUNCH_CHECK();
	d.end((double) p.m * (double) (a.x1 - a.x0) * 12.0);
	if (sharded) {
		d.begin(K_COMM);
		p.dist.allreduce_max_u32(p.f, p.n, d.stream);
		d.end((double) p.n * 8.0);
	}
	p.have_f = true;
}

} // namespace fbg
