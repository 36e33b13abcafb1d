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
	u32 E;         // sorted slots per thread (mpad / tpg)
	u32 groups;    // column groups per CTA
	u64 x0, x1;    // column range [x0, x1)
	int tricks_off, has_ignore;
};

// One column group = tpg threads over mpad = tpg * E sorted slots (E <= 8 consecutive slots
// per thread, held in registers after the sort).
struct FcolCfg { u32 mpad, tpg, groups, threads, E; };
static inline FcolCfg fcol_cfg(u64 m) {
	FcolCfg c;
	c.mpad = 32;
	while (c.mpad < m) c.mpad <<= 1;
	c.tpg = c.mpad / 8;
	if (c.tpg < 32) c.tpg = 32;
	c.E = c.mpad / c.tpg;
	c.threads = c.tpg < 256 ? 256 : c.tpg;
	c.groups = c.threads / c.tpg;
	return c;
}

// Inclusive segmented min over the lanes of a warp, towards higher lanes (UP) or lower
// lanes (!UP): a lane whose flag is set starts a new segment.
template <bool UP> __device__ __forceinline__ void warp_seg_min(u32 &v, bool &fl) {
	const u32 lane = lane_id();
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		u32 ov = UP ? __shfl_up_sync(0xFFFFFFFFu, v, d) : __shfl_down_sync(0xFFFFFFFFu, v, d);
		int of = UP ? __shfl_up_sync(0xFFFFFFFFu, (int) fl, d) : __shfl_down_sync(0xFFFFFFFFu, (int) fl, d);
		const bool in = UP ? lane >= (u32) d : lane + (u32) d < 32u;
		if (in && !fl) { v = ov < v ? ov : v; fl = of != 0; }
	}
}

// dynamic smem per group: key u64[mpad] | A u32[mpad] (PLCP of row i) | RKS u32[mpad] (rank of row i)
__global__ void __launch_bounds__(1024) k2_fcol_kernel(FcolArgs a) {
	extern __shared__ __align__(16) unsigned char k2_raw[];
	const u32 mpad = a.mpad, tpg = a.tpg, E = a.E;
	const u32 grp = threadIdx.x / tpg, tg = threadIdx.x % tpg;
	const u64 per_group = (u64) mpad * 16;
	unsigned char *basep = k2_raw + (u64) grp * per_group;
	u64 *key = reinterpret_cast<u64 *>(basep);
	u32 *A = reinterpret_cast<u32 *>(basep + (u64) mpad * 8);
	u32 *RKS = A + mpad;
	__shared__ u32 red[32];
	__shared__ u32 wl[32], wr[32];      // per warp: segmented aggregate (value), towards the right / the left
	__shared__ u32 wfl[32], wfr[32];    // per warp: a run start / a run end inside the warp
	const u64 m = a.m;
	const u32 wid = threadIdx.x >> 5, lane = lane_id();
	const u32 w_first = (grp * tpg) >> 5, w_last = w_first + (tpg >> 5) - 1;

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
					RKS[i] = rk;
				}
			}
			key[ts_sw(i)] = kk;
		}
		__syncthreads();
		// 2. bitonic sort ascending by (rank, row)
		group_tile_sort(key, mpad, tg, tpg);
		// 3. the thread's E consecutive sorted slots: run flags, LCP[k] and LCP[k+1]
		const u32 j0 = tg * E;
		u32 row[8], lv[8], rv[8];
		u32 smask = 0, emask = 0, vmask = 0;
		u64 kprev = j0 ? key[ts_sw(j0 - 1)] : KEY_INVALID;
		u64 kcur = key[ts_sw(j0)];
#pragma unroll
		for (u32 c = 0; c < 8; c++) {
			if (c < E) {
				const u32 j = j0 + c;
				const u64 kn = j + 1 < mpad ? key[ts_sw(j + 1)] : KEY_INVALID;
				lv[c] = U32_MAX; rv[c] = U32_MAX; row[c] = 0;
				bool s = true, e = true;
				if (kcur != KEY_INVALID) {
					const u32 i = (u32) (kcur & 0xFFFFFFu);
					const u64 r = kcur >> 24;
					row[c] = i;
					vmask |= 1u << c;
					lv[c] = A[i];
					s = (j == 0) || ((kprev >> 24) + 1 != r);
					e = (kn == KEY_INVALID) || ((kn >> 24) != r + 1);
					// LCP[r+1]: inside a run it is the next element's own LCP value; only a run's
					// last element has to look the foreign neighbour up (SA, then PLCP)
					if (!e) rv[c] = A[(u32) (kn & 0xFFFFFFu)];
					else rv[c] = r + 1 < a.N ? a.PLCP[a.SA[r + 1]] : 0u;
				}
				smask |= (u32) s << c; emask |= (u32) e << c;
				kprev = kcur; kcur = kn;
			}
		}
		// 4. segmented prefix-min of lv (run starts) and suffix-min of rv (run ends):
		//    inside the thread, across the warp by shuffles, across the group's warps through
		//    the per-warp aggregates
		u32 runl = U32_MAX, runr = U32_MAX;
#pragma unroll
		for (u32 c = 0; c < 8; c++)
			if (c < E) { runl = ((smask >> c) & 1u) ? lv[c] : (lv[c] < runl ? lv[c] : runl); lv[c] = runl; }
#pragma unroll
		for (int c = 7; c >= 0; c--)
			if ((u32) c < E) { runr = ((emask >> c) & 1u) ? rv[c] : (rv[c] < runr ? rv[c] : runr); rv[c] = runr; }
		u32 sl = runl, sr = runr;
		bool fl = smask != 0, fr = emask != 0;
		warp_seg_min<true>(sl, fl);
		warp_seg_min<false>(sr, fr);
		if (lane == 31) { wl[wid] = sl; wfl[wid] = fl; }
		if (lane == 0) { wr[wid] = sr; wfr[wid] = fr; }
		// what the neighbouring lanes carry into this thread
		u32 pl = __shfl_up_sync(0xFFFFFFFFu, sl, 1), pr = __shfl_down_sync(0xFFFFFFFFu, sr, 1);
		int pfl = __shfl_up_sync(0xFFFFFFFFu, (int) fl, 1), pfr = __shfl_down_sync(0xFFFFFFFFu, (int) fr, 1);
		__syncthreads();
		u32 cl = U32_MAX, cr = U32_MAX;      // carry from the warps before / after this one
		for (u32 w = wid; w > w_first; w--) { u32 q = wl[w - 1]; cl = q < cl ? q : cl; if (wfl[w - 1]) break; }
		for (u32 w = wid; w < w_last; w++) { u32 q = wr[w + 1]; cr = q < cr ? q : cr; if (wfr[w + 1]) break; }
		if (lane > 0) cl = pfl ? pl : (pl < cl ? pl : cl);
		if (lane < 31) cr = pfr ? pr : (pr < cr ? pr : cr);
		// 5. fi per row, max over the column
		u32 fmax = 0;
		bool openl = true;
		u32 closed_r = 0;                    // slots at or after which a run end lies inside the thread
#pragma unroll
		for (int c = 7; c >= 0; c--)
			if ((u32) c < E) { if ((emask >> c) & 1u) closed_r |= (2u << c) - 1u; }
#pragma unroll
		for (u32 c = 0; c < 8; c++) {
			if (c < E) {
				if ((smask >> c) & 1u) openl = false;
				if ((vmask >> c) & 1u) {
					u32 l = lv[c], r = rv[c];
					if (openl) l = cl < l ? cl : l;
					if (!((closed_r >> c) & 1u)) r = cr < r ? cr : r;
					const u32 i = row[c];
					u64 g = (u64) (l > r ? l : r) + 1;                         // fbg.cpp:1656
					u64 gg = (u64) RKS[i] + g;                                  // fbg.cpp:1657
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
			}
		}
		// group reduction (groups are warp multiples)
		fmax = warp_reduce_max(fmax);
		if (lane == 0) red[wid] = fmax;
		__syncthreads();
		if (tg == 0 && live) {
			u32 v = 0;
			for (u32 w = w_first; w <= w_last; w++) v = red[w] > v ? red[w] : v;
			a.f[x] = v > (u32) x ? v : (u32) x;                         // 1618,1681
		}
		__syncthreads();
	}
}

// ---- f(x) from the k-mer list of a "virtual" index (gapless MSA, every k-mer bucket in one
// column, no ignore characters; suffix.cu).  Buckets in rank order: column c[t], LCP h[t] to the
// bucket before (h[nd] = 0).  A run of the reference's coloured leaves is a maximal stretch of
// neighbouring buckets of one column, and every row of bucket t gets the same
//   g = 1 + max( min h[s..t], min h[t+1..e+1] )          (run [s, e]; fbg.cpp:1644-1656)
// because LCP values inside a bucket exceed every h.  Rows are gapless, so rank_i(x) = x,
// rowlen = n, lastcol = n - 1 and col(p + g - 1) = x + g - 1: fi depends on the bucket only
// (fbg.cpp:1657-1666).  Segmented minima by max-scans of (run id, ~h): the run id grows along
// the scan, so the maximum only ever compares values of the current run.
__global__ void k2d_runs_kernel(const u32 *__restrict__ dcol, u32 nd, u32 *__restrict__ seg) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nd) return;
	seg[t] = (t > 0 && dcol[t - 1] != dcol[t]) ? t : 0u; // inclusive max-scan: first bucket of the run
}
__global__ void k2d_keys_kernel(const u32 *__restrict__ seg, const u32 *__restrict__ dlcp, u32 nd, u64 *__restrict__ fw, u64 *__restrict__ bw) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nd) return;
	fw[t] = ((u64) seg[t] << 32) | (u64) (~dlcp[t]);
	bw[nd - 1 - t] = ((u64) (nd - seg[t]) << 32) | (u64) (~dlcp[t + 1]); // backwards the run id must grow too
}
__global__ void k2d_init_kernel(u32 *__restrict__ f, u64 n) {
	u64 x = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (x < n) f[x] = (u32) x; // fbg.cpp:1618
}
__global__ void k2d_f_kernel(const u32 *__restrict__ dcol, const u64 *__restrict__ fw, const u64 *__restrict__ bw, u32 nd, u64 n, int tricks_off, u32 *__restrict__ f) {
	u32 t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nd) return;
	const u64 c = dcol[t];
	if (c >= n || (!tricks_off && c == 0)) return; // '#' and sentinel; rows without a character before x (fbg.cpp:1605-1608)
	const u32 l = ~(u32) fw[t], r = ~(u32) bw[nd - 1 - t];
	const u64 g = (u64) (l > r ? l : r) + 1;
	const u64 gg = c + g;
	const u32 fi = gg > n ? (tricks_off ? (u32) n : (u32) n - 1u) : (u32) (gg - 1);
	atomicMax(&f[c], fi);
}

static void k2_from_dict(Pipeline &p) {
	Dev &d = p.dev;
	const u32 nd = (u32) p.dict_nd;
	d.ensure((u64) nd * 20 + ((u64) nd / SCAN_TILE + 2) * 24 + (1u << 20));
	u64 mk = d.mark();
	u32 *seg = d.alloc<u32>(nd);
	u64 *fw = d.alloc<u64>(nd), *bw = d.alloc<u64>(nd);
	const u32 g = div_up(nd, 256);
	d.begin(K_FCOL);
	k2d_init_kernel<<<div_up(p.n, 256), 256, 0, d.stream>>>(p.f, p.n);
	k2d_runs_kernel<<<g, 256, 0, d.stream>>>(p.dict_col, nd, seg);
	FBG_LAUNCH_CHECK();
	d.end((double) nd * 8.0 + (double) p.n * 4.0);
	d.note_launches(K_FCOL, 1);
	device_scan<OpMaxU32, false>(d, seg, seg, nd);
	d.begin(K_FCOL);
	k2d_keys_kernel<<<g, 256, 0, d.stream>>>(seg, p.dict_lcp, nd, fw, bw);
	FBG_LAUNCH_CHECK();
	d.end((double) nd * 28.0);
	device_scan<OpMaxU64, false>(d, fw, fw, nd);
	device_scan<OpMaxU64, false>(d, bw, bw, nd);
	d.begin(K_FCOL);
	k2d_f_kernel<<<g, 256, 0, d.stream>>>(p.dict_col, fw, bw, nd, p.n, p.opt.disable_tricks ? 1 : 0, p.f);
	FBG_LAUNCH_CHECK();
	d.end((double) nd * 24.0);
	d.release(mk);
	p.have_f = true;
}

void k2_compute_f(Pipeline &p) {
	Dev &d = p.dev;
	if (!p.f) p.f = p.dev.palloc<u32>(p.n + 1);
	if (p.index_virtual) {
		if (p.opt.n_ignore > 0) throw cuda_error("internal: f(x) from the k-mer list with ignore characters set");
		k2_from_dict(p);
		return;
	}
	FcolCfg gc = fcol_cfg(p.m);
	u32 mpad = gc.mpad;
	if (mpad > 8192) throw unsupported_error("more than 8192 rows is not supported by the shared-memory f(x) kernel");
	FcolArgs a;
	a.NG = p.NG; a.IG = p.IG; a.RK = p.RK; a.NIA = p.NIA; a.start = p.start; a.rowlen = p.rowlen; a.lastcol = p.lastcol;
	a.col = p.col; a.ISA = p.ISA; a.SA = p.SA; a.PLCP = p.PLCP; a.f = p.f;
	a.m = p.m; a.n = p.n; a.N = p.N; a.mpad = mpad;
	u32 threads = gc.threads;
	a.tpg = gc.tpg; a.groups = gc.groups; a.E = gc.E;
	a.x0 = 0; a.x1 = p.n;
	const bool sharded = p.dist.world > 1;
	if (sharded) {
		// contiguous column range per GPU; f >= 0 everywhere, so the slices merge by max
		plan_columns(p.n, p.dist.world, p.dist.rank, &a.x0, &a.x1);
		FBG_CUDA(cudaMemsetAsync(p.f, 0, p.n * sizeof(u32), d.stream));
	}
	a.tricks_off = p.opt.disable_tricks ? 1 : 0;
	a.has_ignore = p.opt.n_ignore > 0 ? 1 : 0;
	size_t smem = (size_t) a.groups * mpad * 16;
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
