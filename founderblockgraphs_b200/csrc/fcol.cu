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
	int only_marked; // sorting kernel after the cluster kernel: only the columns it left marked
	const u32 *RC, *AC; // staged form: rank and LCP of row i at column x at [(x - xs0) * m + i] (k2_stage_kernel); null: gathered here
	u64 xs0;
	u32 *dbg;        // FBG_DEBUG: [0] columns left marked, by reason: [1] no cluster, [2] too many rows outside, [3] a row next to the window
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

// fi of row i at column x, given the two minima l = min LCP[a..k], r = min LCP[k+1..b+1] of its run (fbg.cpp:1656-1670)
__device__ __forceinline__ u32 fcol_fi(const FcolArgs &a, u32 i, u32 l, u32 r, u32 rk, u64 x, u64 t, u32 dx) {
	u64 g = (u64) (l > r ? l : r) + 1;                         // fbg.cpp:1656
	u64 gg = (u64) rk + g;                                      // fbg.cpp:1657
	u32 fi;
	if (gg > a.rowlen[i]) fi = a.tricks_off ? (u32) a.n : a.lastcol[i]; // 1659-1665
	else fi = a.col[(u64) a.start[i] + gg - 1];                 // 1666
	if (a.has_ignore) {                                         // 1669-1670
		u32 ig = a.IG[t * a.m + i] >> dx;
		u32 ni = ig ? (u32) x + (__ffs(ig) - 1) : a.NIA[t * a.m + i];
		if (ni < fi) fi = ni;
	}
	return fi;
}

// ---- staging: ISA[p_i(x)] and PLCP[p_i(x)] in COLUMN-major order for a slab of columns.
// The column kernels below need, per column, the rank and the LCP of every row's current suffix: two 4-byte gathers
// per cell, m different 128-byte lines per column and array.  Along a ROW the same reads are consecutive addresses
// (p_i(x + 1) = p_i(x) or p_i(x) + 1), so this kernel reads them the cheap way round -- a warp takes one row and the 32
// columns of one bitmap word: one or two lines per array -- transposes 32 x 32 tiles in shared memory and writes
// lines of 32 consecutive rows of one column.  16 B per cell of streaming traffic instead of 64 B of sectors fetched
// through the column kernel's dependent-load chain.  A row that is not evaluated at x (fbg.cpp:1605-1608: no character
// before the column, tricks on) gets rank U32_MAX.
__global__ void __launch_bounds__(256) k2_stage_kernel(FcolArgs a, u64 xs0, u64 xs1, u32 *__restrict__ RC, u32 *__restrict__ AC) {
	__shared__ u32 tr[32][33], ta[32][33];
	const u32 w = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const u64 m = a.m, i0 = (u64) blockIdx.x * 32, t = (xs0 >> 5) + blockIdx.y, x = t * 32 + lane;
	const u32 below = (1u << lane) - 1u;
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 rr = w + 8 * k;
		const u64 i = i0 + rr;
		u32 r = U32_MAX, v = 0;
		if (i < m && x < xs1) {
			const u32 rk = __ldg(a.RK + t * m + i) + __popc(__ldg(a.NG + t * m + i) & below);
			if (a.tricks_off || rk > 0) {
				const u64 p = (u64) __ldg(a.start + i) + rk;
				r = __ldg(a.ISA + p); v = __ldg(a.PLCP + p);
			}
		}
		tr[rr][lane] = r; ta[rr][lane] = v;
	}
	__syncthreads();
#pragma unroll
	for (u32 k = 0; k < 4; k++) {
		const u32 c = w + 8 * k;
		const u64 xc = t * 32 + c, i = i0 + lane;
		if (i < m && xc < xs1) { RC[(xc - xs0) * m + i] = tr[lane][c]; AC[(xc - xs0) * m + i] = ta[lane][c]; }
	}
}

constexpr u64 FC_CHUNK = 8;           // consecutive columns (column groups) a CTA works through before it moves on
constexpr u32 FC_OCAP = 512;          // cluster path: rows of a column whose rank lies outside the column's cluster
constexpr u32 FC_MARK = 0xFFFFFFFFu;  // f[x] of a column the cluster path could not take (redone by the sorting kernel)

// CLUSTER = false -- dynamic smem per group: key u64[mpad] | A u32[mpad] (PLCP of row i) | RKS u32[mpad] (rank of row i);
//   the active rows are sorted by rank on the swizzled bitonic network.
// CLUSTER = true (one group per CTA, mpad / 8 threads) -- A u32[mpad] | RKS u32[mpad] | TAB u16[mpad] | OKEY u64[FC_OCAP].
//   The rows of an MSA are near copies of each other, so the suffixes that start in one column lie next to
//   each other in the suffix array: the ranks of a column's rows form one cluster about m wide (holes where a
//   suffix of another column sorts in between), plus a few rows elsewhere -- rows inside a gap run (their next
//   character lies in a later column, hence in that column's cluster) and exhausted rows (their '#').  Instead of
//   sorting, every row is PLACED at TAB[rank - base], base = the smallest rank near the median of three
//   probe rows: adjacent slots are consecutive ranks, an empty slot is a foreign suffix, i.e. a run break
//   (fbg.cpp:1633-1641), and the run minima are the same register / shuffle scans over the slots.  Rows outside
//   [base, base + mpad) are few: they go through a 512-key network and walk their (short) runs one thread
//   each.  A column that does not fit the pattern -- no cluster, more than FC_OCAP rows outside, a row
//   right next to the window (its run would cross the edge) -- gets f[x] = FC_MARK and is redone by the
//   sorting form of this kernel (only_marked).
template <bool CLUSTER> __global__ void __launch_bounds__(1024) k2_fcol_kernel(FcolArgs a) {
	extern __shared__ __align__(16) unsigned char k2_raw[];
	const u32 mpad = a.mpad, tpg = a.tpg, E = CLUSTER ? 8u : a.E;
	const u32 grp = threadIdx.x / tpg, tg = threadIdx.x % tpg;
	const u64 per_group = (u64) mpad * 16;
	unsigned char *basep = k2_raw + (CLUSTER ? 0 : (u64) grp * per_group);
	u64 *key = reinterpret_cast<u64 *>(basep);                                             // !CLUSTER
	u32 *A = reinterpret_cast<u32 *>(basep + (CLUSTER ? 0 : (u64) mpad * 8));
	u32 *RKS = A + mpad;
	unsigned short *TAB = reinterpret_cast<unsigned short *>(RKS + mpad);                   // CLUSTER
	u64 *OKEY = reinterpret_cast<u64 *>(basep + (u64) mpad * 10);                           // CLUSTER
	__shared__ u32 red[32];
	__shared__ u32 wl[32], wr[32];      // per warp: segmented aggregate (value), towards the right / the left
	__shared__ u32 wfl[32], wfr[32];    // per warp: a run start / a run end inside the warp
	__shared__ u32 s_probe[3], s_no, s_fb;
	const u64 m = a.m;
	const u32 wid = threadIdx.x >> 5, lane = lane_id();
	const u32 w_first = (grp * tpg) >> 5, w_last = w_first + (tpg >> 5) - 1;

	// A CTA takes chunks of FC_CHUNK consecutive column groups: the text positions of one row at neighbouring columns
	// are neighbours, so the ISA / PLCP / col sectors a column reads are read again by the next seven -- from this
	// SM's L1 / L2 instead of from HBM (with single columns dealt out round-robin, ncu showed 35 GB of DRAM reads
	// for 6 GB of algorithmic bytes on cfg3: every column fetched its sectors anew)
	const u64 ngroups_total = (a.x1 - a.x0 + a.groups - 1) / a.groups;
	for (u64 it = (u64) blockIdx.x * FC_CHUNK; it < ngroups_total; it = (it % FC_CHUNK == FC_CHUNK - 1) ? it + 1 + ((u64) gridDim.x - 1) * FC_CHUNK : it + 1) {
		const u64 xb = a.x0 + it * a.groups;
		const u64 x = xb + grp;
		const bool live = x < a.x1;
		if (a.only_marked && a.f[x] != FC_MARK) continue; // (one group per CTA: uniform)
		const u64 t = x >> 5;
		const u32 dx = (u32) (x & 31), below = (1u << dx) - 1u;
		u32 base = 0, n_out = 0;
		if (!CLUSTER) {
			// 1. gather
			for (u32 i = tg; i < mpad; i += tpg) {
				u64 kk = KEY_INVALID;
				if (live && i < m) {
					u32 rk = a.RK[t * m + i] + __popc(a.NG[t * m + i] & below);
					if (a.RC) {
						const u32 r = a.RC[(x - a.xs0) * m + i];
						if (r != U32_MAX) { kk = ((u64) r << 24) | i; A[i] = a.AC[(x - a.xs0) * m + i]; RKS[i] = rk; }
					} else if (a.tricks_off || rk > 0) {
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
		} else {
			// 1c. gather: this thread's rows are tg, tg + tpg, ... (eight of them); ranks stay in registers
			u32 myr[8];
			if (tg < 3) s_probe[tg] = U32_MAX;
			if (tg == 0) { s_no = 0; s_fb = 0; }
			reinterpret_cast<uint4 *>(TAB)[tg] = make_uint4(0, 0, 0, 0); // eight slots per thread
			__syncthreads();
			const u32 pr0 = (u32) (m / 4), pr1 = (u32) (m / 2), pr2 = (u32) (3 * m / 4);
			// (three batches of loads -- tables, text offsets, then ranks and LCPs -- so that a thread waits for memory
			//  three times per column, not three times per row)
			u32 rkv[8], av[8];
			u64 pv[8];
#pragma unroll
			for (u32 c = 0; c < 8; c++) {
				const u32 i = tg + c * tpg;
				rkv[c] = 0; pv[c] = 0;
				if (live && i < m) { rkv[c] = __ldg(a.RK + t * m + i) + __popc(__ldg(a.NG + t * m + i) & below); pv[c] = __ldg(a.start + i); }
			}
#pragma unroll
			for (u32 c = 0; c < 8; c++) {
				const u32 i = tg + c * tpg;
				myr[c] = U32_MAX; av[c] = 0;
				if (a.RC) { if (live && i < m) { myr[c] = __ldg(a.RC + (x - a.xs0) * m + i); av[c] = __ldg(a.AC + (x - a.xs0) * m + i); } }
				else if (live && i < m && (a.tricks_off || rkv[c] > 0)) { pv[c] += rkv[c]; myr[c] = __ldg(a.ISA + pv[c]); av[c] = __ldg(a.PLCP + pv[c]); }
			}
#pragma unroll
			for (u32 c = 0; c < 8; c++) {
				const u32 i = tg + c * tpg;
				if (i < m) {
					if (myr[c] != U32_MAX) { A[i] = av[c]; RKS[i] = rkv[c]; }
					if (i == pr0) s_probe[0] = myr[c];
					if (i == pr1) s_probe[1] = myr[c];
					if (i == pr2) s_probe[2] = myr[c];
				}
			}
			__syncthreads();
			// 2c. the cluster lies around the median probe: k rows in the mpad ranks below it (the cluster itself is
			//     contiguous up to a few foreign suffixes; stray rows -- an SNP a few columns on puts a row some
			//     thousand ranks away -- only add a few to the count) => window start = pivot - k - slack
			u32 pa = s_probe[0], pb = s_probe[1], pc = s_probe[2];
			const u32 lo3 = pa < pb ? pa : pb, hi3 = pa < pb ? pb : pa;
			const u32 piv = pc < lo3 ? lo3 : (pc > hi3 ? hi3 : pc); // median (inactive rows = U32_MAX sort last)
			u32 kb = 0;
#pragma unroll
			for (u32 c = 0; c < 8; c++) {
				const u32 r = myr[c];
				if (r < piv && piv - r <= mpad) kb++;
			}
			kb = warp_reduce_sum(kb);
			if (lane == 0) red[wid] = kb;
			__syncthreads();
			kb = 0;
			for (u32 w = w_first; w <= w_last; w++) kb += red[w];
			const u32 slack = (mpad - (u32) m) / 4;
			base = piv == U32_MAX ? U32_MAX : (piv > kb + slack ? piv - kb - slack : 0u);
			// 3c. place the rows of the window, list the others
#pragma unroll
			for (u32 c = 0; c < 8; c++) {
				const u32 r = myr[c];
				if (r != U32_MAX) {
					const u32 i = tg + c * tpg;
					const u32 dlt = r - base; // (wraps for r < base)
					if (dlt < mpad) TAB[dlt] = (unsigned short) (i + 1);
					else {
						const u32 q = atomicAdd(&s_no, 1u);
						if (q < FC_OCAP) OKEY[ts_sw(q)] = ((u64) r << 24) | i;
						if (r + 1 == base || dlt == mpad) s_fb = 1u; // next to the window: its run would cross the edge
					}
				}
			}
			__syncthreads();
			n_out = s_no;
			const bool bad = piv == U32_MAX || base == U32_MAX || n_out > FC_OCAP || s_fb != 0;
			__syncthreads(); // (s_no / s_fb / red are rewritten by the next column)
			if (bad) {
				if (tg == 0 && live) {
					a.f[x] = FC_MARK;
					if (a.dbg) { atomicAdd(a.dbg, 1u); if (piv == U32_MAX || base == U32_MAX) atomicAdd(a.dbg + 1, 1u); else if (n_out > FC_OCAP) atomicAdd(a.dbg + 2, 1u); else atomicAdd(a.dbg + 3, 1u); }
				}
				continue;
			}
		}
		// 3. the thread's E consecutive sorted slots: run flags, LCP[k] and LCP[k+1]
		const u32 j0 = tg * E;
		u32 row[8], lv[8], rv[8];
		u32 smask = 0, emask = 0, vmask = 0;
		if (!CLUSTER) {
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
		} else {
			// slot j holds the row whose rank is base + j (0 = none)
			const uint4 tv = reinterpret_cast<const uint4 *>(TAB)[tg];
			const u32 w4[4] = {tv.x, tv.y, tv.z, tv.w};
			u32 prev = j0 ? TAB[j0 - 1] : 0u;
			const u32 nxt8 = j0 + 8 < mpad ? TAB[j0 + 8] : 0u;
#pragma unroll
			for (u32 c = 0; c < 8; c++) {
				const u32 cur = (w4[c >> 1] >> (16 * (c & 1))) & 0xFFFFu;
				const u32 nx = c < 7 ? (w4[(c + 1) >> 1] >> (16 * ((c + 1) & 1))) & 0xFFFFu : nxt8;
				lv[c] = U32_MAX; rv[c] = U32_MAX; row[c] = 0;
				bool s = true, e = true;
				if (cur) {
					const u32 i = cur - 1;
					const u64 r = (u64) base + j0 + c;
					row[c] = i;
					vmask |= 1u << c;
					lv[c] = A[i];
					s = prev == 0;
					e = nx == 0;
					if (!e) rv[c] = A[nx - 1];
					else rv[c] = r + 1 < a.N ? a.PLCP[a.SA[r + 1]] : 0u;
				}
				smask |= (u32) s << c; emask |= (u32) e << c;
				prev = cur;
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
		u32 gl[8], rl[8], fiv[8];
		u64 cp[8];
#pragma unroll
		for (u32 c = 0; c < 8; c++) {
			gl[c] = 0; rl[c] = 0; fiv[c] = 0; cp[c] = 0;
			if (c < E) {
				if ((smask >> c) & 1u) openl = false;
				if ((vmask >> c) & 1u) {
					u32 l = lv[c], r = rv[c];
					if (openl) l = cl < l ? cl : l;
					if (!((closed_r >> c) & 1u)) r = cr < r ? cr : r;
					const u32 i = row[c];
					gl[c] = (l > r ? l : r) + 1;                                  // g, fbg.cpp:1656
					rl[c] = __ldg(a.rowlen + i);
					cp[c] = __ldg(a.start + i);
				}
			}
		}
#pragma unroll
		for (u32 c = 0; c < 8; c++) {
			if (c < E && ((vmask >> c) & 1u)) {
				const u32 i = row[c];
				const u64 gg = (u64) RKS[i] + gl[c];                             // fbg.cpp:1657
				if (gg > rl[c]) fiv[c] = a.tricks_off ? (u32) a.n : __ldg(a.lastcol + i); // 1659-1665
				else fiv[c] = __ldg(a.col + cp[c] + gg - 1);                     // 1666
			}
		}
#pragma unroll
		for (u32 c = 0; c < 8; c++) {
			if (c < E && ((vmask >> c) & 1u)) {
				u32 fi = fiv[c];
				if (a.has_ignore) {                                             // 1669-1670
					const u32 i = row[c];
					u32 ig = a.IG[t * m + i] >> dx;
					u32 ni = ig ? (u32) x + (__ffs(ig) - 1) : a.NIA[t * m + i];
					if (ni < fi) fi = ni;
				}
				fmax = fi > fmax ? fi : fmax;
			}
		}
		// 5c. the rows outside the window: sorted by rank on a small network, one thread per row walks its run
		//     (a few neighbours at most: rows that left the same gap run, the '#' of exhausted rows)
		if (CLUSTER && n_out) {
			u32 onet = 32;
			while (onet < n_out) onet <<= 1;
			for (u32 q = n_out + tg; q < onet; q += tpg) OKEY[ts_sw(q)] = KEY_INVALID;
			__syncthreads();
			group_tile_sort(OKEY, onet, tg, tpg);
			for (u32 o = tg; o < n_out; o += tpg) { // (a group of 256 threads may meet 512 rows)
				const u64 k = OKEY[ts_sw(o)];
				const u32 i = (u32) (k & 0xFFFFFFu);
				u32 l = A[i], r = U32_MAX;
				for (u32 q = o; q > 0; q--) {
					const u64 kp = OKEY[ts_sw(q - 1)], kq = OKEY[ts_sw(q)];
					if ((kp >> 24) + 1 != (kq >> 24)) break;
					const u32 v = A[(u32) (kp & 0xFFFFFFu)];
					l = v < l ? v : l;
				}
				for (u32 q = o;; q++) {
					const u64 rq = OKEY[ts_sw(q)] >> 24;
					const u64 kn = q + 1 < n_out ? OKEY[ts_sw(q + 1)] : KEY_INVALID;
					if (kn != KEY_INVALID && (kn >> 24) == rq + 1) { const u32 v = A[(u32) (kn & 0xFFFFFFu)]; r = v < r ? v : r; }
					else { const u32 v = rq + 1 < a.N ? a.PLCP[a.SA[rq + 1]] : 0u; r = v < r ? v : r; break; }
				}
				const u32 fi = fcol_fi(a, i, l, r, RKS[i], x, t, dx);
				fmax = fi > fmax ? fi : fmax;
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
	a.only_marked = 0;
	a.RC = nullptr; a.AC = nullptr; a.xs0 = 0;
	a.dbg = nullptr;
	const bool debug = getenv("FBG_DEBUG") != nullptr;
	if (debug) { a.dbg = reinterpret_cast<u32 *>(p.scalars + 40); FBG_CUDA(cudaMemsetAsync(a.dbg, 0, 4 * sizeof(u32), d.stream)); }
	size_t smem = (size_t) a.groups * mpad * 16;
	FBG_CUDA(cudaFuncSetAttribute(k2_fcol_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	u64 col_groups = (a.x1 - a.x0 + a.groups - 1) / a.groups;
	if (col_groups == 0) col_groups = 1;
	u64 col_chunks = (col_groups + FC_CHUNK - 1) / FC_CHUNK;
	u64 max_grid = (u64) NUM_SMS * 8;
	(void) col_chunks;
	// many rows: placement by rank instead of a sort (cluster form of the kernel), then the sorting form for the
	// columns that did not fit; the window needs some room beyond the m ranks of the cluster
	const bool cluster = mpad >= 2048 && gc.groups == 1 && gc.E == 8 && p.m * 10 <= (u64) mpad * 9 && !getenv("FBG_FCOL_SORT");
	const size_t smem_c = (size_t) mpad * 10 + (size_t) FC_OCAP * 8;
	if (cluster) FBG_CUDA(cudaFuncSetAttribute(k2_fcol_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem_c));
	// Slabs of columns: ranks and LCPs staged column-major (k2_stage_kernel), then the column kernels on the slab.
	// FBG_FCOL_NO_STAGE=1: the column kernels gather from ISA / PLCP themselves (one slab = the whole range).
	const bool stage = !getenv("FBG_FCOL_NO_STAGE") && a.x1 > a.x0;
	u64 slab = a.x1 - a.x0 ? a.x1 - a.x0 : 1;
	u32 *RC = nullptr, *AC = nullptr;
	u64 mk = 0;
	if (stage) {
		const u64 budget = 1ull << 30; // bytes per staged array
		slab = budget / (4 * p.m) / 256 * 256;
		if (slab < 256) slab = 256;
		if (slab > (1ull << 20)) slab = 1ull << 20; // (grid.y of the staging kernel = slab / 32)
		if (slab > ((a.x1 - a.x0 + 31) / 32 + 1) * 32) slab = ((a.x1 - a.x0 + 31) / 32 + 1) * 32;
		d.ensure(2 * (slab * p.m * 4 + 256) + (1u << 20));
		mk = d.mark();
		RC = d.alloc<u32>(slab * p.m); AC = d.alloc<u32>(slab * p.m);
	}
	const u64 xa = a.x0, xe = a.x1;
	d.begin(K_FCOL);
	for (u64 s0 = stage ? (xa & ~31ull) : xa; s0 < xe; s0 += slab) {
		const u64 s1 = s0 + slab < xe ? s0 + slab : xe;
		a.x0 = s0 < xa ? xa : s0; a.x1 = s1;
		a.RC = RC; a.AC = AC; a.xs0 = s0;
		a.only_marked = 0;
		if (stage) {
			dim3 sg((u32) div_up(p.m, 32), (u32) div_up(s1 - s0, 32));
			k2_stage_kernel<<<sg, 256, 0, d.stream>>>(a, s0, s1, RC, AC);
			d.note_launches(K_FCOL, 1);
		}
		u64 groups_here = (a.x1 - a.x0 + a.groups - 1) / a.groups;
		u64 chunks_here = (groups_here + FC_CHUNK - 1) / FC_CHUNK;
		if (chunks_here == 0) chunks_here = 1;
		const u32 grid_here = (u32) (chunks_here < max_grid ? chunks_here : max_grid);
		if (cluster) {
			k2_fcol_kernel<true><<<grid_here, threads, smem_c, d.stream>>>(a);
			a.only_marked = 1;
			d.note_launches(K_FCOL, 1);
		}
		k2_fcol_kernel<false><<<grid_here, threads, smem, d.stream>>>(a);
		FBG_LAUNCH_CHECK();
		if (s1 < xe) d.note_launches(K_FCOL, 1);
	}
	a.x0 = xa; a.x1 = xe;
	if (cluster && debug) {
		u32 h[4];
		FBG_CUDA(cudaMemcpyAsync(h, a.dbg, sizeof h, cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		fprintf(stderr, "[fbg] K2 cluster form: %u of %llu columns left to the sorting form (no cluster %u, > %u rows outside %u, row next to the window %u)\n",
		        h[0], (unsigned long long) (a.x1 - a.x0), h[1], FC_OCAP, h[2], h[3]);
	}
	d.end((double) p.m * (double) (a.x1 - a.x0) * (stage ? 28.0 : 12.0));
	if (stage) d.release(mk);
	if (sharded) {
		d.begin(K_COMM);
		p.dist.allreduce_max_u32(p.f, p.n, d.stream);
		d.end((double) p.n * 8.0);
	}
	p.have_f = true;
}

} // namespace fbg
