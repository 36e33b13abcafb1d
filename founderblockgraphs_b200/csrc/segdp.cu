// segdp.cu -- K3: segmentation DP + backtrack.
//
// Elastic (fbg.cpp:1940-2039): the reference sweeps j = 1..n with arrival /
// expiry lists; its result has the closed form
//   V_j = {x < j : f[x]+1 <= j, mml[x] <= n},  A = {x in V_j : j-x <= mml[x]},  B = V_j \ A
//   I* = min_{A} mml[x],  S* = j - max B
//   A != {} and (B = {} or I* < S*):  mml[j] = I*, backtrack[j] = max{x in A : mml[x] = I*}
//   else                              mml[j] = S*, backtrack[j] = max B
// Every candidate at distance d = j-x has value >= d, so a downward scan from
// u(j) = max{x : f[x]+1 <= j} that stops once d exceeds the best value seen is
// exact, including the tie rules.  u(j) comes from a scatter-max + prefix-max
// scan.  The recurrence is a chain, so one CTA advances a wavefront: columns whose
// candidates are all final (u(j) < frontier) are evaluated concurrently, one warp each,
// the lanes of a warp sharing the candidate scan of its column.
//
// Non-elastic: v[j] (fbg.cpp:552-611) = max{jp : f_notricks(jp) <= j} through a
// counting scan when f is monotone (always, for gapless rows), else the
// reference's one-directional two-pointer; s/prev DP of fbg.cpp:622-643 as the
// same wavefront; greedy recurrence of fbg.cpp:827-846 for gap_limit != 1.
#include "pipeline.cuh"
#include <cstdlib>
#include <vector>

namespace fbg {

__device__ __forceinline__ u32 ld_vol(const u32 *p) { return *reinterpret_cast<const volatile u32 *>(p); }

// ---------------------------------------------------------------- elastic
__global__ void k3_scatter_u_kernel(const u32 *__restrict__ f, u64 n, u32 *__restrict__ U) {
	u64 x = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (x >= n) return;
	u64 e = (u64) f[x] + 1;
	if (e <= n) atomicMax(&U[e], (u32) x + 1);
}

// The chain only ever looks back about one segment length, so the last DP_RING values
// live in shared memory (slot x & (DP_RING-1)); a slot is trusted while no column that
// aliases it can have been written: x + DP_RING > F + 31.  Older values come from HBM.
// Inputs (U / v, f) are staged DP_CH columns at a time with coalesced loads, so the chain
// itself never waits on an HBM round trip.
constexpr u32 DP_RING = 4096;
constexpr u32 DP_CH = 1024;

// bnd[a] = rev[nb-1-a]; nb read from the device word the DP kernel left
__global__ void k3_reverse_kernel(const u32 *__restrict__ rev, const u64 *__restrict__ out, u32 *__restrict__ bnd) {
	u64 nb = out[0];
	for (u64 a = (u64) blockIdx.x * blockDim.x + threadIdx.x; a < nb; a += (u64) gridDim.x * blockDim.x) bnd[a] = rev[nb - 1 - a];
}

// One CTA of DPW warps.  The DPW columns at the frontier F are examined together: column
// F+w is "ready" when every candidate it can look at is final (u(j) < F), and the ready
// prefix is evaluated at once, one warp per column.  Inside a warp the 32 lanes take 32
// consecutive candidates x = u(j), u(j)-1, ...; a prefix-min over the lanes and two ballots
// replay the reference's sequential scan (stop when the distance exceeds the best value,
// first B candidate ends it and wins ties, largest x wins among equal A candidates).
//
// The chain is broken up by guess-and-verify (k3_segment): the columns are cut into chunks,
// every chunk runs the same wavefront from a local origin t = chunk start - warm-up (values
// relative to a fresh start at t), the running plateau B of the chunks before is folded in
// (guess = max(B, local)), and k3_*_verify_kernel recomputes every column from the guessed
// values of its predecessors, fully in parallel.  A vector that reproduces itself under the
// recurrence IS the solution (induction over j), so a pass without changes proves the guess;
// otherwise the pass is repeated on its own output, and after a few rounds the whole-range
// wavefront (one chunk) is run instead.  Predecessor pointers come from the final pass.
constexpr int DPW = 16;

// Candidate scan of column j over x = uj-1, uj-2, ..., xmin (fbg.cpp:1967-2014 in closed
// form).  ff(x) = f[x], fm(x) = mml[x]; all 32 lanes call; result identical on every lane.
template <class FF, class FM>
__device__ __forceinline__ void elastic_scan(u64 j, u32 uj, u64 xmin, u32 n32, FF ff, FM fm, u32 &best, u32 &bx) {
	const u32 l = lane_id();
	best = U32_MAX; bx = U32_MAX;
	if (uj == 0 || (u64) uj - 1 < xmin) return;
	u64 x0 = (u64) uj - 1;
	for (;;) {
		const bool inb = x0 >= xmin + l;
		const u64 x = inb ? x0 - l : xmin;
		const u32 dd = (u32) (j - x);
		u32 fx = 0, mx = 0;
		if (inb) { fx = ff(x); mx = fm(x); }
		const bool valid = inb && (u64) fx + 1 <= j && mx <= n32;
		const bool isB = valid && dd > mx;            // value j - x
		const u32 a = (valid && !isB) ? mx : U32_MAX; // A candidates: value mml[x]
		const u32 inc = warp_incl_min(a);
		u32 E = __shfl_up_sync(0xffffffffu, inc, 1);  // best value before this lane's candidate
		if (l == 0) E = U32_MAX;
		E = E < best ? E : best;
		const bool stopS = !inb || dd > E;            // the scan ends before this candidate
		const bool stopB = !stopS && isB;             // ... or with it
		const u32 ms = __ballot_sync(0xffffffffu, stopS || stopB);
		const u32 mbits = __ballot_sync(0xffffffffu, stopB);
		if (ms) {
			const u32 k = (u32) (__ffs(ms) - 1);
			if ((mbits >> k) & 1u) { best = __shfl_sync(0xffffffffu, dd, k); bx = (u32) (x0 - k); }
			else {
				const u32 m = __shfl_sync(0xffffffffu, E, k);
				const u32 eq = __ballot_sync(0xffffffffu, a == m) & ((1u << k) - 1u);
				if (m < best) { bx = (u32) (x0 - (u32) (__ffs(eq) - 1)); best = m; }
			}
			return;
		}
		const u32 tail = __shfl_sync(0xffffffffu, inc, 31);
		const u32 eq = __ballot_sync(0xffffffffu, a == tail);
		if (tail < best) { bx = (u32) (x0 - (u32) (__ffs(eq) - 1)); best = tail; }
		if (x0 < xmin + 32) return; // candidate xmin was the last one
		x0 -= 32;
	}
}

// Wavefront over the columns of chunk blockIdx.x: own range [c*C, c*C + C).  A chunk other
// than the first starts from a window of sources [t0, t1] = [c*C - H, c*C - H + H/2], every
// one of them a free block boundary (local value 0), no candidate before t0, and warms up
// over the H/2 columns up to its own range: a single forced boundary would usually sit at a
// column no good segmentation cuts at (f[t] - t can be hundreds of columns in gapped data).
// One chunk (C > n) = the whole DP: then mml/bt are final.  bt may be null (chunked guess).
__global__ void __launch_bounds__(DPW * 32) k3_elastic_dp_kernel(const u32 *__restrict__ f, const u32 *__restrict__ U, u64 n, u64 C, u64 H, u32 chunk0, u32 *mml, u32 *bt, u32 *fail) {
	__shared__ u32 ring[DP_RING];   // mml (local values)
	__shared__ u32 fring[DP_RING];  // f
	__shared__ u32 uch[DP_CH];      // U[base .. base+DP_CH)
	const u32 w = threadIdx.x >> 5, l = lane_id();
	const u32 INV = (u32) n + 1;
	const u64 s = ((u64) blockIdx.x + chunk0) * C; // chunk0: first chunk of this GPU (sharded build)
	const u64 e = s + C < n + 1 ? s + C : n + 1;
	const u64 t = s > H ? s - H : 0;         // first source; nothing before it is looked at
	const u64 t1 = s > H ? t + H / 2 : 0;    // last source
	for (u64 x = t + threadIdx.x; x <= t1; x += DPW * 32) ring[x & (DP_RING - 1)] = 0;
	if (threadIdx.x == 0 && s == 0) { mml[0] = 0; if (bt) bt[0] = 0; }
	u64 base = t1, F = t1 + 1;
	const u64 fst = t1; // f of the columns before it (the sources) is read from global memory
	bool need_stage = true;
	while (F < e) {
		if (need_stage || F + DPW > base + DP_CH) {
			base = need_stage ? t1 : F;
			need_stage = false;
			__syncthreads();
			for (u32 i = threadIdx.x; i < DP_CH; i += DPW * 32) {
				u64 x = base + i;
				uch[i] = x <= n ? U[x] : 0u;
				fring[x & (DP_RING - 1)] = x < n ? f[x] : 0u;
			}
			__syncthreads();
		}
		u32 ujl = 0;
		bool rdy = false;
		if (l < DPW) {
			u64 jj = F + l;
			ujl = jj < e ? uch[jj - base] : 0u;
			rdy = jj < e && (ujl == 0 || (u64) ujl - 1 < F);
		}
		const u32 rb = __ballot_sync(0xffffffffu, rdy);
		const u32 r = (u32) (__ffs(~rb) - 1); // 1 <= r <= DPW: U[j] <= j, so column F is always ready
		if (w < r) {
			const u64 j = F + w;
			const u32 uj = __shfl_sync(0xffffffffu, ujl, w);
			u32 best, bx;
			elastic_scan(j, uj, t, (u32) n,
				[&](u64 x) { return (x >= fst && x + DP_RING >= base + DP_CH) ? fring[x & (DP_RING - 1)] : f[x]; },
				[&](u64 x) {
					if (x + DP_RING > F + DPW - 1) return ring[x & (DP_RING - 1)];
					if (x >= s) return ld_vol(&mml[x]);   // older than the ring: own, already written
					*fail = 1u;                           // warm-up value lost: the guess of this chunk is void
					return 0u;
				}, best, bx);
			if (l == 0) {
				if (bx == U32_MAX) best = INV;
				ring[j & (DP_RING - 1)] = best;
				if (j >= s) { mml[j] = best; if (bt) bt[j] = bx; }
			}
		}
		__syncthreads();
		F += r;
	}
}

// One warp per column: (mml, bt) of column j from the values gin[] of its predecessors.
__global__ void __launch_bounds__(256) k3_elastic_verify_kernel(const u32 *__restrict__ f, const u32 *__restrict__ U, u64 n, const u32 *__restrict__ gin,
                                                                u32 *__restrict__ gout, u32 *__restrict__ bt, u32 *__restrict__ changes, u64 j0, u64 j1) {
	const u64 j = (((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 5) + j0; // columns [j0, j1) of 1..n
	if (j >= j1) return;
	u32 best, bx;
	elastic_scan(j, U[j], 0, (u32) n, [&](u64 x) { return f[x]; }, [&](u64 x) { return gin[x]; }, best, bx);
	if (lane_id() == 0) {
		if (bx == U32_MAX) best = (u32) n + 1;
		gout[j] = best; bt[j] = bx;
		u32 old = gin[j];
		bool same = old == best || (old > (u32) n && best > (u32) n);
		if (!same) atomicAdd(changes, 1u);
	}
}

// per chunk c: q[c] = min of the local values in the w columns up to the next chunk's last
// source (the plateau the next chunk starts from); then B[c] = max_{c' < c} q[c'] (B[0] = 0).
// One WAVE of the whole-range DP: column j only reads columns x < U[j] (U is monotone, U[j] <= j), so once
// the columns below j0 are final, every column with U[j] <= j0 can be computed at once, one warp each.  With
// long blocks (cfg3: no admissible block under ~1400 columns) the whole DP is a few dozen such waves spread
// over all SMs instead of one CTA walking 100 000 columns (17 ms -> about 1 ms).  val is read and written
// in place: reads go to columns < j0, writes to columns >= j0.
__global__ void __launch_bounds__(256) k3_elastic_wave_kernel(const u32 *__restrict__ f, const u32 *__restrict__ U, u64 n, u32 *val, u32 *__restrict__ bt, u64 j0, u64 j1) {
	const u64 j = (((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 5) + j0;
	if (j >= j1) return;
	u32 best, bx;
	elastic_scan(j, U[j], 0, (u32) n, [&](u64 x) { return f[x]; }, [&](u64 x) { return val[x]; }, best, bx);
	if (lane_id() == 0) {
		if (bx == U32_MAX) best = (u32) n + 1;
		val[j] = best; bt[j] = bx;
	}
}

__global__ void __launch_bounds__(1024) k3_plateau_kernel(const u32 *__restrict__ loc, u64 n_cols, u64 C, u64 H, u64 wq, u32 nchunks, u32 limit, bool nonelastic, u32 *__restrict__ B) {
	__shared__ u32 sm[32];
	u32 carry = 0;
	for (u32 c0 = 0; c0 < nchunks; c0 += 1024) {
		u32 c = c0 + threadIdx.x;
		u32 q = 0;
		if (c + 1 < nchunks) {
			u64 tn = (u64) (c + 1) * C - H + H / 2; // last source of chunk c+1, inside chunk c
			u32 mn = U32_MAX;
			for (u64 x = tn > wq ? tn - wq : 0; x <= tn && x < n_cols; x++) {
				if (x < (u64) c * C) continue;
				u32 v = loc[x];
				bool ok = nonelastic ? v != (u32) x + 2 : v <= limit;
				if (ok && v < mn) mn = v;
			}
			q = mn == U32_MAX ? 0u : mn;
		}
		u32 tot;
		u32 ex = block_excl_scan<OpMaxU32>(q, sm, &tot);
		if (c < nchunks) B[c] = ex > carry ? ex : carry;
		carry = tot > carry ? tot : carry;
	}
}

// guess = max(B[chunk], local value); unreachable / invalid markers are kept
__global__ void k3_apply_plateau_kernel(u32 *__restrict__ g, u64 first, u64 n_cols, u64 C, u32 limit, bool nonelastic, const u32 *__restrict__ B) {
	u64 x = first + (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (x >= n_cols) return;
	u32 v = g[x], b = B[x / C];
	bool ok = nonelastic ? v != (u32) x + 2 : v <= limit;
	if (ok && b > v) g[x] = b;
}

// ---- backtrack in parallel
// The chain n -> link(n) -> ... is a list traversal, and one warp walking it through staged
// chunks took 6 ns per column (250 ms on 4*10^7 columns).  Links only point downwards, so:
//   (1) every segment of CH_S columns resolves, for EACH of its columns, where a chain that
//       passes through it leaves the segment (pointer jumping in shared memory);
//   (2) one thread hops from segment to segment (one dependent load per segment) and notes
//       where the chain enters each;
//   (3) the entered segments walk their part of the chain in shared memory and mark it;
//   (4) a scan of the marks places the boundaries in ascending order.
// next(j): elastic  bt[j] (stop at 0 / none);  non-elastic  prev[j] - 1 (stop at prev = 0).
constexpr u32 CH_S = 16384;
constexpr u32 CH_STOP = U32_MAX;
template <bool ELASTIC> __device__ __forceinline__ u32 chain_next(u32 link, u64 j) {
	if (ELASTIC) return (link == 0 || link == U32_MAX || (u64) link >= j) ? CH_STOP : link;
	return (link == 0 || (u64) link - 1 >= j) ? CH_STOP : link - 1;
}

template <bool ELASTIC> __global__ void __launch_bounds__(512) k3_chain_exit_kernel(const u32 *__restrict__ link, u64 ncols, u32 *__restrict__ E) {
	extern __shared__ u32 ch_ex[];
	const u64 s0 = (u64) blockIdx.x * CH_S;
	const u32 len = (u32) (ncols - s0 < CH_S ? ncols - s0 : CH_S);
	for (u32 i = threadIdx.x; i < len; i += blockDim.x) ch_ex[i] = chain_next<ELASTIC>(link[s0 + i], s0 + i);
	__syncthreads();
	for (;;) {
		bool changed = false;
		for (u32 i = threadIdx.x; i < len; i += blockDim.x) {
			const u32 v = ch_ex[i];
			if (v != CH_STOP && (u64) v >= s0) { ch_ex[i] = ch_ex[v - (u32) s0]; changed = true; } // old or new value: both lie on the chain
		}
		if (!__syncthreads_or(changed)) break;
	}
	for (u32 i = threadIdx.x; i < len; i += blockDim.x) E[s0 + i] = ch_ex[i];
}

// ent[seg] = where the chain enters the segment; out[1] = score, out[2] = status
template <bool ELASTIC> __global__ void k3_chain_walk_kernel(const u32 *__restrict__ E, const u32 *__restrict__ val, u64 n, u32 *__restrict__ ent, u64 *out) {
	if (blockIdx.x || threadIdx.x) return;
	u64 cur = ELASTIC ? n : n - 1;
	const u32 sc = val[cur];
	out[1] = sc; out[2] = 0;
	if (!ELASTIC && sc == (u32) n + 1) { out[2] = 1; return; } // no segmentation (fbg.cpp: the DP left its "infinity")
	for (;;) {
		ent[cur / CH_S] = (u32) cur;
		const u32 e = __ldcg(E + cur);
		if (e == CH_STOP) break;
		cur = e;
	}
}

template <bool ELASTIC> __global__ void __launch_bounds__(512) k3_chain_mark_kernel(const u32 *__restrict__ link, u64 ncols, const u32 *__restrict__ ent, u32 *__restrict__ mark) {
	extern __shared__ u32 ch_ex[];
	const u64 s0 = (u64) blockIdx.x * CH_S;
	const u32 len = (u32) (ncols - s0 < CH_S ? ncols - s0 : CH_S);
	const u32 en = ent[blockIdx.x];
	if (en == U32_MAX) return; // the chain does not touch this segment (uniform)
	for (u32 i = threadIdx.x; i < len; i += blockDim.x) ch_ex[i] = chain_next<ELASTIC>(link[s0 + i], s0 + i);
	__syncthreads();
	if (threadIdx.x == 0) {
		u32 j = en;
		for (;;) {
			mark[j] = 1u;
			const u32 nx = ch_ex[j - (u32) s0];
			if (nx == CH_STOP || (u64) nx < s0) break;
			j = nx;
		}
	}
}

// pos = exclusive scan of mark; out[0] = number of boundaries
template <bool ELASTIC> __global__ void k3_chain_emit_kernel(const u32 *__restrict__ mark, const u32 *__restrict__ pos, const u32 *__restrict__ total, u64 ncols, u64 n, u32 *__restrict__ bnd, u64 *out) {
	const u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (j == 0) out[0] = *total;
	if (j >= ncols || !mark[j]) return;
	bnd[pos[j]] = ELASTIC ? (j == n ? (u32) n : (u32) j - 1) : (u32) j; // fbg.cpp:2026-2039: ends of the blocks, the last one = n
}

template <bool ELASTIC> static void dp_backtrack(Pipeline &p, const u32 *link, const u32 *val, u64 *out) {
	Dev &d = p.dev;
	const u64 n = p.n, ncols = ELASTIC ? n + 1 : n;
	const u32 nseg = div_up(ncols, CH_S);
	u32 *E = d.alloc<u32>(ncols + 1), *mark = d.alloc<u32>(ncols + 1), *ent = d.alloc<u32>(nseg + 1);
	u32 *total = reinterpret_cast<u32 *>(p.scalars + 30);
	FBG_CUDA(cudaMemsetAsync(mark, 0, (ncols + 1) * sizeof(u32), d.stream));
	FBG_CUDA(cudaMemsetAsync(ent, 0xFF, ((u64) nseg + 1) * sizeof(u32), d.stream));
	const size_t smem = (size_t) CH_S * sizeof(u32);
	FBG_CUDA(cudaFuncSetAttribute(k3_chain_exit_kernel<ELASTIC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	FBG_CUDA(cudaFuncSetAttribute(k3_chain_mark_kernel<ELASTIC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	d.begin(K_DP);
	k3_chain_exit_kernel<ELASTIC><<<nseg, 512, smem, d.stream>>>(link, ncols, E);
	k3_chain_walk_kernel<ELASTIC><<<1, 32, 0, d.stream>>>(E, val, n, ent, out);
	k3_chain_mark_kernel<ELASTIC><<<nseg, 512, smem, d.stream>>>(link, ncols, ent, mark);
	FBG_LAUNCH_CHECK();
	d.end((double) ncols * 16.0);
	d.note_launches(K_DP, 2);
	u32 *pos = E; // E is dead
	device_scan<OpSumU32, true>(d, mark, pos, ncols, total);
	d.begin(K_DP);
	k3_chain_emit_kernel<ELASTIC><<<div_up(ncols, 256), 256, 0, d.stream>>>(mark, pos, total, ncols, n, p.bnd, out);
	FBG_LAUNCH_CHECK();
	d.end((double) ncols * 8.0);
}

// ------------------------------------------------------------ non-elastic
__global__ void k3_monotone_kernel(const u32 *__restrict__ f, u64 n, u32 *__restrict__ nonmono, u32 *__restrict__ cnt) {
	u64 x = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (x >= n) return;
	u32 fx = f[x];
	if (x + 1 < n && f[x + 1] < fx) *nonmono = 1;
	atomicAdd(&cnt[fx < (u32) n ? fx : (u32) n], 1u);
}

// C = inclusive counts; v[j] = C[j]-1 or j+1
__global__ void k3_v_from_counts_kernel(const u32 *__restrict__ C, u64 n, u32 *__restrict__ v) {
	u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (j < n) v[j] = C[j] ? C[j] - 1 : (u32) j + 1;
}

// reference two-pointer (only when f is not monotone, i.e. gapped rows)
__global__ void k3_v_sequential_kernel(const u32 *__restrict__ f, u64 n, const u32 *__restrict__ nonmono, u32 *__restrict__ v) {
	if (blockIdx.x || threadIdx.x || !*nonmono) return;
	u64 jp = n;
	for (u64 j = n; j-- > 0;) {
		u32 vj = (u32) j + 1;
		for (;;) {
			if (jp < n && f[jp] <= j) { vj = (u32) jp; break; }
			if (jp == 0) break;
			jp--;
		}
		v[j] = vj;
	}
}

// Candidate scan of fbg.cpp:622-643 for column j over jp = vj, vj-1, ..., jmin (jp = jmin
// starts the segmentation: predecessor value 0).  fs(x) = s[x].
template <class FS>
__device__ __forceinline__ void nonelastic_scan(u64 j, u32 vj, u64 jmin, FS fs, u32 &sj, u32 &pj) {
	const u32 l = lane_id();
	sj = (u32) j + 2; pj = (u32) j + 1;
	if (vj > j || vj < jmin) return;
	u64 jp0 = vj;
	for (;;) {
		const bool inb = jp0 >= jmin + l;
		const u64 jp = inb ? jp0 - l : jmin;
		u32 sp = 0;
		if (inb && jp > jmin) sp = fs(jp - 1);
		const bool use = inb && !(jp > jmin && sp == (u32) jp + 1); // predecessors without a segmentation are skipped
		const u32 len = (u32) (j - jp + 1);
		const u32 cand = use ? (sp > len ? sp : len) : U32_MAX;
		u32 inc = warp_incl_min(cand);
		inc = inc < sj ? inc : sj;
		const bool stop = use && inc == len;
		const bool endz = inb && jp == jmin;
		const u32 ms = __ballot_sync(0xffffffffu, stop || endz);
		if (ms) {
			const u32 k = (u32) (__ffs(ms) - 1);
			const u32 m = __shfl_sync(0xffffffffu, inc, k);
			const u32 eq = __ballot_sync(0xffffffffu, cand == m) & ((2u << k) - 1u);
			if (m < sj) { pj = (u32) (jp0 - (u32) (__ffs(eq) - 1)); sj = m; }
			return;
		}
		const u32 tail = __shfl_sync(0xffffffffu, inc, 31);
		const u32 eq = __ballot_sync(0xffffffffu, cand == tail);
		if (tail < sj) { pj = (u32) (jp0 - (u32) (__ffs(eq) - 1)); sj = tail; }
		jp0 -= 32; // no lane reached jmin, so jp0 >= jmin + 32
	}
}

// chunk wavefront of the non-elastic DP; same chunk / origin convention as the elastic one
__global__ void __launch_bounds__(DPW * 32) k3_nonelastic_dp_kernel(const u32 *__restrict__ v, u64 n, u64 C, u64 H, u32 chunk0, u32 *s, u32 *prev, u32 *fail) {
	__shared__ u32 ring[DP_RING];
	__shared__ u32 vch[DP_CH];
	const u32 w = threadIdx.x >> 5, l = lane_id();
	const u64 cs = ((u64) blockIdx.x + chunk0) * C;
	const u64 e = cs + C < n ? cs + C : n;
	// sources: prefixes ending at t0 .. t1 cost nothing (see the elastic kernel); jmin = t0 + 1
	const u64 t0 = cs > H ? cs - H : 0, t1 = cs > H ? t0 + H / 2 : 0;
	const u64 t = cs > H ? t0 + 1 : 0;
	if (cs > H) for (u64 x = t0 + threadIdx.x; x <= t1; x += DPW * 32) ring[x & (DP_RING - 1)] = 0;
	u64 F = cs > H ? t1 + 1 : 0, base = F;
	bool need_stage = true;
	while (F < e) {
		if (need_stage || F + DPW > base + DP_CH) {
			base = F;
			need_stage = false;
			__syncthreads();
			for (u32 i = threadIdx.x; i < DP_CH; i += DPW * 32) vch[i] = base + i < n ? v[base + i] : 0u;
			__syncthreads();
		}
		u32 vjl = 0;
		bool rdy = false;
		if (l < DPW) {
			u64 jj = F + l;
			vjl = jj < e ? vch[jj - base] : 0u;
			rdy = jj < e && (vjl > jj || vjl <= F);
		}
		const u32 rb = __ballot_sync(0xffffffffu, rdy);
		const u32 r = (u32) (__ffs(~rb) - 1);
		if (w < r) {
			const u64 j = F + w;
			const u32 vj = __shfl_sync(0xffffffffu, vjl, w);
			u32 sj, pj;
			nonelastic_scan(j, vj, t, [&](u64 x) {
				if (x + DP_RING > F + DPW - 1) return ring[x & (DP_RING - 1)];
				if (x >= cs) return ld_vol(&s[x]);
				*fail = 1u;
				return 0u;
			}, sj, pj);
			if (l == 0) {
				ring[j & (DP_RING - 1)] = sj;
				if (j >= cs) { s[j] = sj; if (prev) prev[j] = pj; }
			}
		}
		__syncthreads();
		F += r;
	}
}

__global__ void __launch_bounds__(256) k3_nonelastic_verify_kernel(const u32 *__restrict__ v, u64 n, const u32 *__restrict__ gin, u32 *__restrict__ gout, u32 *__restrict__ prev,
                                                                   u32 *__restrict__ changes, u64 j0, u64 j1) {
	const u64 j = (((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 5) + j0; // columns [j0, j1) of 0..n-1
	if (j >= j1) return;
	u32 sj, pj;
	nonelastic_scan(j, v[j], 0, [&](u64 x) { return gin[x]; }, sj, pj);
	if (lane_id() == 0) {
		gout[j] = sj; prev[j] = pj;
		if (gin[j] != sj) atomicAdd(changes, 1u);
	}
}

// fbg.cpp:827-865 (greedy, not optimal by design); sequential by construction
__global__ void k3_greedy_dp_kernel(const u32 *__restrict__ v, u64 n, u32 *s, u32 *prev, u32 *rev, u64 *out) {
	if (blockIdx.x || threadIdx.x) return;
	const u64 INV = n + 1;
	// size_t arithmetic of the reference, including its wrap-around on prev = n+1
	u64 sp = INV, pp = INV; // s[j-1], prev[j-1]
	s[0] = (u32) INV; prev[0] = (u32) INV;
	for (u64 j = 1; j < n; j++) {
		u64 jp = v[j];
		u64 sj = INV, pj = INV;
		if (jp > j) {
			// keep defaults
		} else if (jp == 0) {
			sj = j + 1; pj = 0;
		} else {
			u64 a0 = s[jp - 1], a1 = j - jp + 1;
			u64 a = a0 > a1 ? a0 : a1;
			u64 b1 = j - pp + 1; // wraps like size_t when pp = n+1
			u64 b = sp > b1 ? sp : b1;
			if (a < b) { sj = a; pj = jp; }
			else { sj = b; pj = pp; }
		}
		// the reference stores size_t; values above 2^32 only arise from the wrap and
		// behave as "huge" in later comparisons: clamp to U32_MAX-1 keeps that order
		s[j] = sj > 0xFFFFFFFEull ? 0xFFFFFFFEu : (u32) sj;
		prev[j] = (u32) pj;
		sp = sj; pp = pj;
	}
	u64 sc = n == 1 ? INV : sp;
	out[1] = sc;
	if (sc == INV) { out[0] = 0; out[2] = 1; return; }
	u64 nb = 0, j = n - 1;
	rev[nb++] = (u32) j;
	while (prev[j] != 0) { rev[nb++] = prev[j] - 1; j = prev[j] - 1; }
	out[0] = nb; out[2] = 0;
}

// sum over columns of the shortest admissible last block (j - u(j) resp. j - v[j] + 1): its mean
// says how long a warm-up has to be before chunking is worth trying
__global__ void __launch_bounds__(256) k3_min_block_sum_kernel(const u32 *__restrict__ UV, u64 ncols, bool elastic, u64 *__restrict__ sum) {
	u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	u64 v = 0;
	if (j < ncols) {
		u32 q = UV[j];
		if (elastic) v = q ? j - ((u64) q - 1) : 0;
		else v = q <= j ? j - q + 1 : 0;
	}
	v = warp_reduce_sum(v);
	if (lane_id() == 0 && v) atomicAdd((unsigned long long *) sum, (unsigned long long) v);
}

// Guess-and-verify driver (see the comment above DPW).  `elastic`: columns 0..n (mml), else
// 0..n-1 (s).  Leaves values in val, pointers in ptr.
static void dp_solve(Pipeline &p, bool elastic, const u32 *U_or_v, u32 *val, u32 *ptr) {
	Dev &d = p.dev;
	const u64 n = p.n;
	const u64 ncols = elastic ? n + 1 : n;
	const bool debug = getenv("FBG_DEBUG") != nullptr;
	u32 *flags = reinterpret_cast<u32 *>(p.scalars + 26); // [0] warm-up lost, [1] changes of the last pass
	u32 *h_flags = reinterpret_cast<u32 *>(p.h_scalars + 26);
	auto whole = [&]() {
		if (elastic && !getenv("FBG_DP_NO_WAVES")) {
			// waves of mutually independent columns (k3_elastic_wave_kernel), if they come out wide
			std::vector<u32> hu(n + 1);
			FBG_CUDA(cudaMemcpyAsync(hu.data(), U_or_v, (n + 1) * sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
			FBG_CUDA(cudaStreamSynchronize(d.stream));
			std::vector<u64> cuts; // wave k = columns [cuts[k], cuts[k+1]) of 1..n
			u64 j0 = 1;
			while (j0 <= n && cuts.size() <= 4096) {
				cuts.push_back(j0);
				u64 lo = j0 + 1, hi = n + 1; // first j > j0 with U[j] > j0
				while (lo < hi) { const u64 mid = (lo + hi) >> 1; if (hu[mid] > j0) hi = mid; else lo = mid + 1; }
				j0 = lo;
			}
			if (j0 > n && (cuts.size() * 64 <= n || getenv("FBG_DP_WAVES"))) { // (the variable: test hook, waves on small inputs too)
				cuts.push_back(n + 1);
				FBG_CUDA(cudaMemsetAsync(val, 0, sizeof(u32), d.stream)); // mml[0] = 0
				FBG_CUDA(cudaMemsetAsync(ptr, 0, sizeof(u32), d.stream));
				for (size_t k = 0; k + 1 < cuts.size(); k++)
					k3_elastic_wave_kernel<<<div_up((cuts[k + 1] - cuts[k]) * 32, 256), 256, 0, d.stream>>>(p.f, U_or_v, n, val, ptr, cuts[k], cuts[k + 1]);
				FBG_LAUNCH_CHECK();
				d.note_launches(K_DP, cuts.size() - 1);
				if (debug) fprintf(stderr, "[fbg] K3 whole-range DP in %zu waves of independent columns\n", cuts.size() - 1);
				return;
			}
		}
		if (elastic) k3_elastic_dp_kernel<<<1, DPW * 32, 0, d.stream>>>(p.f, U_or_v, n, ncols + 1, 0, 0u, val, ptr, flags);
		else k3_nonelastic_dp_kernel<<<1, DPW * 32, 0, d.stream>>>(U_or_v, n, ncols + 1, 0, 0u, val, ptr, flags);
		FBG_LAUNCH_CHECK();
	};
	// The warm-up must outlast the plateau (the optimal segment length so far), which is only
	// known afterwards; the mean shortest admissible block is a cheap indicator of it.
	u64 C = p.dp_chunk ? p.dp_chunk : 2048, H = p.dp_chunk ? p.dp_warm : 1024;
	if (!p.dp_chunk && ncols > C) {
		u64 *sum = p.scalars + 28;
		FBG_CUDA(cudaMemsetAsync(sum, 0, sizeof(u64), d.stream));
		k3_min_block_sum_kernel<<<div_up(ncols, 256), 256, 0, d.stream>>>(U_or_v, ncols, elastic, sum);
		FBG_LAUNCH_CHECK();
		FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 28, sum, sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
		FBG_CUDA(cudaStreamSynchronize(d.stream));
		const u64 mean = p.h_scalars[28] / ncols + 1;
		if (mean * 16 > 1024) { C = 8192; H = 4096; }
		if (mean * 16 > 4096) C = ncols + 1; // blocks of hundreds of columns: few chunks would be left, run the wavefront once
		if (debug) fprintf(stderr, "[fbg] K3 mean shortest block %llu columns -> chunk %llu, warm-up %llu\n", (unsigned long long) mean, (unsigned long long) C, (unsigned long long) H);
		d.note_launches(K_DP, 1);
	}
	if (H > C) H = C;
	if (H < 1) H = 1; // at least one source column in front of a chunk
	u64 mk = d.mark();
	u32 *g2 = d.alloc<u32>(ncols + 1);
	u32 *B = d.alloc<u32>(ncols / 16 + 2);
	bool proven = false;
	p.dp_passes = 0;
	std::vector<u32> dbg_guess; u64 dbgC = 0;
	for (int attempt = 0; attempt < 1 && !proven; attempt++) {
		const u32 nchunks = div_up(ncols, C);
		if (nchunks <= 1) break;
		const u64 wq = C - H / 2 < 64 ? C - H / 2 : 64;
		FBG_CUDA(cudaMemsetAsync(flags, 0, 2 * sizeof(u32), d.stream));
		// multi-GPU build: the chunks are independent, every GPU guesses its share of them and the
		// guesses (each column written by one GPU, zero elsewhere) are merged by all-reduce(max)
		u64 c0 = 0, c1 = nchunks;
		const bool sharded = p.dist.world > 1;
		if (sharded) {
			plan_columns(nchunks, p.dist.world, p.dist.rank, &c0, &c1);
			FBG_CUDA(cudaMemsetAsync(val, 0, ncols * sizeof(u32), d.stream));
		}
		if (c1 > c0) {
			if (elastic) k3_elastic_dp_kernel<<<(u32) (c1 - c0), DPW * 32, 0, d.stream>>>(p.f, U_or_v, n, C, H, (u32) c0, val, nullptr, flags);
			else k3_nonelastic_dp_kernel<<<(u32) (c1 - c0), DPW * 32, 0, d.stream>>>(U_or_v, n, C, H, (u32) c0, val, nullptr, flags);
		}
		if (sharded) {
			p.dist.allreduce_max_u32(val, ncols, d.stream);
			p.dist.allreduce_max_u32(flags, 1, d.stream);
		}
		k3_plateau_kernel<<<1, 1024, 0, d.stream>>>(val, ncols, C, H, wq, nchunks, (u32) n, !elastic, B);
		k3_apply_plateau_kernel<<<div_up(ncols - C, 256), 256, 0, d.stream>>>(val, C, ncols, C, (u32) n, !elastic, B);
		FBG_LAUNCH_CHECK();
		d.note_launches(K_DP, 3);
		if (debug) { dbg_guess.resize(ncols); FBG_CUDA(cudaMemcpyAsync(dbg_guess.data(), val, ncols * sizeof(u32), cudaMemcpyDeviceToHost, d.stream)); FBG_CUDA(cudaStreamSynchronize(d.stream)); dbgC = C; }
		u32 *cur = val, *nxt = g2;
		// the verify pass is one independent recomputation per column: every GPU takes a column range
		const u64 jlo = elastic ? 1 : 0, jn = n;
		u64 j0 = jlo, j1 = jlo + jn;
		if (sharded) { plan_columns(jn, p.dist.world, p.dist.rank, &j0, &j1); j0 += jlo; j1 += jlo; }
		const u32 vgrid = div_up((j1 - j0) * 32, 256);
		for (int pass = 0; pass < 6; pass++) {
			FBG_CUDA(cudaMemsetAsync(flags + 1, 0, sizeof(u32), d.stream));
			if (sharded) {
				FBG_CUDA(cudaMemsetAsync(nxt, 0, (ncols + 1) * sizeof(u32), d.stream));
				FBG_CUDA(cudaMemsetAsync(ptr, 0, ncols * sizeof(u32), d.stream));
			}
			if (elastic) {
				FBG_CUDA(cudaMemsetAsync(nxt, 0, sizeof(u32), d.stream)); // mml[0] = 0
				if (j1 > j0) k3_elastic_verify_kernel<<<vgrid, 256, 0, d.stream>>>(p.f, U_or_v, n, cur, nxt, ptr, flags + 1, j0, j1);
			} else {
				if (j1 > j0) k3_nonelastic_verify_kernel<<<vgrid, 256, 0, d.stream>>>(U_or_v, n, cur, nxt, ptr, flags + 1, j0, j1);
			}
			if (sharded) {
				p.dist.allreduce_max_u32(nxt, ncols, d.stream);
				p.dist.allreduce_max_u32(ptr, ncols, d.stream);
				p.dist.allreduce_max_u32(flags + 1, 1, d.stream); // "some GPU saw a change" is all the host asks
			}
			FBG_LAUNCH_CHECK();
			FBG_CUDA(cudaMemcpyAsync(h_flags, flags, 2 * sizeof(u32), cudaMemcpyDeviceToHost, d.stream));
			FBG_CUDA(cudaStreamSynchronize(d.stream));
			d.note_launches(K_DP, 1);
			p.dp_passes++;
			if (debug) fprintf(stderr, "[fbg] K3 attempt %d (chunk %llu, warm-up %llu, %u chunks) pass %d: %u columns changed%s\n", attempt,
			                   (unsigned long long) C, (unsigned long long) H, nchunks, pass, h_flags[1], h_flags[0] ? ", a warm-up window was lost" : "");
			if (h_flags[0]) break; // a chunk lost its warm-up window: no guess to verify
			u32 *tmp = cur; cur = nxt; nxt = tmp;
			if (h_flags[1] == 0) { proven = true; break; }
			if ((u64) h_flags[1] * 64 > ncols) break; // the guess is far off (plateau longer than the warm-up): do not iterate
		}
		if (proven && cur != val) FBG_CUDA(cudaMemcpyAsync(val, cur, ncols * sizeof(u32), cudaMemcpyDeviceToDevice, d.stream));
	}
	if (!proven) { if (p.dp_passes) p.dp_passes = -p.dp_passes; whole(); FBG_CUDA(cudaStreamSynchronize(d.stream)); }
	if (!proven && debug && !dbg_guess.empty()) {
		std::vector<u32> tv(ncols);
		FBG_CUDA(cudaMemcpy(tv.data(), val, ncols * sizeof(u32), cudaMemcpyDeviceToHost));
		int shown = 0;
		for (u64 x = 0; x < ncols && shown < 24; x++) {
			if (tv[x] == dbg_guess[x]) continue;
			u64 y = x;
			while (y < ncols && tv[y] != dbg_guess[y]) y++;
			fprintf(stderr, "[fbg] K3 guess off in columns [%llu, %llu) (chunk %llu, offset %llu): guess %u true %u; before: true %u\n", (unsigned long long) x,
			        (unsigned long long) y, (unsigned long long) (x / dbgC), (unsigned long long) (x % dbgC), dbg_guess[x], tv[x], x ? tv[x - 1] : 0);
			x = y; shown++;
		}
	}
	else if (elastic) FBG_CUDA(cudaMemsetAsync(ptr, 0, sizeof(u32), d.stream)); // bt[0] = 0
	d.release(mk);
}

void k3_segment(Pipeline &p) {
	Dev &d = p.dev;
	const u64 n = p.n;
	if (!p.bnd) p.bnd = p.dev.palloc<u32>(n + 2);
	if (!p.dp_a) { p.dp_a = p.dev.palloc<u32>(n + 2); p.dp_b = p.dev.palloc<u32>(n + 2); p.dp_v = p.dev.palloc<u32>(n + 2); }
	d.ensure((n + 2) * 20 + (n / 16 + 4) * 4 + (n / CH_S + 4) * 4 + (n / SCAN_TILE + 2) * 24 + (1u << 20));
	u64 mk = d.mark();
	u32 *rev = d.alloc<u32>(n + 2);
	u64 *out = p.scalars + 16;
	FBG_CUDA(cudaMemsetAsync(out, 0, 4 * sizeof(u64), d.stream));
	u32 g = div_up(n, 256);
	if (p.opt.elastic) {
		u32 *U = d.alloc<u32>(n + 2);
		FBG_CUDA(cudaMemsetAsync(U, 0, (n + 2) * sizeof(u32), d.stream));
		d.begin(K_DP);
		k3_scatter_u_kernel<<<g, 256, 0, d.stream>>>(p.f, n, U);
		FBG_LAUNCH_CHECK();
		d.end((double) n * 8.0);
		device_scan<OpMaxU32, false>(d, U, U, n + 1);
		d.begin(K_DP);
		dp_solve(p, true, U, p.dp_a, p.dp_b);
		FBG_LAUNCH_CHECK();
		d.end((double) n * 24.0);
		dp_backtrack<true>(p, p.dp_b, p.dp_a, out);
	} else {
		u32 *cnt = d.alloc<u32>(n + 2);
		u32 *nonmono = reinterpret_cast<u32 *>(p.scalars + 24);
		FBG_CUDA(cudaMemsetAsync(cnt, 0, (n + 2) * sizeof(u32), d.stream));
		FBG_CUDA(cudaMemsetAsync(nonmono, 0, sizeof(u64), d.stream));
		d.begin(K_DP);
		k3_monotone_kernel<<<g, 256, 0, d.stream>>>(p.f, n, nonmono, cnt);
		FBG_LAUNCH_CHECK();
		d.end((double) n * 12.0);
		device_scan<OpSumU32, false>(d, cnt, cnt, n + 1);
		d.begin(K_DP);
		k3_v_from_counts_kernel<<<g, 256, 0, d.stream>>>(cnt, n, p.dp_v);
		k3_v_sequential_kernel<<<1, 32, 0, d.stream>>>(p.f, n, nonmono, p.dp_v);
		if (p.opt.gap_limit == 1) {
			dp_solve(p, false, p.dp_v, p.dp_a, p.dp_b);
			FBG_LAUNCH_CHECK();
			d.end((double) n * 24.0);
			d.note_launches(K_DP, 2);
			dp_backtrack<false>(p, p.dp_b, p.dp_a, out);
		} else {
			k3_greedy_dp_kernel<<<1, 32, 0, d.stream>>>(p.dp_v, n, p.dp_a, p.dp_b, rev, out);
			k3_reverse_kernel<<<NUM_SMS, 256, 0, d.stream>>>(rev, out, p.bnd);
			FBG_LAUNCH_CHECK();
			d.end((double) n * 24.0);
			d.note_launches(K_DP, 4);
		}
	}
	FBG_CUDA(cudaMemcpyAsync(p.h_scalars + 16, out, 4 * sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	p.nb = p.h_scalars[16];
	p.score = p.h_scalars[17];
	p.seg_status = (int) p.h_scalars[18];
	p.have_seg = true;
	d.release(mk);
}

} // namespace fbg
