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

// Walks a predecessor chain backwards through chunks staged in shared memory; every lane
// follows the same chain, lane 0 records it.  STEP(b, j): false = stop, else sets the value
// to record and the next j.
template <class STEP> __device__ __forceinline__ u64 dp_chase(const u32 *link, u64 j, u64 nb, u32 *rev, u32 *stage, STEP step) {
	const u32 l = lane_id();
	__threadfence_block();
	for (;;) {
		u64 cb = j + 1 >= DP_CH ? j + 1 - DP_CH : 0;
		__syncwarp();
		for (u32 i = l; i < DP_CH && cb + i <= j; i += 32) stage[i] = __ldcg(&link[cb + i]);
		__syncwarp();
		bool done = false;
		for (;;) {
			u32 rec; u64 nj;
			if (!step(stage[j - cb], rec, nj)) { done = true; break; }
			if (l == 0) rev[nb] = rec;
			nb++;
			j = nj;
			if (j < cb) break;
		}
		if (done) break;
	}
	return nb;
}

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
constexpr int DPW = 16;

// mml/bt have n+1 entries. out[0] = nb, out[1] = mml[n]
__global__ void __launch_bounds__(DPW * 32) k3_elastic_dp_kernel(const u32 *__restrict__ f, const u32 *__restrict__ U, u64 n, u32 *mml, u32 *bt, u32 *rev, u64 *out) {
	__shared__ u32 ring[DP_RING];   // mml
	__shared__ u32 fring[DP_RING];  // f
	__shared__ u32 uch[DP_CH];      // U[base .. base+DP_CH)
	const u32 w = threadIdx.x >> 5, l = lane_id();
	const u32 INV = (u32) n + 1;
	if (threadIdx.x == 0) { mml[0] = 0; bt[0] = 0; ring[0] = 0; }
	u64 base = 0, F = 1;
	bool need_stage = true;
	while (F <= n) {
		if (need_stage || F + DPW > base + DP_CH) {
			base = need_stage ? 0 : F;
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
			ujl = jj <= n ? uch[jj - base] : 0u;
			rdy = jj <= n && (ujl == 0 || (u64) ujl - 1 < F);
		}
		const u32 rb = __ballot_sync(0xffffffffu, rdy);
		const u32 r = (u32) (__ffs(~rb) - 1); // 1 <= r <= DPW: U[j] <= j, so column F is always ready
		if (w < r) {
			const u64 j = F + w;
			const u32 uj = __shfl_sync(0xffffffffu, ujl, w);
			u32 best = U32_MAX, bx = U32_MAX;
			if (uj) {
				u64 x0 = (u64) uj - 1;
				for (;;) {
					const bool inb = x0 >= l;
					const u64 x = inb ? x0 - l : 0;
					const u32 dd = (u32) (j - x);
					u32 fx = 0, mx = 0;
					if (inb) {
						fx = (x + DP_RING >= base + DP_CH) ? fring[x & (DP_RING - 1)] : f[x];
						mx = (x + DP_RING > F + DPW - 1) ? ring[x & (DP_RING - 1)] : ld_vol(&mml[x]);
					}
					const bool valid = inb && (u64) fx + 1 <= j && mx <= (u32) n;
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
						break;
					}
					const u32 tail = __shfl_sync(0xffffffffu, inc, 31);
					const u32 eq = __ballot_sync(0xffffffffu, a == tail);
					if (tail < best) { bx = (u32) (x0 - (u32) (__ffs(eq) - 1)); best = tail; }
					if (x0 < 32) break; // candidate 0 was the last one
					x0 -= 32;
				}
			}
			if (l == 0) {
				if (bx == U32_MAX) best = INV;
				mml[j] = best; bt[j] = bx;
				ring[j & (DP_RING - 1)] = best;
			}
		}
		__syncthreads();
		F += r;
	}
	if (w != 0) return;
	// backtrack, fbg.cpp:2026-2039
	if (l == 0) rev[0] = (u32) n;
	u64 nb = dp_chase(bt, n, 1, rev, uch, [](u32 b, u32 &rec, u64 &nj) {
		if (b == 0 || b == U32_MAX) return false;
		rec = b - 1; nj = b;
		return true;
	});
	if (l == 0) { out[0] = nb; out[1] = ld_vol(&mml[n]); }
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

// fbg.cpp:622-643 on the same block wavefront; lanes take jp = v[j], v[j]-1, ...;
// out[0]=nb, out[1]=s[n-1], out[2]=status
__global__ void __launch_bounds__(DPW * 32) k3_nonelastic_dp_kernel(const u32 *__restrict__ v, u64 n, u32 *s, u32 *prev, u32 *rev, u64 *out) {
	__shared__ u32 ring[DP_RING];
	__shared__ u32 vch[DP_CH];
	const u32 w = threadIdx.x >> 5, l = lane_id();
	u64 F = 0, base = 0;
	bool need_stage = true;
	while (F < n) {
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
			vjl = jj < n ? vch[jj - base] : 0u;
			rdy = jj < n && (vjl > jj || vjl <= F);
		}
		const u32 rb = __ballot_sync(0xffffffffu, rdy);
		const u32 r = (u32) (__ffs(~rb) - 1);
		if (w < r) {
			const u64 j = F + w;
			const u32 vj = __shfl_sync(0xffffffffu, vjl, w);
			u32 sj = (u32) j + 2, pj = (u32) j + 1;
			if (vj <= j) {
				u64 jp0 = vj;
				for (;;) {
					const bool inb = jp0 >= l;
					const u64 jp = inb ? jp0 - l : 0;
					u32 sp = 0;
					if (inb && jp) { u64 x = jp - 1; sp = (x + DP_RING > F + DPW - 1) ? ring[x & (DP_RING - 1)] : ld_vol(&s[x]); }
					const bool use = inb && !(jp != 0 && sp == (u32) jp + 1); // predecessors without a segmentation are skipped
					const u32 len = (u32) (j - jp + 1);
					const u32 cand = use ? (sp > len ? sp : len) : U32_MAX;
					u32 inc = warp_incl_min(cand);
					inc = inc < sj ? inc : sj;
					const bool stop = use && inc == len;
					const bool endz = inb && jp == 0;
					const u32 ms = __ballot_sync(0xffffffffu, stop || endz);
					if (ms) {
						const u32 k = (u32) (__ffs(ms) - 1);
						const u32 m = __shfl_sync(0xffffffffu, inc, k);
						const u32 eq = __ballot_sync(0xffffffffu, cand == m) & ((2u << k) - 1u);
						if (m < sj) { pj = (u32) (jp0 - (u32) (__ffs(eq) - 1)); sj = m; }
						break;
					}
					const u32 tail = __shfl_sync(0xffffffffu, inc, 31);
					const u32 eq = __ballot_sync(0xffffffffu, cand == tail);
					if (tail < sj) { pj = (u32) (jp0 - (u32) (__ffs(eq) - 1)); sj = tail; }
					jp0 -= 32; // no lane reached jp = 0, so jp0 >= 32
				}
			}
			if (l == 0) { s[j] = sj; prev[j] = pj; ring[j & (DP_RING - 1)] = sj; }
		}
		__syncthreads();
		F += r;
	}
	if (w != 0) return;
	__threadfence_block();
	u32 sc = ld_vol(&s[n - 1]);
	if (sc == (u32) n + 1) { if (l == 0) { out[0] = 0; out[1] = sc; out[2] = 1; } return; }
	if (l == 0) rev[0] = (u32) (n - 1);
	u64 nb = dp_chase(prev, n - 1, 1, rev, vch, [](u32 pv, u32 &rec, u64 &nj) {
		if (pv == 0) return false;
		rec = pv - 1; nj = (u64) pv - 1;
		return true;
	});
	if (l == 0) { out[0] = nb; out[1] = sc; out[2] = 0; }
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

void k3_segment(Pipeline &p) {
	Dev &d = p.dev;
	const u64 n = p.n;
	if (!p.bnd) p.bnd = p.dev.palloc<u32>(n + 2);
	if (!p.dp_a) { p.dp_a = p.dev.palloc<u32>(n + 2); p.dp_b = p.dev.palloc<u32>(n + 2); p.dp_v = p.dev.palloc<u32>(n + 2); }
	d.ensure((n + 2) * 8 + (n / SCAN_TILE + 2) * 8 + (1u << 20));
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
		k3_elastic_dp_kernel<<<1, DPW * 32, 0, d.stream>>>(p.f, U, n, p.dp_a, p.dp_b, rev, out);
		k3_reverse_kernel<<<NUM_SMS, 256, 0, d.stream>>>(rev, out, p.bnd);
		FBG_LAUNCH_CHECK();
		d.end((double) n * 24.0);
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
		if (p.opt.gap_limit == 1)
			k3_nonelastic_dp_kernel<<<1, DPW * 32, 0, d.stream>>>(p.dp_v, n, p.dp_a, p.dp_b, rev, out);
		else
			k3_greedy_dp_kernel<<<1, 32, 0, d.stream>>>(p.dp_v, n, p.dp_a, p.dp_b, rev, out);
		k3_reverse_kernel<<<NUM_SMS, 256, 0, d.stream>>>(rev, out, p.bnd);
		FBG_LAUNCH_CHECK();
		d.end((double) n * 24.0);
		d.note_launches(K_DP, 2);
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
