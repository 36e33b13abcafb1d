// pack.cu -- K0: MSA bytes -> bitmaps, per-row rank tables, gap-stripped text.
//
// Replaces, on the device, the reference's text writer (fbg.cpp:372-386) and the
// per-row bit-vector / rank / select preprocessing of
// segment_elastic_minmaxlength (fbg.cpp:1845-1917):
//   indexedrows[i] + rank_support  ->  NG bitmap + RK table (rank at 32-col tiles)
//   rowsignore[i] + rank/select    ->  IG bitmap + NIA table (next ignore column)
//   indexedrows_ss[i].select(r)    ->  col[start_i + r - 1]
//   the two "MSA contains ..." counters (fbg.cpp:1858-1859)
#include "pipeline.cuh"

namespace fbg {

constexpr int CH = 128; // 32-column tiles per scan chunk

__device__ __forceinline__ void masks32(const uint4 a, const uint4 b, const IgnoreSet &ign, u32 &ng, u32 &ig) {
	u32 w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
	ng = 0; ig = 0;
#pragma unroll
	for (int q = 0; q < 8; q++) {
#pragma unroll
		for (int s = 0; s < 4; s++) {
			u32 c = (w[q] >> (8 * s)) & 0xFFu;
			u32 bit = 1u << (q * 4 + s);
			if (c != 0x2Du) ng |= bit;
			if ((ign.bits[c >> 5] >> (c & 31u)) & 1u) ig |= bit;
		}
	}
}

// grid (ceil(W/32), ceil(m/32)), block (32,8): 32 rows x 32 tiles per CTA, transposed through smem
__global__ void __launch_bounds__(256) k0_mask_kernel(const u8 *__restrict__ msa, u64 pitch, u64 m, u64 n, u64 W, IgnoreSet ign, u32 *__restrict__ NG, u32 *__restrict__ IG) {
	__shared__ u32 sng[32][33], sig[32][33];
	const u32 tx = threadIdx.x, ty = threadIdx.y;
	const u64 t0 = (u64) blockIdx.x * 32, r0 = (u64) blockIdx.y * 32;
	for (u32 rr = ty; rr < 32; rr += 8) {
		u64 row = r0 + rr, t = t0 + tx;
		u32 ng = 0, ig = 0;
		if (row < m && t < W) {
			const u8 *src = msa + row * pitch + t * 32;
			uint4 a = ld_nc_u4(src), b = ld_nc_u4(src + 16);
			masks32(a, b, ign, ng, ig);
			// columns >= n are padding ('-'): never non-gap; mask ignore bits too
			u64 rem = n - t * 32;
			if (rem < 32) { u32 keep = (1u << rem) - 1u; ng &= keep; ig &= keep; }
		}
		sng[rr][tx] = ng; sig[rr][tx] = ig;
	}
	__syncthreads();
	for (u32 ww = ty; ww < 32; ww += 8) {
		u64 row = r0 + tx, t = t0 + ww;
		if (row < m && t < W) { NG[t * m + row] = sng[tx][ww]; IG[t * m + row] = sig[tx][ww]; }
	}
}

// thread per (chunk, row); row fastest so that tile-major reads coalesce
__global__ void k0_chunk_sum_kernel(const u32 *__restrict__ NG, const u32 *__restrict__ IG, u64 m, u64 W, u64 NC,
                                    u32 *__restrict__ CS, u32 *__restrict__ CL, u32 *__restrict__ CI, u32 *__restrict__ CG) {
	u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (gid >= NC * m) return;
	u64 c = gid / m, i = gid % m;
	u64 t1 = (c + 1) * CH < W ? (c + 1) * CH : W;
	u32 s = 0, last = 0, firsti = U32_MAX, g = 0;
	for (u64 t = c * CH; t < t1; t++) {
		u32 ng = NG[t * m + i], ig = IG[t * m + i];
		s += __popc(ng); g += __popc(ig);
		if (ng) last = (u32) (t * 32) + (32 - __clz(ng));         // last non-gap column + 1
		if (ig && firsti == U32_MAX) firsti = (u32) (t * 32) + (__ffs(ig) - 1);
	}
	CS[gid] = s; CL[gid] = last; CI[gid] = firsti; CG[gid] = g;
}

// thread per row: exclusive prefix of chunk sums, suffix "next ignore column", row totals
__global__ void k0_chunk_scan_kernel(u64 m, u64 NC, u32 *__restrict__ CS, const u32 *__restrict__ CL, u32 *__restrict__ CI, const u32 *__restrict__ CG,
                                     u32 *__restrict__ rowlen, u32 *__restrict__ lastcol, u64 *__restrict__ totals) {
	u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	u64 nong = 0, ign = 0;
	if (i < m) {
		u32 run = 0, last = 0;
		for (u64 c = 0; c < NC; c++) {
			u32 s = CS[c * m + i];
			CS[c * m + i] = run; run += s;
			u32 l = CL[c * m + i]; if (l) last = l;
			ign += CG[c * m + i];
		}
		rowlen[i] = run; lastcol[i] = last ? last - 1 : 0;
		nong = run;
		u32 nxt = U32_MAX;
		for (u64 c = NC; c-- > 0;) { u32 v = CI[c * m + i]; CI[c * m + i] = nxt; if (v != U32_MAX) nxt = v; }
	}
	// block reduction then one atomic per block
	nong = warp_reduce_sum(nong); ign = warp_reduce_sum(ign);
	if (lane_id() == 0 && (nong | ign)) {
		atomicAdd((unsigned long long *) &totals[0], (unsigned long long) nong);
		atomicAdd((unsigned long long *) &totals[1], (unsigned long long) ign);
	}
}

__global__ void k0_chunk_apply_kernel(const u32 *__restrict__ NG, const u32 *__restrict__ IG, u64 m, u64 W, u64 NC,
                                      const u32 *__restrict__ CSX, const u32 *__restrict__ CNI, u32 *__restrict__ RK, u32 *__restrict__ NIA) {
	u64 gid = (u64) blockIdx.x * blockDim.x + threadIdx.x;
	if (gid >= NC * m) return;
	u64 c = gid / m, i = gid % m;
	u64 t1 = (c + 1) * CH < W ? (c + 1) * CH : W;
	u32 run = CSX[gid];
	for (u64 t = c * CH; t < t1; t++) { RK[t * m + i] = run; run += __popc(NG[t * m + i]); }
	u32 nxt = CNI[gid];
	for (u64 t = t1; t-- > c * CH;) {
		NIA[t * m + i] = nxt;
		u32 ig = IG[t * m + i];
		if (ig) nxt = (u32) (t * 32) + (__ffs(ig) - 1);
	}
}

// start[i] = sum_{i'<i} (rowlen[i'] + 1); start[m] = total; N = start[m] + 1.  One CTA,
// chunks of 1024 rows scanned in shared memory.
__global__ void __launch_bounds__(1024) k0_starts_kernel(const u32 *__restrict__ rowlen, u64 m, u32 *__restrict__ start, u64 *__restrict__ totals) {
	__shared__ u64 sm[32];
	u64 carry = 0;
	for (u64 base = 0; base < m; base += 1024) {
		u64 i = base + threadIdx.x;
		u64 v = i < m ? (u64) rowlen[i] + 1 : 0;
		u64 tot;
		u64 ex = block_excl_scan<OpSumU64>(v, sm, &tot);
		if (i < m) start[i] = (u32) (carry + ex);
		carry += tot;
	}
	if (threadIdx.x == 0) { start[m] = (u32) carry; totals[2] = carry + 1; }
}

// one thread per (row, 32-column tile), one warp per 1024 columns of a row: the lanes compact their non-gap bytes and
// the columns of these into the warp's shared-memory strip at (text position - the warp's first text position),
// the warp writes the strip out with consecutive lanes on consecutive addresses.  (The first form let every lane
// write its own 32 bytes / 32 words straight to T and col: 64 store instructions per lane, each touching 32
// different sectors per warp -- 4.8 ms on cfg3 for 3 GB of traffic.)
__global__ void __launch_bounds__(256) k0_text_kernel(const u8 *__restrict__ msa, u64 pitch, u64 m, u64 n, u64 W, u64 wb,
                                                      const u32 *__restrict__ RK, const u32 *__restrict__ start, const u32 *__restrict__ rowlen,
                                                      u8 *__restrict__ T, u32 *__restrict__ col, u64 N) {
	__shared__ u8 sT[8][1024];
	__shared__ u32 sC[8][1024];
	const u32 wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
	u64 row = blockIdx.x / wb;
	u64 t = (blockIdx.x % wb) * 256 + threadIdx.x;
	if (t - lane >= W) return; // the whole warp lies behind the last tile
	const bool act = t < W;
	u32 w[8] = {0x2D2D2D2Du, 0x2D2D2D2Du, 0x2D2D2D2Du, 0x2D2D2D2Du, 0x2D2D2D2Du, 0x2D2D2D2Du, 0x2D2D2D2Du, 0x2D2D2D2Du};
	u64 p = 0;
	u64 lim = 0;
	if (act) {
		const u8 *src = msa + row * pitch + t * 32;
		uint4 a = ld_nc_u4(src), b = ld_nc_u4(src + 16);
		w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
		p = (u64) start[row] + RK[t * m + row];
		lim = n - t * 32 < 32 ? n - t * 32 : 32;
	}
	const u64 p0 = __shfl_sync(0xffffffffu, p, 0);
	u32 o = act ? (u32) (p - p0) : 0u;
#pragma unroll
	for (int q = 0; q < 8; q++) {
#pragma unroll
		for (int s = 0; s < 4; s++) {
			u32 c = (w[q] >> (8 * s)) & 0xFFu;
			u32 x = q * 4 + s;
			if (x < lim && c != 0x2Du) { sT[wid][o] = (u8) c; sC[wid][o] = (u32) (t * 32) + x; o++; }
		}
	}
	const u32 total = __reduce_max_sync(0xffffffffu, o); // (text positions grow with the lane)
	__syncwarp();
	for (u32 j = lane; j < total; j += 32) { T[p0 + j] = sT[wid][j]; col[p0 + j] = sC[wid][j]; }
	if (t == W - 1) {
		u64 e = (u64) start[row] + rowlen[row];
		T[e] = (u8) '#'; col[e] = (u32) n;
		if (row == m - 1) { T[N - 1] = 0; col[N - 1] = (u32) n + 1; }
	}
}

// gapless MSA: the text is the rows themselves with a '#' after each, position p = i (n + 1) + x.
// One thread per four text positions, coalesced on both sides.
__global__ void __launch_bounds__(256) k0_text_gapless_kernel(const u8 *__restrict__ msa, u64 pitch, u64 m, u64 n, u8 *__restrict__ T, u32 *__restrict__ col, u64 N) {
	const u64 p0 = ((u64) blockIdx.x * blockDim.x + threadIdx.x) * 4;
	if (p0 >= N) return;
	u64 i = p0 / (n + 1), x = p0 % (n + 1);
	u32 tv = 0, cv[4];
#pragma unroll
	for (int q = 0; q < 4; q++) {
		u32 c = 0, cl = (u32) n + 1; // past the end / the final sentinel
		if (p0 + q < N - 1) {
			if (x < n) { c = msa[i * pitch + x]; cl = (u32) x; x++; }
			else { c = (u32) '#'; cl = (u32) n; x = 0; i++; }
		}
		tv |= c << (8 * q);
		cv[q] = cl;
	}
	// T and col are 256-byte aligned and hold at least N + 4 entries
	*reinterpret_cast<u32 *>(T + p0) = tv;
	*reinterpret_cast<uint4 *>(col + p0) = make_uint4(cv[0], cv[1], cv[2], cv[3]);
}

static void alloc_msa(Pipeline &p, u64 m, u64 n) {
	p.free_all();
	p.m = m; p.n = n; p.pitch = round_up(n, 32); p.W = p.pitch / 32;
	if (persistent_need(m, n) > p.dev.pers_cap) FBG_CUDA(cudaStreamSynchronize(p.dev.stream)); // about to free the old arena
	p.dev.reserve_pers(persistent_need(m, n));
	p.msa = p.dev.palloc<u8>(m * p.pitch);
	if (!p.scalars) FBG_CUDA(cudaMalloc(&p.scalars, 64 * sizeof(u64)));
	if (!p.h_scalars) FBG_CUDA(cudaMallocHost(&p.h_scalars, 64 * sizeof(u64)));
	if (p.pitch != n) FBG_CUDA(cudaMemsetAsync(p.msa, 0x2D, m * p.pitch, p.dev.stream));
}

void k0_upload_host(Pipeline &p, const u8 *rows, u64 m, u64 n) {
	alloc_msa(p, m, n);
	FBG_CUDA(cudaMemcpy2DAsync(p.msa, p.pitch, rows, n, n, m, cudaMemcpyHostToDevice, p.dev.stream));
	p.have_msa = true;
}

void k0_upload_device(Pipeline &p, const u8 *drows, u64 m, u64 n) {
	alloc_msa(p, m, n);
	FBG_CUDA(cudaMemcpy2DAsync(p.msa, p.pitch, drows, n, n, m, cudaMemcpyDeviceToDevice, p.dev.stream));
	p.have_msa = true;
}

void k0_pack(Pipeline &p) {
	Dev &d = p.dev;
	const u64 m = p.m, n = p.n, W = p.W;
	const u64 NC = (W + CH - 1) / CH;
	if (!p.NG) { // (packing the same MSA again, e.g. after the ignore characters changed, re-uses the arrays)
		p.NG = p.dev.palloc<u32>(W * m); p.IG = p.dev.palloc<u32>(W * m);
		p.RK = p.dev.palloc<u32>(W * m); p.NIA = p.dev.palloc<u32>(W * m);
		p.start = p.dev.palloc<u32>(m + 1); p.rowlen = p.dev.palloc<u32>(m); p.lastcol = p.dev.palloc<u32>(m);
	}
	IgnoreSet ign;
	for (int q = 0; q < 8; q++) ign.bits[q] = p.opt.ignore_bits[q];

	d.ensure(NC * m * 16 + (1u << 20));
	u64 mk = d.mark();
	u32 *CS = d.alloc<u32>(NC * m), *CL = d.alloc<u32>(NC * m), *CI = d.alloc<u32>(NC * m), *CG = d.alloc<u32>(NC * m);
	FBG_CUDA(cudaMemsetAsync(p.scalars, 0, 8 * sizeof(u64), d.stream));
	d.begin(K_PACK);
	dim3 g1(div_up(W, 32), div_up(m, 32)), b1(32, 8);
	k0_mask_kernel<<<g1, b1, 0, d.stream>>>(p.msa, p.pitch, m, n, W, ign, p.NG, p.IG);
	k0_chunk_sum_kernel<<<div_up(NC * m, 256), 256, 0, d.stream>>>(p.NG, p.IG, m, W, NC, CS, CL, CI, CG);
	k0_chunk_scan_kernel<<<div_up(m, 128), 128, 0, d.stream>>>(m, NC, CS, CL, CI, CG, p.rowlen, p.lastcol, p.scalars);
	k0_chunk_apply_kernel<<<div_up(NC * m, 256), 256, 0, d.stream>>>(p.NG, p.IG, m, W, NC, CS, CI, p.RK, p.NIA);
	k0_starts_kernel<<<1, 1024, 0, d.stream>>>(p.rowlen, m, p.start, p.scalars);
	FBG_LAUNCH_CHECK();
	FBG_CUDA(cudaMemcpyAsync(p.h_scalars, p.scalars, 8 * sizeof(u64), cudaMemcpyDeviceToHost, d.stream));
	FBG_CUDA(cudaStreamSynchronize(d.stream));
	u64 nongap = p.h_scalars[0];
	p.toignore = p.h_scalars[1];
	p.gaps = m * n - nongap;
	p.N = p.h_scalars[2];
	if (p.N >= 0xFFFFFFF0ull)
		throw unsupported_error("text length " + std::to_string(p.N) + " needs suffix indices wider than 32 bits, which this build does not have (single-GPU and sharded alike); "
		                        "--heuristic-subset=ROWNUM builds the graph from row chunks that fit");
	if (!p.T) { p.T = p.dev.palloc<u8>(p.N + 64); p.col = p.dev.palloc<u32>(p.N + 8); }
	FBG_CUDA(cudaMemsetAsync(p.T + p.N, 0, 64, d.stream)); // zero tail so k-mer windows can over-read
	u64 wb = div_up(W, 256);
	if (p.gaps == 0) k0_text_gapless_kernel<<<div_up((p.N + 3) / 4, 256), 256, 0, d.stream>>>(p.msa, p.pitch, m, n, p.T, p.col, p.N);
	else k0_text_kernel<<<(u32) (wb * m), 256, 0, d.stream>>>(p.msa, p.pitch, m, n, W, wb, p.RK, p.start, p.rowlen, p.T, p.col, p.N);
	FBG_LAUNCH_CHECK();
	d.end((double) m * n * 2.0 + (double) p.N * 5.0 + (double) W * m * 4.0 * 6.0);
	d.note_launches(K_PACK, 5);
	d.release(mk);
}

} // namespace fbg
