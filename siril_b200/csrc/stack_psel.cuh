// stack_psel.cuh — kernel of the SELECTION engine (engine 6): mean with no / percentile / sigma
// rejection and the median stack WITHOUT sorting.  psel.cuh explains the method; this file is
// phase A on the device and the glue to phase B.
//
// One CTA owns SUB sub-tiles of 32 consecutive pixels of one layer (tile load: TMA, one 3-D box per
// sub-tile and <= 256 frames, 128B swizzle; or coalesced row loads with x shifts / 16-bit samples —
// stack_sorted.cuh's loaders).  A group of T = 2^LT lanes owns one pixel: lane t walks the samples
// of rows t, t + T, ... of the pixel's column (bank-conflict free, as the lane-sort engine); a
// sub-tile is T warp-rounds of 32/T pixels and the CTA has SUB*T warps, one round each.
//  phase A  psel_group (psel.cuh) drives the streaming passes below.  The samples STAY in shared
//           memory and every pass is a compact loop of eight loads and eight bodies (an earlier
//           version kept them in registers with every pass unrolled: 19 k SASS lines, and the
//           instruction cache, not the pipes, set the pace): sums (LDS, F2F.F64, DADD, DFMA, FMNMX3),
//           count (LDS, 2 FADD, 2 LEA.HI: sign bits added up, no compare), extract (LDS, 2 FADD,
//           LEA.HI, 2 FSETP, predicated STS + pointer bump into a per-lane list).  Scalars of a
//           pixel are combined over its T lanes with shuffles.
//  phase B  B1, still the group: the lists are dealt out evenly over the lanes, sorted in registers
//           (the lane-sort engine's networks) and written over the first rows of the pixel's column;
//           B2, after a barrier: thread i of the CTA finishes pixel i (psel_pixel: sorted.cuh's
//           rejection code through the partial-rank accessor).
#pragma once
#include "stack_lsort.cuh"
#include "psel.cuh"

namespace sgpu {

__device__ __forceinline__ float lds_f32(unsigned a) { float x; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(a)); return x; }
__device__ __forceinline__ void sts_f32(unsigned a, float x) { asm volatile("st.shared.f32 [%0], %1;" :: "r"(a), "f"(x) : "memory"); }

// the streaming passes of one lane (psel_group's `P`)
template <int LT>
struct PselLane {
	static constexpr int T = 1 << LT;
	static constexpr int NA = (T >= 8) ? 1 : 8 / T;	// distinct values of row & 7 among a lane's rows
	unsigned la[NA];	// shared-memory address of (row q*T + t, px): rows 8 apart share the swizzle term
	int t, N;
	unsigned list0;		// shared-memory address of this thread's first list slot
	unsigned lstride;	// bytes between its slots
	unsigned lcap;		// slots
	int cnt;		// entries after the last extraction
	unsigned pk;		// of which: centre | centre below p1 << 10 | low tail << 20 (classify)
	float *rec;		// split form: where this pixel's entries go (global memory), or null
	const double2 *coef;	// per-frame {scale, offset | mul} when the tile holds raw samples, else null
	int norm_mode;		// 1 additive family, 2 multiplicative family (with coef)

	__device__ __forceinline__ bool any(bool b) const { return __any_sync(0xffffffffu, b); }
	template <typename V> __device__ __forceinline__ V gsum(V x) const {
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x += __shfl_xor_sync(0xffffffffu, x, m);
		return x;
	}
	__device__ __forceinline__ float gmin(float x) const {
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x = fminf(x, __shfl_xor_sync(0xffffffffu, x, m));
		return x;
	}
	__device__ __forceinline__ float gmax(float x) const {
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, m));
		return x;
	}

	// f(k, x, row, address) for every sample of the lane, eight at a time: loads first, then the
	// bodies; k = 0..7 is the position inside the batch (compile-time: accumulator chains alternate on it)
	template <class F> __device__ __forceinline__ void each(F f) const {
		const int nfull = N / (8 * T);
		unsigned off = 0;
#pragma unroll 1
		for (int it = 0; it < nfull; it++, off += 1024u * T) {
			float x[8];
#pragma unroll
			for (int k = 0; k < 8; k++) x[k] = lds_f32(la[k % NA] + off + (unsigned) ((k - k % NA) * T * 128));
#pragma unroll
			for (int k = 0; k < 8; k++) f(k, x[k], it * 8 * T + k * T + t, la[k % NA] + off + (unsigned) ((k - k % NA) * T * 128));
		}
		// the rest (< 8 T rows): batches of NA samples, one of each swizzle term
		for (int row = nfull * 8 * T + t; row < N; row += NA * T, off += (unsigned) (NA * T * 128)) {
#pragma unroll
			for (int q = 0; q < NA; q++) if (row + q * T < N) {
				const unsigned a = la[q] + off;
				f(q, lds_f32(a), row + q * T, a);
			}
		}
	}

	__device__ __forceinline__ void sums(int is_mean, int *n, double *S1, double *S2, float *mn, float *mx) {
		float lo = SGPU_INF, hi = -SGPU_INF;
		if (!is_mean) {
			// median stack: no sample is dropped and the sums only PLACE the windows (no result depends on them), so float
			// accumulators do: FADD + FFMA per sample instead of F2F.F64 + DADD + DFMA, the first of which runs on the
			// conversion pipe at an eighth of the rate and bounds this pass in mean mode
			float a1 = 0.f, a2 = 0.f, b1 = 0.f, b2 = 0.f;
			if (coef) {
				each([&](int k, float x, int row, unsigned a) {
					const double2 c = coef[row];
					x = norm_mode == 1 ? normalize_sample<float, 1>(x, c.x, c.y) : normalize_sample<float, 2>(x, c.x, c.y);
					sts_f32(a, x);
					if (k & 1) { b1 += x; b2 = fmaf(x, x, b2); } else { a1 += x; a2 = fmaf(x, x, a2); }
					lo = fminf(lo, x); hi = fmaxf(hi, x);
				});
			} else {
				each([&](int k, float x, int, unsigned) {
					if (k & 1) { b1 += x; b2 = fmaf(x, x, b2); } else { a1 += x; a2 = fmaf(x, x, a2); }
					lo = fminf(lo, x); hi = fmaxf(hi, x);
				});
			}
			*S1 = (double) gsum(a1 + b1); *S2 = (double) gsum(a2 + b2);
			*n = N; *mn = gmin(lo); *mx = gmax(hi);
			return;
		}
		double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
		if (coef) {	// raw samples in the tile: normalize (median_and_mean.c:1644-1686) and park the result
			each([&](int k, float x, int row, unsigned a) {
				const double2 c = coef[row];
				x = norm_mode == 1 ? normalize_sample<float, 1>(x, c.x, c.y) : normalize_sample<float, 2>(x, c.x, c.y);
				sts_f32(a, x);
				const double xd = (double) x;
				if (k & 1) { b1 += xd; b2 = fma(xd, xd, b2); } else { a1 += xd; a2 = fma(xd, xd, a2); }
				lo = fminf(lo, x); hi = fmaxf(hi, x);
			});
		} else {
			each([&](int k, float x, int, unsigned) {
				const double xd = (double) x;
				if (k & 1) { b1 += xd; b2 = fma(xd, xd, b2); } else { a1 += xd; a2 = fma(xd, xd, a2); }
				lo = fminf(lo, x); hi = fmaxf(hi, x);		// NaN is ignored
			});
		}
		*S1 = gsum(a1 + b1); *S2 = gsum(a2 + b2);
		lo = gmin(lo); hi = gmax(hi);
		int nz = 0;
		if (is_mean && any(lo <= 0.f && hi >= 0.f)) {	// nulls: drop them (rejection_float.c:128-135)
			const bool mine = lo <= 0.f && hi >= 0.f;
			lo = SGPU_INF; hi = -SGPU_INF;
			each([&](int, float x, int, unsigned a) {
				if (mine && x == 0.f) { x = PSEL_NAN; sts_f32(a, x); nz++; }
				lo = fminf(lo, x); hi = fmaxf(hi, x);
			});
			nz = gsum(nz); lo = gmin(lo); hi = gmax(hi);
		}
		*n = N - nz; *mn = lo; *mx = hi;
	}

	__device__ __forceinline__ void shifted_sums(double K, double *S1, double *S2) {
		double a1 = 0.0, a2 = 0.0;
		each([&](int, float x, int, unsigned) {
			const double y = (x == x) ? (double) x - K : 0.0;
			a1 += y; a2 = fma(y, y, a2);
		});
		*S1 = gsum(a1); *S2 = gsum(a2);
	}

	// sign bits added up: x < mean, |x - mean| < hw (NaN: neither)
	__device__ __forceinline__ void count(float mean, float hw, int *c_neg, int *c_in) {
		unsigned a0 = 0, a1 = 0, b0 = 0, b1 = 0;
		each([&](int k, float x, int, unsigned) {
			const float d = x - mean;
			const float u = fabsf(d) - hw;
			if (k & 1) { a1 += __float_as_uint(d) >> 31; b1 += __float_as_uint(u) >> 31; }
			else { a0 += __float_as_uint(d) >> 31; b0 += __float_as_uint(u) >> 31; }
		});
		const unsigned pk = gsum((a0 + a1) | ((b0 + b1) << 16));	// stacks are at most 1280 deep
		*c_neg = (int) (pk & 0xffffu); *c_in = (int) (pk >> 16);
	}

	__device__ __forceinline__ void extract(bool on, const PselPlan &pl, int *below, bool *over) {
		unsigned ptr = list0;
		const unsigned lim = list0 + (lcap - 8u) * lstride;	// a batch of 8 samples starts only with 8 free slots
		float w = on ? pl.w : -1.f, rT = on ? pl.rT : FLT_MAX;
		const float p1 = pl.p1, cT = pl.cT;
		const unsigned ls = lstride;
		unsigned c0 = 0, c1 = 0;
		bool ov = false;
		each([&](int k, float x, int, unsigned) {
			if (k == 0 && ptr > lim) { w = -1.f; rT = FLT_MAX; ov = true; }
			const float d1 = x - p1, d2 = x - cT;
			if (k & 1) c1 += __float_as_uint(d1) >> 31; else c0 += __float_as_uint(d1) >> 31;
			asm volatile("{\n\t.reg .pred q;\n\tsetp.gt.f32 q, %2, %3;\n\tsetp.lt.or.f32 q, %4, %5, q;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, %6;\n\t}"
				: "+r"(ptr) : "f"(x), "f"(fabsf(d2)), "f"(rT), "f"(fabsf(d1)), "f"(w), "r"(ls) : "memory");
		});
		if (on) cnt = (int) ((ptr - list0) / ls);	// a group that is done keeps its lists while others extract again
		*below = (int) gsum(c0 + c1);
		*over = gsum((int) ov) != 0;
	}

	__device__ __forceinline__ void classify(const PselPlan &pl, int *cc, int *ccb, int *nl, int *m) {
		pk = 0;
		int pre = cnt;		// entries of the lanes before this one (split form: the entries go to the pixel's record on the way;
					// an extraction that is repeated simply writes them again)
#pragma unroll
		for (int d = 1; d < T; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, pre, d, T); if (t >= d) pre += o; }
		pre -= cnt;
		float *out = rec ? rec + PSEL_REC_HDR + pre : nullptr;
		const int room = PSEL_REC_ENTRIES - pre;
		for (int k = 0; k < cnt; k++) {
			const float x = lds_f32(list0 + (unsigned) k * lstride);
			const float d = x - pl.p1;
			const bool in = fabsf(d) < pl.w, neg = d < 0.f;
			pk += in ? (neg ? 1u + (1u << 10) : 1u) : (neg ? (1u << 20) : 0u);
			if (out && k < room) out[k] = x;
		}
		// fields of 10 bits: a class with more than 1023 entries is far past what phase B sorts
		const unsigned a = gsum(pk & 0x3ffu), b = gsum((pk >> 10) & 0x3ffu), l = gsum(pk >> 20);
		*cc = (int) a; *ccb = (int) b; *nl = (int) l; *m = gsum(cnt);
	}
};

// a pixel's column inside the 128B-swizzled tile: row i, as the TMA wrote it (B1 sorts into the
// column while other groups of the sub-tile still read theirs, so the swizzle stays)
struct PselCol {
	unsigned a0, c;	// address of (row 0, chunk 0) + (px & 3) * 4; c = px >> 2
	__device__ __forceinline__ PselCol(const float *srows, int px) : a0(smem_u32(srows) + ((unsigned) (px & 3) << 2)), c((unsigned) px >> 2) { }
	__device__ __forceinline__ unsigned at(int i) const { return a0 + (unsigned) i * 128u + ((((unsigned) i ^ c) & 7u) << 4); }
	__device__ __forceinline__ float get(int i) const { return lds_f32(at(i)); }
	__device__ __forceinline__ void set(int i, float x) const { sts_f32(at(i), x); }
};

// register slots per lane for the sort of the lists (network sizes of the lane-sort engine): 64 entries
// in all, and 96 and more for the rare long list
constexpr int psel_sort_slots(int T) { return T == 1 ? 64 : (T == 2 ? 32 : (T == 4 ? 16 : (T == 8 ? 8 : 4))); }
constexpr int psel_sort_slots2(int T) { return T == 1 ? 80 : (T == 2 ? 48 : (T == 4 ? 24 : (T == 8 ? 16 : 8))); }

// B2 of one pixel: rejection and combine on the sorted rows, result and rejection-map stores (by the
// lanes with `store`), hand-over list, rejection counts into the CTA's counters
template <class Store>
__device__ __forceinline__ void psel_finish(const StackParams &p, const Store &col, const PselHdr &h, int layer, size_t pix, size_t npix,
		int N, bool store, int lane, unsigned *cta_rej) {
	int rej[2] = { 0, 0 };
	if (pix < npix) {
		int status;
		double result;
		const int x = (int) (pix % p.W);
		const double *w = (p.is_mean && p.weights) ? p.weights + (size_t) layer * N : nullptr;
		if (p.dbg_stop == 5) { result = h.S1; status = FAST_DONE; rej[0] = h.nl & 1; rej[1] = h.cc & 1; }	// timing: stores without the rejection
		else if (p.is_u16) {
			auto orig = [&](int ff) { return (float) load_norm<u16>(p, ff, layer, pix, x); };
			result = psel_pixel<true>(col, h, N, p.is_mean, p.rc, w, orig, rej, &status, p.rcp_tab);
		} else {
			auto orig = [&](int ff) { return load_norm<float>(p, ff, layer, pix, x); };
			result = psel_pixel<false>(col, h, N, p.is_mean, p.rc, w, orig, rej, &status, p.rcp_tab);
		}
		if (p.dbg_stop == 4 && result != 123.456) return;	// timing: the rejection without the stores
		if (!store) rej[0] = rej[1] = 0;
		else if (status == FAST_DONE) {
			if (p.is_u16) store_result<u16>(p, (size_t) layer * p.out_plane + pix, result, rej);
			else store_result<float>(p, (size_t) layer * p.out_plane + pix, result, rej);
		} else {
			rej[0] = rej[1] = 0;
			p.list[atomicAdd(p.list_count, 1u)] = ((unsigned) layer << 30) | (unsigned) pix;
		}
	}
	if (p.is_mean && p.rc.type != SGPU_NO_REJEC) {
		__syncwarp();
		unsigned lo = rej[0], hi = rej[1];
		lo = __reduce_add_sync(0xffffffffu, lo); hi = __reduce_add_sync(0xffffffffu, hi);
		if (lane == 0 && (lo | hi)) {
			if (cta_rej) { atomicAdd(&cta_rej[0], lo); atomicAdd(&cta_rej[1], hi); }
			else {	// (split form: the few warps that finish their pixels here; the CTA's other warps have left)
				atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
				atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
			}
		}
	}
}

// rows [0, m) of the column, dealt out over the group's lanes (lane t: rows t, t + T, ...), sorted, and
// written lane-major (rank t * RB + k) into `out`
template <int LT, int RB, class Out>
__device__ __forceinline__ void psel_sort_into(const PselCol &col, const Out &out, int t, int m, unsigned gmask) {
	constexpr int T = 1 << LT;
	float e[RB];
#pragma unroll
	for (int k = 0; k < RB; k++) { e[k] = SGPU_INF; if (t + T * k < m) e[k] = col.get(t + T * k); }
	__syncwarp();
	LocalNet<RB>::template run<RB>(e);
	if constexpr (LT > 0) group_merge<LT, RB>(e, t, gmask);
#pragma unroll
	for (int k = 0; k < RB; k++) if (t * RB + k < m) out.set(t * RB + k, e[k]);
}

// One CTA per tile.  (A persistent form — CTAs walking several tiles, the tile buffer handed to the TMA for the
// next tile as soon as B1 had moved the interesting samples out, B2 running under that load — was built and
// measured SLOWER on the target, 54 against 48 ms: the barrier it needs couples the CTA's warps, which otherwise
// never meet, and four independent CTAs per SM already overlap each other's loads.)
template <int LT, int SUB>
__global__ void __launch_bounds__((SUB << LT) * 32, (SUB << LT) <= 4 ? 4 : ((SUB << LT) <= 8 ? 2 : 1))
stack_psel_kernel(const StackParams p, const __grid_constant__ CUtensorMap tmap) {
	constexpr int T = 1 << LT, NW = SUB << LT, NT = NW * 32, PXC = SUB * TILE_PX;
	extern __shared__ __align__(1024) unsigned char smem_raw[];
	const int N = p.N;
	float *rows = (float *) smem_raw;	// [SUB][rows8][32]; the swizzle needs it 1024-byte aligned: the kernel has no static shared memory before it
	if (smem_u32(smem_raw) & 1023u) __trap();
	const int rows8 = (p.tile_rows + 9) & ~7;	// two rows more than the stack is deep: phase B adds two stand-ins to the sorted lists
	const size_t sub_floats = (size_t) rows8 * TILE_PX;	// sub-tiles start on 1024-byte swizzle atoms
	unsigned long long *bar = (unsigned long long *) (rows + SUB * sub_floats);
	unsigned *cta_rej = (unsigned *) (bar + 1);	// rejections of the CTA's pixels, low / high: one pair of global atomics per CTA
	double2 *coef = (double2 *) (bar + 2);		// [N] {scale, offset | mul} when the tile holds raw samples
	const bool raw_norm = p.use_tma && p.normalize != SGPU_NO_NORM;
	constexpr int LCAP = psel_list_cap(T);
	float *lists = (float *) (coef + (raw_norm ? N : 0));	// [LCAP][NT]
	const int layer = (int) blockIdx.y + p.layer_base, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t npix = (size_t) p.H * p.W, pix0 = p.pix_base + (size_t) blockIdx.x * PXC;

	if (threadIdx.x == 0) { cta_rej[0] = 0; cta_rej[1] = 0; }
	if (p.use_tma) {
		if (threadIdx.x == 0) mbar_init(bar, 1);
		__syncthreads();
		if (threadIdx.x == 0) {
			int nsub = 0;	// sub-tiles that start inside the image (the others are never looked at)
			for (int sb = 0; sb < SUB; sb++) if (pix0 + (size_t) sb * TILE_PX < npix) nsub++;
			mbar_expect_tx(bar, (unsigned) (nsub * (size_t) p.tile_rows * TILE_PX * sizeof(float)));
			for (int sb = 0; sb < nsub; sb++)
				for (int r0 = 0; r0 < p.tile_rows; r0 += p.box_rows)
					tma_load_3d(rows + sb * sub_floats + (size_t) r0 * TILE_PX, &tmap, (int) (pix0 + sb * TILE_PX), layer, r0, bar);
		}
		if (raw_norm) {	// per-frame coefficients, staged while the tile is in flight
			const bool add = p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING;
			const double *c0 = p.scale + layer * N, *c1 = (add ? p.offset : p.mul) + layer * N;
			for (int i = threadIdx.x; i < N; i += NT) coef[i] = make_double2(c0[i], c1[i]);
			__syncthreads();
		}
		mbar_wait(bar, 0);
	} else {
		for (int sb = 0; sb < SUB; sb++) {
			if (p.is_u16) load_tile_any<u16, NW>(p, rows + sb * sub_floats, layer, pix0 + sb * TILE_PX, npix, warp, lane);
			else load_tile_any<float, NW>(p, rows + sb * sub_floats, layer, pix0 + sb * TILE_PX, npix, warp, lane);
		}
		__syncthreads();
	}
	if (p.dbg_stop == 1) return;

	// ---- phase A: warp w works on sub-tile w % SUB, round w / SUB ----
	{
		PselLane<LT> ln;
		const int t = lane & (T - 1), g = lane >> LT;
		ln.t = t; ln.N = N;
		const int sb = warp % SUB, r = warp / SUB;
		const float *srows = rows + sb * sub_floats;	// 1024-byte aligned
		const int px = lsort_pixel<LT>(r, g);
		{	// address of (row q*T + t, px): the 128B swizzle XORs the 16-byte chunk index with row & 7
			const unsigned a0 = smem_u32(srows) + (((unsigned) (t * 128) | ((unsigned) (((px >> 2) ^ t) & 7) << 4) | ((unsigned) (px & 3) << 2)));
#pragma unroll
			for (int q = 0; q < PselLane<LT>::NA; q++) ln.la[q] = (a0 ^ ((unsigned) ((q * T) & 7) << 4)) + (unsigned) (q * T * 128);
		}
		ln.coef = raw_norm ? coef : nullptr;
		ln.norm_mode = (p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING) ? 1 : 2;
		ln.list0 = smem_u32(lists + threadIdx.x); ln.lstride = NT * 4; ln.lcap = LCAP; ln.cnt = 0;
		{
			const size_t pixg = pix0 + (size_t) (sb * TILE_PX + px);
			ln.rec = (p.split && pixg < npix) ? p.records + (pixg - p.pix_base) * PSEL_REC_WORDS : nullptr;
		}
		const int rt = p.is_mean ? p.rc.type : SGPU_NO_REJEC;
		PselHdr h;
		psel_group(ln, p.is_mean, rt == SGPU_SIGMA || rt == SGPU_PERCENTILE, rt == SGPU_PERCENTILE ? 1 : 0, p.rc.s0, p.rc.s1, h);
		h.tid0 = (int) threadIdx.x - t;
		if (p.dbg_stop == 2) return;
		if (p.split) {
			// ---- split form: the pixel's header and its (unsorted) entries go to global memory; psel_finish_kernel
			// sorts them and runs the rejection with one thread per pixel at full occupancy, which this kernel,
			// whose CTAs are sized by the tile they hold, cannot offer to the scalar tail of a pixel ----
			const size_t pix = pix0 + (size_t) (sb * TILE_PX + px);
			const bool valid = pix < npix;
			if (!ln.any(valid)) return;	// (a warp past the image's last pixel)
			const bool live = !(h.flags & (PSEL_WEIRD | PSEL_HANDOVER | PSEL_TRIVIAL));
			// a pixel with more entries than a record holds: its whole warp stays here and finishes below (B1, B2)
			const bool stay = ln.any(live && h.m > PSEL_REC_ENTRIES);
			float *rec = p.records + (pix - p.pix_base) * PSEL_REC_WORDS;
			const uint4 q0 = make_uint4((unsigned) __double2loint(h.S1), (unsigned) __double2hiint(h.S1), (unsigned) __double2loint(h.S2), (unsigned) __double2hiint(h.S2));
			const uint4 q1 = make_uint4((unsigned) __double2loint(h.K), (unsigned) __double2hiint(h.K), __float_as_uint(h.mn), __float_as_uint(h.mx));
			const uint4 q2 = make_uint4(__float_as_uint(h.blo), __float_as_uint(h.bhi), (unsigned) h.n, (unsigned) h.cl);
			const uint4 q3 = make_uint4((unsigned) h.nl, (unsigned) h.cc, (unsigned) h.m, (unsigned) (h.flags | (stay ? PSEL_DONE : 0)));
			uint4 *rq = (uint4 *) rec;
			if (!valid) { }
			else if (T >= 4) { if (t == 0) rq[0] = q0; else if (t == 1) rq[1] = q1; else if (t == 2) rq[2] = q2; else if (t == 3) rq[3] = q3; }
			else if (T == 2) { if (t == 0) { rq[0] = q0; rq[2] = q2; } else { rq[1] = q1; rq[3] = q3; } }
			else { rq[0] = q0; rq[1] = q1; rq[2] = q2; rq[3] = q3; }
			if (!stay) return;	// (the entries went to the record while they were classified)
		}
		// ---- B1: the group's entries, dealt out evenly over its lanes, sorted in registers (the
		// lane-sort engine's networks) and written over the first rows of the pixel's (now dead) column ----
		constexpr int RB = psel_sort_slots(T), RB2 = psel_sort_slots2(T);
		const bool live = !(h.flags & (PSEL_WEIRD | PSEL_HANDOVER | PSEL_TRIVIAL));
		if (live && h.m > T * RB2) h.flags |= PSEL_HANDOVER;	// more than the networks take
		const bool sorting = live && h.m <= T * RB2;
		const PselCol col(srows, px);
		const int mine = sorting ? ln.cnt : 0, m = sorting ? h.m : 0;
		int pre = mine;		// entries of the lanes before this one
#pragma unroll
		for (int d = 1; d < T; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, pre, d, T); if (t >= d) pre += o; }
		pre -= mine;
		for (int k = 0; k < mine; k++) col.set(pre + k, lds_f32(ln.list0 + (unsigned) k * ln.lstride));
		__syncwarp();
		if (!ln.any(m > T * RB)) psel_sort_into<LT, RB>(col, col, t, m, 0xffffffffu);
		else psel_sort_into<LT, RB2>(col, col, t, m, 0xffffffffu);	// a long list somewhere in the warp: the larger network
		__syncwarp();
		if (p.dbg_stop == 3) return;
		// ---- B2, still the group: every lane of it runs the pixel's rejection on the sorted rows (the same
		// instructions for all T lanes: no barrier, the header never leaves the registers); lane 0 stores.
		// (One thread per pixel after a CTA barrier issues a quarter of the instructions on a quarter of the
		// warps: measured the same, 46.8 against 47.8 ms.) ----
		psel_finish(p, col, h, layer, pix0 + (size_t) (sb * TILE_PX + px), npix, N, t == 0, lane, p.split ? nullptr : cta_rej);
	}
	if (!p.split && p.is_mean && p.rc.type != SGPU_NO_REJEC) {
		__syncthreads();
		if (threadIdx.x == 0 && (cta_rej[0] | cta_rej[1])) {
			atomicAdd(&p.irej[layer * 2], (unsigned long long) cta_rej[0]);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) cta_rej[1]);
		}
	}
}

}  // namespace sgpu

template <int LT, int SUB>
int sgpu_launch_psel_variant(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm);

#define SGPU_PSEL_VARIANT(LT, SUB) \
	template <> int sgpu_launch_psel_variant<LT, SUB>(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm) { \
		auto kern = sgpu::stack_psel_kernel<LT, SUB>; \
		const CUtensorMap tm = tmap ? *tmap : CUtensorMap{}; \
		if (shm > 48 * 1024) SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm)); \
		kern<<<grid, (SUB << LT) * 32, shm, ctx->stream>>>(p, tm); \
		ctx->launches++; \
		SGPU_CUDA(ctx, cudaGetLastError()); \
		return SGPU_OK; \
	}
