// stack_psel.cuh — kernel of the SELECTION engine (engine 6): mean with no / percentile / sigma
// rejection and the median stack WITHOUT sorting.  psel.cuh explains the method; this file is
// phase A on the device and the glue to phase B.
//
// One CTA owns SUB sub-tiles of 32 consecutive pixels of one layer (tile load: TMA, one 3-D box per
// sub-tile and <= 256 frames, 128B swizzle; or coalesced row loads with x shifts / 16-bit samples —
// stack_sorted.cuh's loaders).  A group of T = 2^LT lanes owns one pixel: lane t takes the samples
// of rows j*T + t (bank-conflict free, as the lane-sort engine) into R registers and KEEPS them
// there; a sub-tile is T warp-rounds of 32/T pixels and the CTA has SUB*T warps, one round each.
//  phase A  psel_group (psel.cuh) drives the streaming passes below over the registers: sums
//           (F2F.F64 + DADD + DFMA + 2 FMNMX per sample), count (2 FADD + 2 LEA.HI: sign bits added
//           up, no compare), extract (2 FADD, LEA.HI, 2 FSETP, predicated STS + pointer bump).
//           Scalars of a pixel are combined over its T lanes with shuffles.  The extracted samples
//           go to per-thread lists laid over the (by then dead) upper rows of the tile when the
//           stack is deep enough, else to shared memory of their own.
//  phase B  B1, still the group: the lists are sorted in registers (the lane-sort engine's networks)
//           into the first rows of the pixel's column; B2, after a barrier: thread i of the CTA finishes pixel
//           i (psel_pixel: sorted.cuh's rejection code through the partial-rank accessor).
#pragma once
#include "stack_lsort.cuh"
#include "psel.cuh"

namespace sgpu {

// the streaming passes of one lane (psel_group's `P`)
template <int LT, int R, int RF>
struct PselLane {
	static constexpr int T = 1 << LT;
	float v[R];		// samples j*T + t; NaN: absent (beyond the stack, or a null in mean mode)
	int t, N;
	unsigned gmask;
	unsigned list0;		// shared-memory address of this thread's first list slot
	unsigned lstride;	// bytes between its slots
	unsigned lcap;		// slots
	int cnt;		// entries after the last extraction

	__device__ __forceinline__ bool any(bool b) const { return __any_sync(0xffffffffu, b); }
	template <typename V> __device__ __forceinline__ V gsum(V x) const {
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x += __shfl_xor_sync(gmask, x, m);
		return x;
	}
	__device__ __forceinline__ float gmin(float x) const {
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x = fminf(x, __shfl_xor_sync(gmask, x, m));
		return x;
	}
	__device__ __forceinline__ float gmax(float x) const {
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x = fmaxf(x, __shfl_xor_sync(gmask, x, m));
		return x;
	}

	__device__ __forceinline__ void sums(int is_mean, int *n, double *S1, double *S2, float *mn, float *mx) {
		double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
		float lo = SGPU_INF, hi = -SGPU_INF;
#pragma unroll
		for (int j = 0; j < R; j++) {
			const float x = v[j];
			double xd = (double) x;
			if (j >= RF && j * T + t >= N) xd = 0.0;	// beyond the stack (a NaN among the samples must poison the sums)
			if (j & 1) { b1 += xd; b2 = fma(xd, xd, b2); } else { a1 += xd; a2 = fma(xd, xd, a2); }
			lo = fminf(lo, x); hi = fmaxf(hi, x);		// NaN is ignored
		}
		*S1 = gsum(a1 + b1); *S2 = gsum(a2 + b2);
		lo = gmin(lo); hi = gmax(hi);
		int nz = 0;
		if (is_mean && any(lo <= 0.f && hi >= 0.f)) {	// nulls: drop them (rejection_float.c:128-135)
			const bool mine = lo <= 0.f && hi >= 0.f;
			lo = SGPU_INF; hi = -SGPU_INF;
#pragma unroll
			for (int j = 0; j < R; j++) {
				if (mine && v[j] == 0.f) { v[j] = PSEL_NAN; nz++; }
				lo = fminf(lo, v[j]); hi = fmaxf(hi, v[j]);
			}
			nz = gsum(nz); lo = gmin(lo); hi = gmax(hi);
		}
		*n = N - nz; *mn = lo; *mx = hi;
	}

	__device__ __forceinline__ void shifted_sums(double K, double *S1, double *S2) {
		double a1 = 0.0, a2 = 0.0;
#pragma unroll
		for (int j = 0; j < R; j++) {
			const float x = v[j];
			const double y = (x == x) ? (double) x - K : 0.0;
			a1 += y; a2 = fma(y, y, a2);
		}
		*S1 = gsum(a1); *S2 = gsum(a2);
	}

	// sign bits added up: x < mean, |x - mean| < hw (NaN: neither)
	__device__ __forceinline__ void count(float mean, float hw, int *c_neg, int *c_in) {
		unsigned a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll
		for (int j = 0; j < R; j++) {
			const float d = v[j] - mean;
			const float u = fabsf(d) - hw;
			if (j & 1) { a1 += __float_as_uint(d) >> 31; b1 += __float_as_uint(u) >> 31; }
			else { a0 += __float_as_uint(d) >> 31; b0 += __float_as_uint(u) >> 31; }
		}
		const unsigned pk = gsum((a0 + a1) | ((b0 + b1) << 16));	// stacks are at most 2560 deep
		*c_neg = (int) (pk & 0xffffu); *c_in = (int) (pk >> 16);
	}

	__device__ __forceinline__ void extract(bool on, const PselPlan &pl, int *below, bool *over) {
		unsigned ptr = list0;
		const unsigned lim = list0 + (lcap - 8u) * lstride;	// a chunk of 8 samples starts only with 8 free slots
		float w = on ? pl.w : -1.f, rT = on ? pl.rT : FLT_MAX;
		const float p1 = pl.p1, cT = pl.cT;
		unsigned c0 = 0, c1 = 0;
		bool ov = false;
#pragma unroll
		for (int j = 0; j < R; j++) {
			if (j % 8 == 0 && j > 0) { if (ptr > lim) { w = -1.f; rT = FLT_MAX; ov = true; } }
			const float x = v[j];
			const float d1 = x - p1, d2 = x - cT;
			if (j & 1) c1 += __float_as_uint(d1) >> 31; else c0 += __float_as_uint(d1) >> 31;
			asm volatile("{\n\t.reg .pred q;\n\tsetp.gt.f32 q, %2, %3;\n\tsetp.lt.or.f32 q, %4, %5, q;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, %6;\n\t}"
				: "+r"(ptr) : "f"(x), "f"(fabsf(d2)), "f"(rT), "f"(fabsf(d1)), "f"(w), "r"(lstride) : "memory");
		}
		if (on) cnt = (int) ((ptr - list0) / lstride);	// a group that is done keeps its lists while others extract again
		*below = (int) gsum(c0 + c1);
		*over = gsum((int) ov) != 0;
	}

	__device__ __forceinline__ void classify(const PselPlan &pl, int *cc, int *ccb, int *nl, int *m) const {
		unsigned pk = 0;	// centre | centre below p1 << 10 | low tail << 20
		for (int k = 0; k < cnt; k++) {
			float x;
			asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(list0 + (unsigned) k * lstride));
			const float d = x - pl.p1;
			const bool in = fabsf(d) < pl.w, neg = d < 0.f;
			pk += in ? (neg ? 1u + (1u << 10) : 1u) : (neg ? (1u << 20) : 0u);
		}
		// fields of 10 bits: a class with more than 1023 entries is far past what phase B sorts
		const unsigned a = gsum(pk & 0x3ffu), b = gsum((pk >> 10) & 0x3ffu), l = gsum(pk >> 20);
		*cc = (int) a; *ccb = (int) b; *nl = (int) l; *m = gsum(cnt);
	}
};

// phase B's views of shared memory.  The tile is dead once every sample sits in a register, so a
// pixel's column is addressed plainly (row i at i * 128 + px * 4 bytes: bank = px whatever the row)
struct PselCol {
	unsigned a0;	// shared-memory address of row 0
	__device__ __forceinline__ float get(int i) const { float x; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(a0 + (unsigned) i * 128u)); return x; }
	__device__ __forceinline__ void set(int i, float x) const { asm volatile("st.shared.f32 [%0], %1;" :: "r"(a0 + (unsigned) i * 128u), "f"(x) : "memory"); }
};

// register slots per lane for the sort of the lists (stack_lsort.cuh's network sizes)
constexpr int psel_sort_slots(int T) { return T == 1 ? 50 : (T == 2 ? 32 : (T == 4 ? 24 : (T == 8 ? 16 : 8))); }

constexpr int psel_min_blocks(int nt, int r) {
	const int regs = (r + 53 + 7) & ~7, b = 65536 / (nt * regs);
	return b < 1 ? 1 : (b > 6 ? 6 : b);
}

template <int LT, int R, int RF, int SUB>
__global__ void __launch_bounds__((SUB << LT) * 32, psel_min_blocks((SUB << LT) * 32, R))
stack_psel_kernel(const StackParams p, const __grid_constant__ CUtensorMap tmap) {
	constexpr int T = 1 << LT, NW = SUB << LT, NT = NW * 32, PXC = SUB * TILE_PX;
	extern __shared__ __align__(1024) unsigned char smem_raw[];
	const int N = p.N;
	float *rows = (float *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);	// [SUB][tile_rows8][32]
	const int rows8 = (p.tile_rows + 9) & ~7;	// two rows more than the stack is deep: phase B adds two stand-ins to the sorted lists
	const size_t sub_floats = (size_t) rows8 * TILE_PX;	// sub-tiles start on 1024-byte swizzle atoms
	PselHdr *hdr = (PselHdr *) (rows + SUB * sub_floats);
	unsigned long long *bar = (unsigned long long *) (hdr + PXC);
	double2 *coef = (double2 *) (bar + 2);		// [N] {scale, offset | mul} when the tile holds raw samples
	// lists: over the tile rows the sorted lists of phase B cannot reach, when those suffice
	constexpr int LCAP = psel_list_cap(T);
	const bool overlay = (size_t) (rows8 - PSEL_ROWS) * TILE_PX * SUB >= (size_t) LCAP * NT && rows8 > PSEL_ROWS;
	float *lists = overlay ? nullptr : (float *) (coef + (p.normalize != SGPU_NO_NORM && p.use_tma ? N : 0));
	const int layer = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t npix = (size_t) p.H * p.W, pix0 = (size_t) blockIdx.x * PXC;

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
		if (p.normalize != SGPU_NO_NORM) {	// per-frame coefficients, staged while the tile is in flight
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
	const int tma_mode = !p.use_tma ? 0 : (p.normalize == SGPU_NO_NORM ? 0 : ((p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING) ? 1 : 2));

	// ---- phase A: warp w works on sub-tile w % SUB, round w / SUB ----
	{
		PselLane<LT, R, RF> ln;
		const int t = lane & (T - 1), g = lane >> LT;
		ln.t = t; ln.N = N;
		ln.gmask = (T == 32) ? 0xffffffffu : (((1u << T) - 1u) << (lane & ~(T - 1)));
		const int sb = warp % SUB, r = warp / SUB;
		float *srows = rows + sb * sub_floats;	// 1024-byte aligned
		constexpr int NA = (T >= 8) ? 1 : 8 / T;	// distinct values of (j*T) & 7
		const int px = lsort_pixel<LT>(r, g);
		// byte address of (row j*T + t, px) = (la[k] + j*T*128) with k = ((j*T) & 7) / T: the 128B swizzle
		// XORs the 16-byte chunk index with row & 7 = ((j*T) & 7) | (t & 7) [T <= 8], t & 7 [T > 8]
		const char *la[NA];
		{
			const unsigned a0 = (unsigned) (t * 128) | ((unsigned) (((px >> 2) ^ t) & 7) << 4) | ((unsigned) (px & 3) << 2);
#pragma unroll
			for (int k = 0; k < NA; k++) la[k] = (const char *) srows + (a0 ^ ((unsigned) ((k * T) & 7) << 4));
		}
#pragma unroll
		for (int j = 0; j < R; j++) {
			float x = PSEL_NAN;
			const int i = j * T + t;
			if (j < RF || i < N) {	// rows below RF are full for every stack depth this variant serves
				x = *(const float *) (la[(((j * T) & 7) / T) % NA] + j * T * 128);
				if (tma_mode) {
					const double2 c = coef[i];
					x = tma_mode == 1 ? normalize_sample<float, 1>(x, c.x, c.y) : normalize_sample<float, 2>(x, c.x, c.y);
				}
			}
			ln.v[j] = x;
		}
		__syncthreads();	// every sample of the tile is in a register now: its rows become scratch (sorted lists below PSEL_ROWS, sample lists above when they fit)
		// lists: one array [LCAP][NT] when separate, [LCAP][32 T] per sub-tile when laid over the tile
		unsigned l0, lstr;
		if (overlay) {
			constexpr int per_sub = 32 * T;		// threads that work on one sub-tile
			const int q = r * 32 + lane;		// this thread among them
			l0 = smem_u32(rows + sb * sub_floats + (size_t) PSEL_ROWS * TILE_PX + q); lstr = per_sub * 4;
		} else { l0 = smem_u32(lists + threadIdx.x); lstr = NT * 4; }
		ln.list0 = l0; ln.lstride = lstr; ln.lcap = LCAP; ln.cnt = 0;
		const int rt = p.is_mean ? p.rc.type : SGPU_NO_REJEC;
		PselHdr h;
		psel_group(ln, p.is_mean, rt == SGPU_SIGMA || rt == SGPU_PERCENTILE, rt == SGPU_PERCENTILE, p.rc.s0, p.rc.s1, h);
		h.tid0 = (int) threadIdx.x - t;
		// ---- B1: the group sorts its lists in registers (the lane-sort engine's networks: every lane
		// sorts its own entries, then the runs are merged lane-major) into the first rows of the column ----
		constexpr int RE = psel_sort_slots(T);
		const bool sorting = !(h.flags & (PSEL_WEIRD | PSEL_HANDOVER | PSEL_TRIVIAL));
		if (ln.gsum((int) (sorting && ln.cnt > RE)) > 0) h.flags |= PSEL_HANDOVER;	// a lane with more entries than sort slots
		if (t == 0) hdr[sb * TILE_PX + px] = h;
		{
			float e[RE];
			const int mine = sorting ? ln.cnt : 0;
#pragma unroll
			for (int k = 0; k < RE; k++) {
				e[k] = SGPU_INF;
				if (k < mine) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(e[k]) : "r"(l0 + (unsigned) k * lstr));
			}
			LocalNet<RE>::template run<RE>(e);
			if constexpr (LT > 0) group_merge<LT, RE>(e, t, ln.gmask);
			__syncwarp();
			const unsigned c0 = smem_u32(srows + px) + (unsigned) (t * RE) * 128u;
			const int lim = sorting ? h.m - t * RE : 0;	// rank t*RE + k goes to row t*RE + k
#pragma unroll
			for (int k = 0; k < RE; k++) if (k < lim) asm volatile("st.shared.f32 [%0], %1;" :: "r"(c0 + (unsigned) k * 128u), "f"(e[k]) : "memory");
		}
	}
	__syncthreads();
	if (threadIdx.x >= PXC) return;

	// ---- phase B: thread i of the CTA owns pixel i of the tile ----
	const int sb = threadIdx.x >> 5;
	const size_t pix = pix0 + threadIdx.x;
	int rej[2] = { 0, 0 };
	if (pix < npix) {
		const PselHdr h = hdr[threadIdx.x];
		const PselCol col{ smem_u32(rows + sb * sub_floats + lane) };
		int status;
		double result;
		const int x = (int) (pix % p.W);
		const double *w = (p.is_mean && p.weights) ? p.weights + (size_t) layer * N : nullptr;
		if (p.is_u16) {
			auto orig = [&](int ff) { return (float) load_norm<u16>(p, ff, layer, pix, x); };
			result = psel_pixel<true>(col, h, N, p.is_mean, p.rc, w, orig, rej, &status);
		} else {
			auto orig = [&](int ff) { return load_norm<float>(p, ff, layer, pix, x); };
			result = psel_pixel<false>(col, h, N, p.is_mean, p.rc, w, orig, rej, &status);
		}
		if (status == FAST_DONE) {
			if (p.is_u16) store_result<u16>(p, (size_t) layer * npix + pix, result, rej);
			else store_result<float>(p, (size_t) layer * npix + pix, result, rej);
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
			atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
		}
	}
}

}  // namespace sgpu

template <int LT, int R, int RF, int SUB>
int sgpu_launch_psel_variant(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm);

#define SGPU_PSEL_VARIANT(LT, R, RF, SUB) \
	template <> int sgpu_launch_psel_variant<LT, R, RF, SUB>(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm) { \
		auto kern = sgpu::stack_psel_kernel<LT, R, RF, SUB>; \
		const CUtensorMap tm = tmap ? *tmap : CUtensorMap{}; \
		if (shm > 48 * 1024) SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm)); \
		kern<<<grid, (SUB << LT) * 32, shm, ctx->stream>>>(p, tm); \
		ctx->launches++; \
		SGPU_CUDA(ctx, cudaGetLastError()); \
		return SGPU_OK; \
	}
