// stack_reg.cuh — kernel of the REGISTER engine: mean with sigma / winsorized / percentile / no
// rejection and the median stack on float frames, without sorting (per-pixel logic: reg.cuh).
//
// One CTA of T warps owns a tile of 32 consecutive pixels of one layer.
//  load    the (32 px x N frames) tile comes in through TMA (one 3-D box per <= 256 frames,
//          128-byte rows, hardware 128B swizzle) or, with per-frame x shifts, through coalesced
//          128-byte row loads; every input sample crosses HBM exactly once.
//  gather  a group of T = 2^LT lanes owns one pixel: lane t takes the samples of rows j*T + t into
//          registers (bank-conflict free through the swizzle), normalizing in double as
//          median_and_mean.c:1644-1686 does when the tile holds raw samples.
//  reduce  everything else happens on those registers (reg.cuh); the pixel's column of the tile
//          becomes the lanes' scratch lists for the median window.
//  store   the group's first lane writes the pixel; pixels that cannot be decided identically
//          to the reference are appended to the list the exact engine processes afterwards.
#pragma once
#include "stack_sorted.cuh"
#include "reg.cuh"

namespace sgpu {

// pixel (within the tile) of group g of warp w: chosen so that the T lanes of a group and the
// 32/T groups of a warp hit 32 different banks when they read rows j*T + t
template <int LT>
__device__ __forceinline__ int reg_pixel(int warp, int g) {
	if (LT == 3) return (warp << 2) | g;					// 8 warps x 4 px
	if (LT == 2) return ((warp + 4 * (g >> 2)) << 2) | (g & 3);		// 4 warps x 8 px: 16-byte chunks w and w + 4
	return 0;
}

// lane-private list entry e lives in row e*T + t of the pixel's column.  Byte address =
// row*128 + (((px >> 2) ^ (row & 7)) << 4) + (px & 3)*4 with row & 7 = ((e*T) & 7) | t (T <= 8):
// la = t*128 + (((px >> 2) ^ t) & 7) << 4 | (px & 3) << 2 is fixed per lane, the rest depends on e only
template <int T>
struct TileSt {
	char *la; float *wb;
	__device__ __forceinline__ float *at(int e) const { return (float *) ((char *) ((uintptr_t) la ^ (uintptr_t) (((e * T) & 7) << 4)) + e * (T * 128)); }
	__device__ __forceinline__ void put(int e, float x) { *at(e) = x; }
	__device__ __forceinline__ float get(int e) const { return *at(e); }
	__device__ __forceinline__ float &w(int i) { return wb[i]; }
	__device__ __forceinline__ float w(int i) const { return wb[i]; }
	__device__ __forceinline__ float &sw(int i) { return wb[REG_WCAP + i]; }
	__device__ __forceinline__ float sw(int i) const { return wb[REG_WCAP + i]; }
};

template <int LT, int RE, int RF, int MINB, bool WINSOR>
__global__ void __launch_bounds__(32 << LT, MINB) stack_reg_kernel(const StackParams p, const __grid_constant__ CUtensorMap tmap) {
	constexpr int T = 1 << LT, Q = 32 >> LT;
	extern __shared__ __align__(1024) unsigned char smem_raw[];
	const int N = p.N;
	float *rows = (float *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
	float *wbuf = rows + (size_t) p.tile_rows * TILE_PX;			// [32 px][2 * REG_WCAP]
	double2 *coef = (double2 *) (wbuf + TILE_PX * 2 * REG_WCAP);		// [N] {scale, offset | mul} when the tile holds raw samples
	unsigned long long *bar = (unsigned long long *) (coef + (p.use_tma && p.normalize != SGPU_NO_NORM ? N : 0));
	const int layer = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t npix = (size_t) p.H * p.W, pix0 = (size_t) blockIdx.x * TILE_PX;
	const int tma_mode = !p.use_tma ? 0 : (p.normalize == SGPU_NO_NORM ? 0 : ((p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING) ? 1 : 2));

	{	// TMA only (frames with per-frame x shifts go to the lane-sort engine)
		if (threadIdx.x == 0) mbar_init(bar, 1);
		__syncthreads();
		if (threadIdx.x == 0) {
			mbar_expect_tx(bar, (unsigned) (p.tile_rows * TILE_PX * sizeof(float)));
			for (int r0 = 0; r0 < p.tile_rows; r0 += p.box_rows)
				tma_load_3d(rows + (size_t) r0 * TILE_PX, &tmap, (int) pix0, layer, r0, bar);
		}
		if (tma_mode) {	// per-frame coefficients, staged while the tile is in flight
			const double *c0 = p.scale + layer * N, *c1 = (tma_mode == 1 ? p.offset : p.mul) + layer * N;
			for (int i = threadIdx.x; i < N; i += 32 * T) coef[i] = make_double2(c0[i], c1[i]);
			__syncthreads();
		}
		mbar_wait(bar, 0);
	}

	const int t = lane & (T - 1), gi = lane >> LT;
	const int px = reg_pixel<LT>(warp, gi);
	Group<T> g; g.t = t; g.mask = (T == 32) ? 0xffffffffu : (((1u << T) - 1u) << (lane & ~(T - 1)));
	TileSt<T> st; st.wb = wbuf + px * 2 * REG_WCAP;
	st.la = (char *) rows + ((unsigned) (t * 128) | ((unsigned) (((px >> 2) ^ t) & 7) << 4) | ((unsigned) (px & 3) << 2));

	// ---- gather: samples j*T + t -> registers ----
	float v[RE];
	int nullish = 0;
#pragma unroll
	for (int j = 0; j < RE; j++) {
		const int i = j * T + t;
		float x = reg_nan();
		if (j < RF || i < N) {
			x = st.get(j);
			if (tma_mode) { const double2 c = coef[i]; x = tma_mode == 1 ? normalize_sample<float, 1>(x, c.x, c.y) : normalize_sample<float, 2>(x, c.x, c.y); }
			nullish |= (x == 0.f);
		}
		v[j] = x;
	}
	__syncwarp();	// the column is scratch from here on; only this pixel's group touches it

	const size_t pix = pix0 + px;
	int rej[2] = { 0, 0 };
	int status = FAST_DONE;
	double result = 0.0;
	if (pix < npix) {
		if (p.is_mean) {
			if (g.any(nullish)) status = FAST_FALLBACK;	// null samples are dropped first (rejection_float.c:128-135): exact engine
			else status = reg_mean_and_reject<T, RE, RF, WINSOR>(v, g, N, p.rc, st, rej, &result);
		} else status = reg_median<T, RE, RF>(v, g, N, st, &result);
		if (t == 0) {
			if (status == FAST_DONE) store_result<float>(p, (size_t) layer * npix + pix, result, rej);
			else p.list[atomicAdd(p.list_count, 1u)] = ((unsigned) layer << 30) | (unsigned) pix;
		}
	}
	if (p.is_mean && p.rc.type != SGPU_NO_REJEC) {
		const bool mine = (t == 0 && pix < npix && status == FAST_DONE);
		unsigned lo = mine ? rej[0] : 0, hi = mine ? rej[1] : 0;
		__syncwarp();
		lo = __reduce_add_sync(0xffffffffu, lo); hi = __reduce_add_sync(0xffffffffu, hi);
		if (lane == 0 && (lo | hi)) {
			atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
		}
	}
}

}  // namespace sgpu

template <int LT, int RE, int RF, int MINB, bool WINSOR>
int sgpu_launch_reg_variant(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm);

#define SGPU_REG_VARIANT(LT, RE, RF, MINB, WINSOR) \
	template <> int sgpu_launch_reg_variant<LT, RE, RF, MINB, WINSOR>(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm) { \
		auto kern = sgpu::stack_reg_kernel<LT, RE, RF, MINB, WINSOR>; \
		const CUtensorMap tm = tmap ? *tmap : CUtensorMap{}; \
		if (shm > 48 * 1024) SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm)); \
		kern<<<grid, 32 << LT, shm, ctx->stream>>>(p, tm); \
		ctx->launches++; \
		SGPU_CUDA(ctx, cudaGetLastError()); \
		return SGPU_OK; \
	}
