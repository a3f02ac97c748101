// stack_sorted.cuh — kernel of the sorted engine: the fast path for mean-with-rejection (none,
// percentile, sigma, winsorized[f32], linear-fit, GESDT) and for the median stack.
//
// One CTA (4 warps) owns a tile of 32 consecutive pixels of one layer; its shared-memory tile is
// rows[i][32 px] (128 bytes per row) with the 16-byte chunks of a row XOR-swizzled by (i & 7).
//  load   every warp streams whole 128-byte rows: lane = pixel, one frame per instruction, so the
//         frame-major block is read with fully coalesced HBM requests exactly once.  Each sample
//         is normalized in double as median_and_mean.c:1644-1686 does and parked as a float in
//         row f of the tile.
//  sort   (phase A) a group of T = 2^LT lanes owns one pixel and holds its stack in registers, RE
//         samples per lane: slot j of lane t is row j*T + t (bank-conflict free through the
//         swizzle), padded with +inf.  A bitonic network in the all-ascending ("flip") form
//         sorts the T*RE slots: compare-exchanges between registers of one lane are two FMNMX,
//         those between lanes one SHFL + one FMNMX per slot.  Because every comparator is
//         ascending the +inf padding never moves, so comparators that touch a register index
//         >= RE are dropped at compile time and RE need not be a power of two.  Zeros (null
//         pixels, rejection_float.c:128-135) are turned into +inf first in mean mode, so they
//         sort to the end and drop out.  The sorted stack overwrites the pixel's column: row i
//         now holds rank i.  The group also leaves sum(x-K), sum((x-K)^2) of the valid samples.
//  reject (phase B) warp 0 runs one THREAD per pixel over the sorted stacks (sorted.cuh) and
//         writes the 32 results / rejection-map entries as coalesced rows.  Pixels it cannot
//         decide identically to the reference are appended to a list that the exact engine
//         (stack_exact.cu, LIST mode) processes right after on the same stream.
#pragma once
#include <cuda.h>
#include "kparams.cuh"
#include "sorted.cuh"

namespace sgpu {

constexpr int TILE_PX = 32;
#define SGPU_INF __int_as_float(0x7f800000)

__device__ __forceinline__ int sw_idx(int row, int px) { return row * TILE_PX + ((((px >> 2) ^ (row & 7)) << 2) | (px & 3)); }

// ---- TMA (cp.async.bulk.tensor) + mbarrier, raw PTX ----
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
	// bounded: a tile that never arrives (bad tensor map) must fault, not hang the GPU
	for (unsigned spin = 0; spin < (1u << 24); spin++) {
		unsigned done;
		asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
		if (done) return;
	}
	__trap();
}
// box (32 px, 1 layer, rows) of the [frame][layer][pixel] tensor at (pix0, layer, row0)
__device__ __forceinline__ void tma_load_3d(void *dst, const void *tmap, int c0, int c1, int c2, unsigned long long *bar) {
	asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
		:: "r"(smem_u32(dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}

constexpr int pow2_ceil(int x) { int p = 1; while (p < x) p <<= 1; return p; }

// one compare-exchange step between lanes t and t ^ lm: the lane with `keep_min` keeps the
// minimum of each slot pair, the other the maximum
template <int RE>
__device__ __forceinline__ void lane_step(float (&v)[RE], int lm, bool keep_min) {
#pragma unroll
	for (int j = 0; j < RE; j++) {
		const float r = __shfl_xor_sync(0xffffffffu, v[j], lm);
		v[j] = keep_min ? fminf(v[j], r) : fmaxf(v[j], r);
	}
}

// ascending sort of the T*RE slots held by the T lanes of a group; slot (j, t) has index j*T + t
template <int LT, int RE>
__device__ __forceinline__ void group_sort(float (&v)[RE], int t) {
	constexpr int T = 1 << LT, NS = pow2_ceil(RE * T);
#pragma unroll
	for (int k = 2; k <= NS; k <<= 1) {
		// flip step: index i against i ^ (k - 1), minimum to the lower index
		if (k <= T) {
			lane_step<RE>(v, k - 1, (t & (k >> 1)) == 0);
		} else {
			const int jm = (k >> LT) - 1, hb = k >> (LT + 1);	// register flip mask, the bit that tells lower from upper
#pragma unroll
			for (int j = 0; j < RE; j++) {
				if ((j & hb) == 0 && (j ^ jm) < RE) {
					const int pj = j ^ jm;
					const float a = v[j], b = v[pj];
					const float x = __shfl_xor_sync(0xffffffffu, b, T - 1);	// partner lane's upper slot
					const float y = __shfl_xor_sync(0xffffffffu, a, T - 1);	// partner lane's lower slot
					v[j] = fminf(a, x); v[pj] = fmaxf(b, y);
				}
			}
		}
#pragma unroll
		for (int s = k >> 2; s > 0; s >>= 1) {
			if (s >= T) {
				const int js = s >> LT;
#pragma unroll
				for (int j = 0; j < RE; j++) {
					if ((j & js) == 0 && (j | js) < RE) {
						const float a = v[j], b = v[j | js];
						v[j] = fminf(a, b); v[j | js] = fmaxf(a, b);
					}
				}
			} else {
				lane_step<RE>(v, s, (t & s) == 0);
			}
		}
	}
}

struct TileStack {	// one pixel's sorted stack inside the tile
	float *rows; int px;
	__device__ __forceinline__ float operator()(int i) const { return rows[sw_idx(i, px)]; }
	__device__ __forceinline__ void set(int i, float v) const { rows[sw_idx(i, px)] = v; }
};

// normalization of one raw sample (median_and_mean.c:1644-1686); MODE 0 none, 1 additive family,
// 2 multiplicative family; c0 = scale, c1 = offset or mul
template <typename T, int MODE>
__device__ __forceinline__ float normalize_sample(T v, double c0, double c1) {
	if (MODE == 0) return (float) v;
	if (MODE == 1) {
		if (v == (T) 0) return 0.f;
		const double tmp = (double) v * c0;
		return sizeof(T) == 4 ? (float) (tmp - c1) : (float) round_to_word(tmp - c1);
	}
	const double tmp = (double) v * c0;
	return sizeof(T) == 4 ? (float) (tmp * c1) : (float) round_to_word(tmp * c1);
}

// load + normalize the tile: lane = pixel, warp w takes frames w, w + 4, ...  All raw loads of a
// batch are issued before the first one is used, so U of them are in flight per thread.
template <typename T, int MODE, bool SHIFT, int SORT_WARPS>
__device__ __forceinline__ void load_tile(const StackParams &p, float *rows, int layer, size_t pix0, size_t npix, int warp, int lane) {
	const size_t pix = pix0 + lane;
	const bool in = pix < npix;
	const int x = in ? (int) (pix % p.W) : 0;
	const int N = p.N;
	const size_t fstride = (size_t) p.C * npix;	// samples between the same pixel of consecutive frames
	const T *src = (const T *) p.frames + (size_t) layer * npix + (in ? pix : 0);
	const double *c0 = p.scale + layer * N, *c1 = (MODE == 1 ? p.offset : p.mul) + layer * N;
	constexpr int U = 10;
	for (int f0 = warp; f0 < N; f0 += U * SORT_WARPS) {
		T raw[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			const int f = f0 + u * SORT_WARPS;
			raw[u] = (T) 0;
			if (f < N) {
				if (SHIFT) {
					const int sx = p.shiftx[f], xs = x - sx;
					if (in && xs >= 0 && xs < p.W) raw[u] = __ldg(src + (size_t) f * fstride - sx);
				} else if (in) raw[u] = __ldg(src + (size_t) f * fstride);
			}
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
			const int f = f0 + u * SORT_WARPS;
			if (f < N) rows[sw_idx(f, lane)] = normalize_sample<T, MODE>(raw[u], MODE ? c0[f] : 0.0, MODE ? c1[f] : 0.0);
		}
	}
}

template <typename T, int NW>
__device__ __forceinline__ void load_tile_any(const StackParams &p, float *rows, int layer, size_t pix0, size_t npix, int warp, int lane) {
	const int mode = p.normalize == SGPU_NO_NORM ? 0 : ((p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING) ? 1 : 2);
	if (p.shiftx) {
		if (mode == 0) load_tile<T, 0, true, NW>(p, rows, layer, pix0, npix, warp, lane);
		else if (mode == 1) load_tile<T, 1, true, NW>(p, rows, layer, pix0, npix, warp, lane);
		else load_tile<T, 2, true, NW>(p, rows, layer, pix0, npix, warp, lane);
	} else {
		if (mode == 0) load_tile<T, 0, false, NW>(p, rows, layer, pix0, npix, warp, lane);
		else if (mode == 1) load_tile<T, 1, false, NW>(p, rows, layer, pix0, npix, warp, lane);
		else load_tile<T, 2, false, NW>(p, rows, layer, pix0, npix, warp, lane);
	}
}

template <int LT, int RE, int SORT_WARPS>
__global__ void __launch_bounds__(SORT_WARPS * 32, (RE > 44 ? 12 : (RE > 25 ? 16 : 24)) / SORT_WARPS) stack_sorted_kernel(const StackParams p, const __grid_constant__ CUtensorMap tmap) {
	constexpr int TL = 1 << LT, Q = 32 >> LT;	// lanes per pixel, pixels per warp and round
	extern __shared__ __align__(1024) unsigned char smem_raw[];	// rows[..][32] (1024-aligned: TMA swizzle), Sums[32], cnt[32], bad[32], mbarrier
	const int N = p.N;
	float *rows = (float *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
	Sums *sums = (Sums *) (rows + (size_t) p.tile_rows * TILE_PX);
	int *cnt = (int *) (sums + TILE_PX), *bad = cnt + TILE_PX;
	unsigned long long *bar = (unsigned long long *) (bad + TILE_PX);
	const int layer = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t npix = (size_t) p.H * p.W, pix0 = (size_t) blockIdx.x * TILE_PX;

	if (p.use_tma) {
		// one thread has the TMA unit fetch the whole (32 px x N frames) tile: raw samples, rows
		// swizzled exactly as sw_idx() expects; normalization happens when phase A reads them
		if (threadIdx.x == 0) mbar_init(bar, 1);
		__syncthreads();
		if (threadIdx.x == 0) {
			mbar_expect_tx(bar, (unsigned) (p.tile_rows * TILE_PX * sizeof(float)));
			for (int r0 = 0; r0 < p.tile_rows; r0 += p.box_rows)
				tma_load_3d(rows + (size_t) r0 * TILE_PX, &tmap, (int) pix0, layer, r0, bar);
		}
		mbar_wait(bar, 0);
	} else {
		if (p.is_u16) load_tile_any<u16, SORT_WARPS>(p, rows, layer, pix0, npix, warp, lane);
		else load_tile_any<float, SORT_WARPS>(p, rows, layer, pix0, npix, warp, lane);
		__syncthreads();
	}
	const int tma_mode = !p.use_tma ? 0 : (p.normalize == SGPU_NO_NORM ? 0 : ((p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING) ? 1 : 2));
	const double *nc0 = p.scale + layer * N, *nc1 = (tma_mode == 1 ? p.offset : p.mul) + layer * N;

	// ---- phase A: sort, Q pixels per warp and round ----
	{
		const int t = lane & (TL - 1), g = lane >> LT;
#pragma unroll 1
		for (int r = 0; r < TILE_PX / (Q * SORT_WARPS); r++) {
			const int px = r * (Q * SORT_WARPS) + warp * Q + g;
			float v[RE];
			int n = 0, weird = 0;
#pragma unroll
			for (int j = 0; j < RE; j++) {
				const int i = j * TL + t;
				float x = SGPU_INF;
				if (i < N) {
					x = rows[sw_idx(i, px)];
					if (tma_mode == 1) x = normalize_sample<float, 1>(x, nc0[i], nc1[i]);
					else if (tma_mode == 2) x = normalize_sample<float, 2>(x, nc0[i], nc1[i]);
					if (!(fabsf(x) < SGPU_INF)) { weird = 1; x = SGPU_INF; }
					else if (p.is_mean && x == 0.f) x = SGPU_INF;
					else n++;
				}
				v[j] = x;
			}
#pragma unroll
			for (int m = 1; m < TL; m <<= 1) { n += __shfl_xor_sync(0xffffffffu, n, m); weird |= __shfl_xor_sync(0xffffffffu, weird, m); }
			group_sort<LT, RE>(v, t);
			if (p.need_sums) {	// sum(x - K), sum((x - K)^2) over the n valid samples; K: the sample of rank n / 2 (warp-uniform branch: shuffles inside)
				const int mid = n >> 1;
				float Kf = 0.f;
#pragma unroll
				for (int j = 0; j < RE; j++) if (j == (mid >> LT)) Kf = v[j];
				Kf = __shfl_sync(0xffffffffu, Kf, (lane & ~(TL - 1)) | (mid & (TL - 1)));
				const double K = n > 0 ? (double) Kf : 0.0;
				double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
#pragma unroll
				for (int j = 0; j < RE; j++) {
					const double y = (j * TL + t < n) ? (double) v[j] - K : 0.0;
					if (j & 1) { b1 += y; b2 = fma(y, y, b2); } else { a1 += y; a2 = fma(y, y, a2); }
				}
				a1 += b1; a2 += b2;
#pragma unroll
				for (int m = 1; m < TL; m <<= 1) { a1 += __shfl_xor_sync(0xffffffffu, a1, m); a2 += __shfl_xor_sync(0xffffffffu, a2, m); }
				if (t == 0) { sums[px].K = K; sums[px].S1 = a1; sums[px].S2 = a2; }
			}
			__syncwarp();
			if (t == 0) { cnt[px] = n; bad[px] = weird; }
#pragma unroll
			for (int j = 0; j < RE; j++) { const int i = j * TL + t; if (i < N) rows[sw_idx(i, px)] = v[j]; }
		}
	}
	__syncthreads();
	if (warp != 0) return;

	// ---- phase B: one thread per pixel ----
	const size_t pix = pix0 + lane;
	int rej[2] = { 0, 0 };
	if (pix < npix) {
		TileStack st{ rows, lane };
		const int n = cnt[lane];
		int status = bad[lane] ? FAST_FALLBACK : FAST_DONE;
		double result = 0.0;
		if (status == FAST_DONE) {
			if (p.is_mean) {
				const int x = (int) (pix % p.W);
				const double *fw = p.weights ? p.weights + (size_t) layer * N : nullptr;
				// frame weights, and the pixel's feather-mask samples where the block carries masks (median_and_mean.c:1060-1066)
				const PixWeights w = { fw, p.mask ? p.mask + pix : nullptr, nullptr, npix, 0, p.shiftx };
				const Sums *S0 = (p.need_sums && n > 0) ? &sums[lane] : nullptr;
				if (p.is_u16) {
					auto orig = [&](int ff) { return (float) load_norm<u16>(p, ff, layer, pix, x); };
					result = fast_mean_and_reject<true>(st, n, N, p.rc, w, orig, S0, rej, &status);
				} else {
					auto orig = [&](int ff) { return load_norm<float>(p, ff, layer, pix, x); };
					result = fast_mean_and_reject<false>(st, n, N, p.rc, w, orig, S0, rej, &status);
				}
			} else {	// all N samples, zeros included (median_and_mean.c:1711-1715)
				result = p.is_u16 ? median_sorted<true>(st, 0, N) : median_sorted<false>(st, 0, N);
			}
		}
		if (status == FAST_DONE) {
			if (p.is_u16) store_result<u16>(p, (size_t) layer * p.out_plane + pix, result, rej);
			else store_result<float>(p, (size_t) layer * p.out_plane + pix, result, rej);
		} else {
			rej[0] = rej[1] = 0;
			p.list[atomicAdd(p.list_count, 1u)] = ((unsigned) layer << 30) | (unsigned) pix;
		}
	}
	if (p.is_mean && p.rc.type != SGPU_NO_REJEC) {
		unsigned lo = rej[0], hi = rej[1];
		lo = __reduce_add_sync(0xffffffffu, lo); hi = __reduce_add_sync(0xffffffffu, hi);
		if (lane == 0 && (lo | hi)) {
			atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
		}
	}
}

}  // namespace sgpu

// one instantiation per (lanes per pixel, slots per lane, warps per CTA); defined in stack_sorted_inst*.cu
template <int LT, int RE, int NW>
int sgpu_launch_sorted_variant(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm);

#define SGPU_SORTED_VARIANT(LT, RE, NW) \
	template <> int sgpu_launch_sorted_variant<LT, RE, NW>(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm) { \
		auto kern = sgpu::stack_sorted_kernel<LT, RE, NW>; \
		const CUtensorMap tm = tmap ? *tmap : CUtensorMap{}; \
		if (shm > 48 * 1024) SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm)); \
		kern<<<grid, NW * 32, shm, ctx->stream>>>(p, tm); \
		ctx->launches++; \
		SGPU_CUDA(ctx, cudaGetLastError()); \
		return SGPU_OK; \
	}
