// stack_sorted.cu — the sorted engine: the fast path for mean-with-rejection (none, percentile,
// sigma, winsorized[f32], linear-fit, GESDT) and for the median stack.
//
// One CTA (4 warps) owns a tile of 32 consecutive pixels of one layer.
//  load   every warp streams whole 128-byte rows: lane = pixel, one frame per instruction, so the
//         frame-major block is read with fully coalesced HBM requests exactly once.  Each sample
//         is normalized in double as median_and_mean.c:1644-1686 does and parked, as a float, in
//         shared memory tile[pixel][frame] (odd row pitch: conflict-free).
//  sort   (phase A) a group of T = 2^LT lanes owns one pixel and holds its stack in registers,
//         R = 2^LR samples per lane, padded with +inf.  A bitonic network over the T*R slots
//         sorts it: compare-exchanges at distance < R are register-to-register, the others are
//         one SHFL.xor + one FMNMX per slot.  Zeros (null pixels, rejection_float.c:128-135) are
//         turned into +inf first in mean mode, so they sort to the end and drop out.
//         The sorted stack goes back to the pixel's shared-memory row (rank i at i + q*(i>>LR):
//         the skew keeps the 8 lanes' stores on different banks).
//  reject (phase B) warp 0 runs one THREAD per pixel over the sorted stacks (sorted.cuh) and
//         writes the 32 results / rejection-map entries as coalesced rows.  Pixels it cannot
//         decide identically to the reference are appended to a list that the exact engine
//         (stack_exact.cu, LIST mode) processes right after on the same stream.
#include "kparams.cuh"
#include "sorted.cuh"

namespace sgpu {

constexpr int TILE_PX = 32;
constexpr int SORT_WARPS = 4;

// ascending bitonic sort of the T*R slots (slot i = t*R + j) held by the T lanes of a group
template <int LT, int LR>
__device__ __forceinline__ void group_sort(float (&v)[1 << LR], int t) {
	constexpr int R = 1 << LR, NS = R << LT;
#pragma unroll
	for (int k = 2; k <= NS; k <<= 1) {
		// direction of the size-k blocks: ascending where bit k of the slot index is clear
		const bool up_lane = (k >= NS) ? true : (k >= R ? ((t & (k >> LR)) == 0) : true);
#pragma unroll
		for (int s = k >> 1; s > 0; s >>= 1) {
			if (s >= R) {	// partner slot lives in lane t ^ (s / R)
				const int lm = s >> LR;
				const bool keep_min = ((t & lm) == 0) == up_lane;
#pragma unroll
				for (int j = 0; j < R; j++) {
					const float b = __shfl_xor_sync(0xffffffffu, v[j], lm);
					v[j] = keep_min ? fminf(v[j], b) : fmaxf(v[j], b);
				}
			} else {
#pragma unroll
				for (int j = 0; j < R; j++) {
					if ((j & s) == 0) {
						const bool up = (k < R) ? ((j & k) == 0) : up_lane;
						const float a = v[j], b = v[j | s];
						v[j] = up ? fminf(a, b) : fmaxf(a, b);
						v[j | s] = up ? fmaxf(a, b) : fminf(a, b);
					}
				}
			}
		}
	}
}

template <int LT, int LR>
struct SmemStack {	// a pixel's sorted stack in shared memory
	float *row;
	__device__ __forceinline__ float operator()(int i) const { return row[i + ((i >> LR) << (5 - LT))]; }
	__device__ __forceinline__ void set(int i, float v) const { row[i + ((i >> LR) << (5 - LT))] = v; }
};

template <typename T, int LT, int LR>
__global__ void __launch_bounds__(SORT_WARPS * 32, (LR >= 6 ? 3 : (LR == 5 ? 4 : 6))) stack_sorted_kernel(const StackParams p) {
	constexpr int TL = 1 << LT, R = 1 << LR, Q = 32 >> LT;	// lanes per pixel, slots per lane, pixels per warp
	constexpr bool U16 = sizeof(T) == 2;
	extern __shared__ __align__(16) float tile[];	// [32][ld], then int cnt[32], int bad[32]
	const int ld = p.ld, N = p.N;
	int *cnt = (int *) (tile + TILE_PX * ld), *bad = cnt + TILE_PX;
	const int layer = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t npix = (size_t) p.H * p.W, pix0 = (size_t) blockIdx.x * TILE_PX;

	// ---- load: lane = pixel, warp w takes frames w, w + 4, ... ----
	{
		const size_t pix = pix0 + lane;
		const bool in = pix < npix;
		const int x = in ? (int) (pix % p.W) : 0;
		float *dst = tile + lane * ld;
		int f = warp;
		for (; f + 3 * SORT_WARPS < N; f += 4 * SORT_WARPS) {	// 4 independent loads in flight
			T v0 = (T) 0, v1 = (T) 0, v2 = (T) 0, v3 = (T) 0;
			if (in) {
				v0 = load_norm<T>(p, f, layer, pix, x); v1 = load_norm<T>(p, f + SORT_WARPS, layer, pix, x);
				v2 = load_norm<T>(p, f + 2 * SORT_WARPS, layer, pix, x); v3 = load_norm<T>(p, f + 3 * SORT_WARPS, layer, pix, x);
			}
			dst[f] = (float) v0; dst[f + SORT_WARPS] = (float) v1; dst[f + 2 * SORT_WARPS] = (float) v2; dst[f + 3 * SORT_WARPS] = (float) v3;
		}
		for (; f < N; f += SORT_WARPS) dst[f] = in ? (float) load_norm<T>(p, f, layer, pix, x) : 0.f;
	}
	__syncthreads();

	// ---- phase A: sort, Q pixels per warp and round ----
	{
		const int t = lane & (TL - 1), g = lane >> LT;
#pragma unroll 1
		for (int r = 0; r < TILE_PX / (Q * SORT_WARPS); r++) {
			const int px = r * (Q * SORT_WARPS) + warp * Q + g;
			float *row = tile + px * ld;
			float v[R];
			int n = 0, weird = 0;
#pragma unroll
			for (int j = 0; j < R; j++) {	// frame 32*(j/Q) + Q*t + j%Q: the group's lanes hit distinct banks
				const int f = 32 * (j / Q) + Q * t + (j % Q);
				float x = __int_as_float(0x7f800000);
				if (f < N) {
					x = row[f];
					if (!(fabsf(x) < __int_as_float(0x7f800000))) { weird = 1; x = __int_as_float(0x7f800000); }
					else if (p.is_mean && x == 0.f) x = __int_as_float(0x7f800000);
					else n++;
				}
				v[j] = x;
			}
#pragma unroll
			for (int m = 1; m < TL; m <<= 1) { n += __shfl_xor_sync(0xffffffffu, n, m); weird |= __shfl_xor_sync(0xffffffffu, weird, m); }
			group_sort<LT, LR>(v, t);
			__syncwarp();
			if (t == 0) { cnt[px] = n; bad[px] = weird; }
			SmemStack<LT, LR> st{ row };
#pragma unroll
			for (int j = 0; j < R; j++) { const int i = t * R + j; if (i < N) st.set(i, v[j]); }
		}
	}
	__syncthreads();
	if (warp != 0) return;

	// ---- phase B: one thread per pixel ----
	const size_t pix = pix0 + lane;
	int rej[2] = { 0, 0 };
	if (pix < npix) {
		SmemStack<LT, LR> st{ tile + lane * ld };
		const int n = cnt[lane];
		int status = bad[lane] ? FAST_FALLBACK : FAST_DONE;
		double result = 0.0;
		if (status == FAST_DONE) {
			if (p.is_mean) {
				const int x = (int) (pix % p.W);
				const double *w = p.weights ? p.weights + (size_t) layer * N : nullptr;
				auto orig = [&](int ff) { return (float) load_norm<T>(p, ff, layer, pix, x); };
				result = fast_mean_and_reject<U16>(st, n, N, p.rc, w, orig, rej, &status);
			} else {
				result = median_sorted<U16>(st, 0, N);	// all N samples, zeros included (:1711-1715)
			}
		}
		if (status == FAST_DONE) store_result<T>(p, (size_t) layer * npix + pix, result, rej);
		else {
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

using namespace sgpu;

bool sgpu_sorted_supports(const sgpu_block_desc &d, const sgpu_stack_args *a) {
	if (d.nb_frames > 2048 || (size_t) d.height * d.width >= (1u << 30)) return false;
	if (!a->is_mean) return true;
	switch (a->type_of_rejection) {
	case SGPU_NO_REJEC: case SGPU_PERCENTILE: case SGPU_SIGMA: case SGPU_LINEARFIT: case SGPU_GESDT: return true;
	case SGPU_WINSORIZED: return d.dtype == SGPU_F32;
	default: return false;	// MAD, SIGMEDIAN: exact engine
	}
}

template <typename T, int LT, int LR>
static int launch_sorted(sgpu_ctx *ctx, const StackParams &p, dim3 grid, size_t shm) {
	auto kern = stack_sorted_kernel<T, LT, LR>;
	if (shm > 48 * 1024)
		SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm));
	kern<<<grid, SORT_WARPS * 32, shm, ctx->stream>>>(p);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}

template <typename T>
static int dispatch_sorted(sgpu_ctx *ctx, const StackParams &p, dim3 grid, size_t shm) {
	const int N = p.N;
	if (N <= 32) return launch_sorted<T, 3, 2>(ctx, p, grid, shm);
	if (N <= 64) return launch_sorted<T, 3, 3>(ctx, p, grid, shm);
	if (N <= 128) return launch_sorted<T, 3, 4>(ctx, p, grid, shm);
	if (N <= 256) return launch_sorted<T, 3, 5>(ctx, p, grid, shm);
	if (N <= 512) return launch_sorted<T, 3, 6>(ctx, p, grid, shm);
	if (N <= 1024) return launch_sorted<T, 4, 6>(ctx, p, grid, shm);
	return launch_sorted<T, 5, 6>(ctx, p, grid, shm);
}

int sgpu_launch_sorted(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej) {
	const sgpu_block_desc &d = blk->d;
	const size_t npix = (size_t) d.height * d.width;
	if (npix == 0) return SGPU_OK;
	StackParams p;
	sgpu_fill_stack_params(p, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	// fallback list: one counter + one entry per pixel at most, at the start of the scratch area
	const size_t list_bytes = ((npix * d.nb_layers + 1) * sizeof(unsigned) + 255) & ~(size_t) 255;
	int rv = sgpu_ensure_list(ctx, list_bytes); if (rv) return rv;
	p.list_count = (unsigned *) ctx->d_list; p.list = p.list_count + 1;
	SGPU_CUDA(ctx, cudaMemsetAsync(p.list_count, 0, sizeof(unsigned), ctx->stream));
	p.ld = (d.nb_frames + 32) | 1;
	const size_t shm = (size_t) TILE_PX * p.ld * sizeof(float) + 2 * TILE_PX * sizeof(int);
	if (shm > ctx->smem_optin) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "stack too deep for the sorted engine");
	dim3 grid((unsigned) ((npix + TILE_PX - 1) / TILE_PX), (unsigned) d.nb_layers);
	rv = (d.dtype == SGPU_F32) ? dispatch_sorted<float>(ctx, p, grid, shm) : dispatch_sorted<u16>(ctx, p, grid, shm);
	if (rv) return rv;
	return sgpu_launch_exact_list(ctx, blk, p);
}
