// simple.cu — the sum / max / min stackers on a resident block (K5):
//   sum_stacking_image_hook + sum_stacking_finalize_hook (src/stacking/sum.c:107-237, :240-372)
//   minmax_stacking_image_hook (src/stacking/stackminmax.c:49-84, :103-184)
// One thread per output pixel walks the frames of the block (each frame plane is read once,
// coalesced; eight loads in flight per thread).  A pixel that falls outside a frame after the
// registration shift is SKIPPED (sum.c:190-193, stackminmax.c:164-166), it is not a zero: the
// block's rows were shifted (and zero filled) at upload, so the kernel re-derives which rows are
// real from shifty, the block's first image row and the frame height.
// Sums: double for float samples, 64-bit integers for WORD samples, frames added in index order
// (the reference adds them with `omp atomic` in whatever order its workers finish).  The output
// conversion needs the maximum over the whole image, so it is a second call (row shards reduce
// the maximum in between).
#include <algorithm>
#include <cfloat>
#include <cstdlib>
#include "common.cuh"

namespace sgpu {

typedef uint16_t u16;

struct SimpleGeom {
	int N, C, H, W;		// block: frames, layers, rows held, width
	int row0, frame_h;	// image row of block row 0, full frame height
	const int *shiftx, *shifty;	// [N] device, or nullptr
};

__device__ __forceinline__ u16 round_to_word_d(double x) {	// proto.h:232-237
	x = x + 0.5;
	x = (x > 65535.0) ? 65535.0 : x;
	x = (x < 0.0) ? 0.0 : x;
	return (u16) x;
}

// OP 0: sum (Acc = double | unsigned long long), 1: max, 2: min (Acc = T)
template <typename T, typename Acc, int OP>
__global__ void __launch_bounds__(256) simple_stack_kernel(const T *frames, SimpleGeom g, Acc *out, double *blockmax) {
	const size_t np = (size_t) g.H * g.W, total = np * g.C;
	const size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	Acc acc;
	if (OP == 0) acc = (Acc) 0;
	else if (OP == 1) acc = (Acc) 0;
	else acc = sizeof(T) == 4 ? (Acc) 1.0f : (Acc) 65535;
	if (idx < total) {
		const int layer = (int) (idx / np);
		const size_t pix = idx - (size_t) layer * np;
		const int r = (int) (pix / g.W), x = (int) (pix - (size_t) r * g.W);
		const size_t fstride = np * g.C;
		const T *src = frames + (size_t) layer * np + (size_t) r * g.W;
		for (int f0 = 0; f0 < g.N; f0 += 8) {
			T v[8]; bool ok[8];
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int f = f0 + k;
				ok[k] = false;
				if (f < g.N) {
					const int sx = g.shiftx ? g.shiftx[f] : 0, sy = g.shifty ? g.shifty[f] : 0;
					const int nx = x - sx, ny = g.row0 + r - sy;
					ok[k] = nx >= 0 && nx < g.W && ny >= 0 && ny < g.frame_h;
					if (ok[k]) v[k] = src[(size_t) f * fstride + nx];
				}
			}
#pragma unroll
			for (int k = 0; k < 8; k++) if (ok[k]) {
				if (OP == 0) acc += (Acc) v[k];
				else if (OP == 1) { if (v[k] > (T) acc) acc = (Acc) v[k]; }
				else { if (v[k] < (T) acc) acc = (Acc) v[k]; }
			}
		}
		out[idx] = acc;
	}
	if (OP == 0) {	// maximum of the sums of this CTA (sum.c:256-278 starts from 0)
		__shared__ double sh[8];
		double m = idx < total ? (double) acc : 0.0;
		for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
		if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
		__syncthreads();
		if (threadIdx.x == 0) { for (int k = 1; k < 8; k++) m = fmax(m, sh[k]); blockmax[blockIdx.x] = m; }
	}
}

// The same walk for V = 16 / sizeof(T) adjacent pixels of one row per thread: one 16-byte load per
// frame and thread instead of a 2- or 4-byte one (the WORD kernel above moves 64 bytes per warp
// load).  Needs no x shift, a width that is a multiple of V and a 16-byte aligned block; each
// pixel's samples are still combined in frame order, so the results are the scalar kernel's.
template <typename T, typename Acc, int OP>
__global__ void __launch_bounds__(256) simple_stack_vec_kernel(const T *frames, SimpleGeom g, Acc *out, double *blockmax) {
	constexpr int V = 16 / sizeof(T), F = sizeof(T) == 4 ? 8 : 4;
	const size_t np = (size_t) g.H * g.W, total = np * g.C, nv = total / V;
	const size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	Acc acc[V];
#pragma unroll
	for (int j = 0; j < V; j++) acc[j] = OP == 2 ? (sizeof(T) == 4 ? (Acc) 1.0f : (Acc) 65535) : (Acc) 0;
	if (idx < nv) {
		const size_t e = idx * V;
		const int layer = (int) (e / np);
		const int r = (int) ((e - (size_t) layer * np) / g.W);
		const size_t fstride = np * g.C;
		const T *src = frames + e;
		for (int f0 = 0; f0 < g.N; f0 += F) {
			uint4 q[F]; bool ok[F];
#pragma unroll
			for (int k = 0; k < F; k++) {
				const int f = f0 + k;
				ok[k] = false;
				if (f < g.N) {
					const int ny = g.row0 + r - (g.shifty ? g.shifty[f] : 0);
					ok[k] = ny >= 0 && ny < g.frame_h;
					if (ok[k]) q[k] = __ldg((const uint4 *) (src + (size_t) f * fstride));
				}
			}
#pragma unroll
			for (int k = 0; k < F; k++) if (ok[k]) {
				T v[V];
				memcpy(v, &q[k], 16);
#pragma unroll
				for (int j = 0; j < V; j++) {
					if (OP == 0) acc[j] += (Acc) v[j];
					else if (OP == 1) { if (v[j] > (T) acc[j]) acc[j] = (Acc) v[j]; }
					else { if (v[j] < (T) acc[j]) acc[j] = (Acc) v[j]; }
				}
			}
		}
		uint4 *o = (uint4 *) (out + e);		// V * sizeof(Acc) bytes, 16-byte aligned
#pragma unroll
		for (int j = 0; j < (int) (V * sizeof(Acc) / 16); j++) { uint4 w; memcpy(&w, (const char *) acc + 16 * j, 16); o[j] = w; }
	}
	if (OP == 0) {
		__shared__ double sh[8];
		double m = 0.0;
		if (idx < nv) {
#pragma unroll
			for (int j = 0; j < V; j++) m = fmax(m, (double) acc[j]);
		}
		for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
		if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
		__syncthreads();
		if (threadIdx.x == 0) { for (int k = 1; k < 8; k++) m = fmax(m, sh[k]); blockmax[blockIdx.x] = m; }
	}
}

// sum.c:307-343: float = (float)(sum * (1 / max)); WORD = round(sum) or round(sum * (65535 / max))
template <typename Acc, typename O>
__global__ void sum_convert_kernel(const Acc *acc, size_t n, double ratio, int scaled, O *out) {
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
		const double s = (double) acc[i];
		if (sizeof(O) == 4) out[i] = (O) (float) (s * ratio);
		else out[i] = (O) (scaled ? round_to_word_d(s * ratio) : round_to_word_d(s));
	}
}

}  // namespace sgpu

using namespace sgpu;

extern "C" int sgpu_simple_block(sgpu_ctx *ctx, sgpu_block *blk, int op, const int *shiftx, const int *shifty,
		int row0, int frame_height, void *d_acc, double *max_out) {
	if (!ctx || !blk || !d_acc) return SGPU_GENERIC_ERROR;
	if (op < SGPU_SIMPLE_SUM || op > SGPU_SIMPLE_MIN) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "bad simple stacking operation %d", op);
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const sgpu_block_desc &d = blk->d;
	const size_t total = (size_t) d.nb_layers * d.height * d.width;
	if (total == 0 || d.nb_frames <= 0) return SGPU_OK;
	// 16 bytes per thread and frame when nothing moves along x (simple_stack_vec_kernel)
	bool xshift = false;
	if (shiftx) for (int f = 0; f < d.nb_frames; f++) xshift |= shiftx[f] != 0;
	const int V = d.dtype == SGPU_F32 ? 4 : 8;
	const bool vec = !xshift && d.width % V == 0 && ((uintptr_t) blk->dev & 15) == 0 && ((uintptr_t) d_acc & 15) == 0 &&
			!getenv("SGPU_SIMPLE_SCALAR");
	const unsigned blocks = (unsigned) (((vec ? total / V : total) + 255) / 256);
	SimpleGeom g{ d.nb_frames, d.nb_layers, d.height, d.width, row0, frame_height, nullptr, nullptr };
	// shifts: through the pinned parameter arena (wait for earlier users of the arena first)
	const size_t sbytes = (size_t) d.nb_frames * sizeof(int);
	if ((shiftx || shifty) && 2 * sbytes > ctx->d_params_cap) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "too many frames for the parameter arena");
	if (shiftx || shifty) {
		SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		if (shiftx) { memcpy(ctx->h_params, shiftx, sbytes); g.shiftx = (const int *) ctx->d_params; }
		if (shifty) { memcpy((char *) ctx->h_params + sbytes, shifty, sbytes); g.shifty = (const int *) ((char *) ctx->d_params + sbytes); }
		SGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_params, ctx->h_params, 2 * sbytes, cudaMemcpyHostToDevice, ctx->stream));
	}
	double *d_max = nullptr;
	if (op == SGPU_SIMPLE_SUM) {
		int rv = sgpu_ensure_scratch(ctx, (size_t) blocks * sizeof(double)); if (rv) return rv;
		d_max = (double *) ctx->d_scratch;
	}
	const bool f32 = d.dtype == SGPU_F32;
#define LAUNCH(T, Acc, OP) do { if (vec) simple_stack_vec_kernel<T, Acc, OP><<<blocks, 256, 0, ctx->stream>>>((const T *) blk->dev, g, (Acc *) d_acc, d_max); \
	else simple_stack_kernel<T, Acc, OP><<<blocks, 256, 0, ctx->stream>>>((const T *) blk->dev, g, (Acc *) d_acc, d_max); } while (0)
	if (op == SGPU_SIMPLE_SUM) { if (f32) LAUNCH(float, double, 0); else LAUNCH(u16, unsigned long long, 0); }
	else if (op == SGPU_SIMPLE_MAX) { if (f32) LAUNCH(float, float, 1); else LAUNCH(u16, u16, 1); }
	else { if (f32) LAUNCH(float, float, 2); else LAUNCH(u16, u16, 2); }
#undef LAUNCH
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	if (op == SGPU_SIMPLE_SUM && max_out) {
		std::vector<double> part(blocks);
		SGPU_CUDA(ctx, cudaMemcpyAsync(part.data(), d_max, (size_t) blocks * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
		SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		double m = 0.0;
		for (double v : part) m = std::max(m, v);
		*max_out = m;
	}
	return SGPU_OK;
}

extern "C" int sgpu_sum_convert(sgpu_ctx *ctx, const void *d_acc, int acc_is_float, size_t n, double max,
		int output_32bits, void *d_out) {
	if (!ctx || !d_acc || !d_out) return SGPU_GENERIC_ERROR;
	if (!(max > 0.0)) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "sum stacking: the sum image is empty");	// sum.c:266,275
	if (acc_is_float && !output_32bits) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "sum stacking: float sequences stack to 32 bits");
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const int blocks = ctx->sm_count * 8;
	if (output_32bits) {
		const double ratio = 1.0 / max;
		if (acc_is_float) sum_convert_kernel<double, float><<<blocks, 256, 0, ctx->stream>>>((const double *) d_acc, n, ratio, 1, (float *) d_out);
		else sum_convert_kernel<unsigned long long, float><<<blocks, 256, 0, ctx->stream>>>((const unsigned long long *) d_acc, n, ratio, 1, (float *) d_out);
	} else {
		const int scaled = max > 65535.0;
		const double ratio = scaled ? 65535.0 / max : 1.0;
		sum_convert_kernel<unsigned long long, u16><<<blocks, 256, 0, ctx->stream>>>((const unsigned long long *) d_acc, n, ratio, scaled, (u16 *) d_out);
	}
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}
