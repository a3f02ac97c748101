// misc.cu — output renormalization (norm_to_0_1_range, median_and_mean.c:557-582) and the
// device-side synthetic frame generator used by bench.py / tests.
#include <cfloat>
#include <cstdio>
#include <mutex>
#include "common.cuh"

namespace sgpu {

// min / max over non-zero samples; one partial per CTA, reduced by the host (<= 1184 pairs)
__global__ void minmax_nonzero_kernel(const float *img, size_t n, size_t first, float2 *partial) {
	float mn = FLT_MAX, mx = -FLT_MAX;
	for (size_t i = first + (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
		const float t = img[i];
		if (t == 0.f) continue;
		if (t < mn) mn = t;
		if (t > mx) mx = t;
	}
	for (int o = 16; o; o >>= 1) {
		mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
		mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
	}
	__shared__ float smn[32], smx[32];
	if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int w = 1; w < (int) (blockDim.x >> 5); w++) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
		partial[blockIdx.x] = make_float2(mn, mx);
	}
}

// ---- the same reduction with NO allocation: its scratch is this file-scope array (64 KiB held by every
// GPU that loads the library, until exit).  Slot 0: the result; slots 1..: one partial per chunk of
// SGPU_STATIC_CHUNK pixels.  Every slot a call reads was written by that call: nothing has to start
// at zero.  All calls share the array: two calls overlapping on the GPU race.
#define SGPU_STATIC_CHUNK 65536
__device__ __align__(256) float2 g_static_scratch[SGPU_STATIC_SCRATCH_BYTES / sizeof(float2)];

__global__ void __launch_bounds__(256) minmax_nonzero_chunk_kernel(const float *img, size_t n, size_t first, float2 *scratch) {
	float mn = FLT_MAX, mx = -FLT_MAX;
	const size_t lo = first + (size_t) blockIdx.x * SGPU_STATIC_CHUNK;
	const size_t hi = lo + SGPU_STATIC_CHUNK < n ? lo + SGPU_STATIC_CHUNK : n;
	for (size_t i = lo + threadIdx.x; i < hi; i += 256) {
		const float t = img[i];
		if (t == 0.f) continue;
		if (t < mn) mn = t;
		if (t > mx) mx = t;
	}
	for (int o = 16; o; o >>= 1) {
		mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
		mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
	}
	__shared__ float smn[8], smx[8];
	if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int w = 1; w < 8; w++) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
		scratch[1 + blockIdx.x] = make_float2(mn, mx);
	}
}

__global__ void __launch_bounds__(256) minmax_final_kernel(float2 *scratch, int nparts) {
	float mn = FLT_MAX, mx = -FLT_MAX;
	for (int i = threadIdx.x; i < nparts; i += 256) { const float2 p = scratch[1 + i]; mn = fminf(mn, p.x); mx = fmaxf(mx, p.y); }
	for (int o = 16; o; o >>= 1) {
		mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
		mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
	}
	__shared__ float smn[8], smx[8];
	if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int w = 1; w < 8; w++) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
		scratch[0] = make_float2(mn, mx);
	}
}

// device address of g_static_scratch, per device: in host code the array's NAME is only a stand-in
static float2 *static_scratch_of(sgpu_ctx *ctx) {
	static std::mutex mu;
	static float2 *addr[SGPU_STATIC_MAX_DEVICES];
	if (ctx->device < 0 || ctx->device >= SGPU_STATIC_MAX_DEVICES) {
		sgpu_fail(ctx, SGPU_GENERIC_ERROR, "device %d is beyond the %d devices the static scratch is tracked for", ctx->device, SGPU_STATIC_MAX_DEVICES);
		return nullptr;
	}
	std::lock_guard<std::mutex> lk(mu);
	if (!addr[ctx->device]) {	// set-up, once per device, with that device current
		void *p = nullptr;
		cudaError_t e = cudaSetDevice(ctx->device);
		if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_static_scratch);
		if (e != cudaSuccess) {
			sgpu_cuda_fail(ctx, e, "cudaGetSymbolAddress(g_static_scratch)");
			return nullptr;
		}
		addr[ctx->device] = (float2 *) p;
	}
	return addr[ctx->device];
}

__global__ void rescale_kernel(float *img, size_t n, float mini, float maxi) {
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
		const float v = img[i];
		img[i] = (v == 0.f) ? 0.f : (v - mini) / (maxi - mini);
	}
}

// counter-based generator: 64-bit mix of (seed, frame, layer, row, col, stream)
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
	z += 0x9e3779b97f4a7c15ull;
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	return z ^ (z >> 31);
}
__device__ __forceinline__ float u01(uint64_t h) { return ((h >> 40) + 0.5f) * (1.0f / 16777216.0f); }

template <typename T>
__global__ void synth_kernel(T *dst, int N, int C, int H, int W, uint64_t seed, int row0, int fullH) {
	const size_t plane = (size_t) H * W;
	const size_t total = plane * C * N;
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t) gridDim.x * blockDim.x) {
		const int x = (int) (i % W);
		const int r = (int) ((i / W) % H) + row0;
		const int l = (int) ((i / plane) % C);
		const int f = (int) (i / (plane * C));
		const uint64_t kf = mix64(seed * 0x100000001b3ull + (uint64_t) f * 7919u + l);
		const float g = 2.f * u01(mix64(kf ^ 1)) - 1.f, h = 2.f * u01(mix64(kf ^ 2)) - 1.f;
		const float sky = 0.10f * (1.f + 0.02f * g), gain = 1.f + 0.05f * h;
		const float fx = (float) x / W - 0.5f, fy = (float) r / fullH - 0.5f;
		const float S = 0.3f * __expf(-(fx * fx + fy * fy) * 8.f) * (0.6f + 0.4f * __sinf(12.f * fx) * __cosf(9.f * fy)) ;
		const uint64_t kp = mix64(kf + ((uint64_t) r << 32) + (uint64_t) x);
		const float u1 = u01(mix64(kp ^ 3)), u2 = u01(mix64(kp ^ 4));
		const float n = sqrtf(-2.f * __logf(u1)) * __cosf(6.2831853f * u2);
		float v = gain * (sky + fabsf(S)) + 0.01f * n;
		const float uo = u01(mix64(kp ^ 5));
		if (uo < 1e-3f) v += 0.2f + 0.4f * u01(mix64(kp ^ 6));
		else if (uo < 1.2e-3f) v -= 0.05f;
		v = fminf(fmaxf(v, 1e-4f), 1.f);
		if (sizeof(T) == 4) dst[i] = (T) v;
		else { float q = rintf(v * 65535.f); dst[i] = (T) (q < 1.f ? 1.f : q); }
	}
}

// feather-mask normalization (median_and_mean.c:513-521): distances above feather_dist become 1, the others go
// through the smoothstep table of blending.c:34-50 (1001 points, looked up by truncation) — the table entry is
// recomputed here with the same float operations instead of being stored; zeros stay zero
__global__ void mask_ramp_kernel(float *mask, size_t n, float distf, float invdistf) {
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
		const float v = mask[i];
		if (v == 0.f) continue;	// `if (data->mask[frame][i])`: NaN is true and takes the ramp like any value
		float o = 1.f;
		if (!(v > distf)) {
			const int index = (int) ((v * invdistf) * 1000);	// get_ramped_value: (int)(val * RAMP_PACE)
			const float r = (float) index * (1.f / (float) 1000);
			o = r * r * r * (6.f * r * r - 15.f * r + 10.f);
		}
		mask[i] = o;
	}
}

}  // namespace sgpu

using namespace sgpu;

int sgpu_launch_mask_ramp(sgpu_ctx *ctx, float *d_mask, size_t n, int feather_dist) {
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const float distf = (float) feather_dist, invdistf = 1.f / distf;
	mask_ramp_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(d_mask, n, distf, invdistf);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}

extern "C" {

int sgpu_minmax_nonzero(sgpu_ctx *ctx, const float *d_img, size_t n, int skip_first, float minmax[2]) {
	if (!ctx || !d_img || !minmax) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const int blocks = ctx->sm_count * 8;
	int rv = sgpu_ensure_scratch(ctx, blocks * sizeof(float2)); if (rv) return rv;
	minmax_nonzero_kernel<<<blocks, 256, 0, ctx->stream>>>(d_img, n, skip_first ? 1 : 0, (float2 *) ctx->d_scratch);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	std::vector<float2> part(blocks);
	SGPU_CUDA(ctx, cudaMemcpyAsync(part.data(), ctx->d_scratch, blocks * sizeof(float2), cudaMemcpyDeviceToHost, ctx->stream));
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	float mn = FLT_MAX, mx = -FLT_MAX;
	for (auto &p : part) { if (p.x < mn) mn = p.x; if (p.y > mx) mx = p.y; }
	minmax[0] = mn; minmax[1] = mx;
	return SGPU_OK;
}

int sgpu_minmax_nonzero_static(sgpu_ctx *ctx, const float *d_img, size_t n, int skip_first, float minmax[2]) {
	if (!ctx || !d_img || !minmax) return SGPU_GENERIC_ERROR;
	if (n > (size_t) SGPU_STATIC_MAX_PIXELS) {
		return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "sgpu_minmax_nonzero_static: %zu pixels are beyond the static scratch's limit of %zu "
				"(%d bytes, one 8-byte partial per %d pixels): use sgpu_minmax_nonzero", n, (size_t) SGPU_STATIC_MAX_PIXELS,
				SGPU_STATIC_SCRATCH_BYTES, SGPU_STATIC_CHUNK);
	}
	float2 *scratch = static_scratch_of(ctx);
	if (!scratch) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const size_t first = skip_first ? 1 : 0;
	const int nparts = n > first ? (int) ((n - first + SGPU_STATIC_CHUNK - 1) / SGPU_STATIC_CHUNK) : 0;
	if (nparts) minmax_nonzero_chunk_kernel<<<nparts, 256, 0, ctx->stream>>>(d_img, n, first, scratch);
	minmax_final_kernel<<<1, 256, 0, ctx->stream>>>(scratch, nparts);
	ctx->launches += nparts ? 2 : 1;
	SGPU_CUDA(ctx, cudaGetLastError());
	float2 r;
	SGPU_CUDA(ctx, cudaMemcpyAsync(&r, scratch, sizeof r, cudaMemcpyDeviceToHost, ctx->stream));
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	minmax[0] = r.x; minmax[1] = r.y;
	return SGPU_OK;
}

int sgpu_rescale_0_1(sgpu_ctx *ctx, float *d_img, size_t n, float mini, float maxi) {
	if (!ctx || !d_img) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	rescale_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(d_img, n, mini, maxi);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}

int sgpu_synth_fill(sgpu_ctx *ctx, sgpu_block *blk, uint64_t seed, int row0, int full_height) {
	if (!ctx || !blk) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const sgpu_block_desc &d = blk->d;
	const int blocks = ctx->sm_count * 16;
	if (d.dtype == SGPU_F32)
		synth_kernel<float><<<blocks, 256, 0, ctx->stream>>>((float *) blk->dev, d.nb_frames, d.nb_layers, d.height, d.width, seed, row0, full_height);
	else
		synth_kernel<uint16_t><<<blocks, 256, 0, ctx->stream>>>((uint16_t *) blk->dev, d.nb_frames, d.nb_layers, d.height, d.width, seed, row0, full_height);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}

}  // extern "C"
