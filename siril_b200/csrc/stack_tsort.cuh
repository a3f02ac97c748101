// stack_tsort.cuh — kernel of the THREAD-SORT engine: stacks of up to 128 frames, every rejection
// type the sorted engine covers, both sample types.
//
// One THREAD owns one pixel; a CTA of 128 threads owns 128 consecutive pixels of one layer.
//  load    frame by frame, lane = pixel: every warp instruction reads one 128-byte (64-byte for
//          16-bit samples) row segment of the frame-major block, fully coalesced, each input
//          sample crosses HBM exactly once; all NS loads of a thread are independent, so a warp
//          keeps NS requests in flight.  Samples are normalized in double exactly as
//          median_and_mean.c:1644-1686 does and kept in REGISTERS; the running sums
//          sum(x-K), sum((x-K)^2) (double) are taken on the way.
//  sort    Batcher's odd-even merge sort as straight-line code on the thread's NS registers
//          (gen_sortnet.py; two FMNMX per comparator, no shuffles, no shared memory).  Null
//          samples (mean mode, rejection_float.c:128-135) and padding are +inf and end up last.
//  reject  the sorted stack is written to the thread's shared-memory column [rank][128]
//          (conflict free) and the per-pixel rejection / combine of the sorted engine
//          (sorted.cuh) runs on it, one thread per pixel, all threads busy.  Pixels it cannot
//          decide identically to the reference go to the exact engine's list.
#pragma once
#include "kparams.cuh"
#include "sorted.cuh"

namespace sgpu {

constexpr int TS_BD = 128;	// threads (= pixels) per CTA
#ifndef SGPU_INF
#define SGPU_INF __int_as_float(0x7f800000)
#endif

template <int NS> struct SortNet { static __device__ __forceinline__ void run(float (&v)[NS]); };

#define CE(a, b) { const float lo__ = fminf(v[a], v[b]); v[b] = fmaxf(v[a], v[b]); v[a] = lo__; }
#define SGPU_SORTNET(NS_) template <> __device__ __forceinline__ void SortNet<NS_>::run(float (&v)[NS_])

struct ColStack {	// one pixel's sorted stack: column of the [rank][TS_BD] tile
	float *p;
	__device__ __forceinline__ float operator()(int i) const { return p[i * TS_BD]; }
	__device__ __forceinline__ void set(int i, float x) const { p[i * TS_BD] = x; }
};

template <typename T, int MODE>
__device__ __forceinline__ float ts_normalize(T v, double c0, double c1) {
	if (MODE == 0) return (float) v;
	if (MODE == 1) {
		if (v == (T) 0) return 0.f;
		const double tmp = (double) v * c0;
		return sizeof(T) == 4 ? (float) (tmp - c1) : (float) round_to_word(tmp - c1);
	}
	const double tmp = (double) v * c0;
	return sizeof(T) == 4 ? (float) (tmp * c1) : (float) round_to_word(tmp * c1);
}

// load + normalize the thread's stack into registers
template <typename T, int NS, int MODE>
__device__ __forceinline__ void ts_load(const StackParams &p, float (&v)[NS], int layer, size_t pix, size_t npix, bool in) {
	const int N = p.N;
	const size_t fstride = (size_t) p.C * npix;
	const T *src = (const T *) p.frames + (size_t) layer * npix + (in ? pix : 0);
	const double *c0 = p.scale + layer * N, *c1 = (MODE == 1 ? p.offset : p.mul) + layer * N;
	T raw[NS];
	if (p.shiftx) {
		const int x = in ? (int) (pix % p.W) : 0;
#pragma unroll
		for (int f = 0; f < NS; f++) {
			raw[f] = (T) 0;
			if (f < N && in) { const int sx = p.shiftx[f], xs = x - sx; if (xs >= 0 && xs < p.W) raw[f] = __ldg(src + (size_t) f * fstride - sx); }
		}
	} else {
#pragma unroll
		for (int f = 0; f < NS; f++) { raw[f] = (T) 0; if (f < N && in) raw[f] = __ldg(src + (size_t) f * fstride); }
	}
#pragma unroll
	for (int f = 0; f < NS; f++) v[f] = (f < N) ? ts_normalize<T, MODE>(raw[f], MODE ? c0[f] : 0.0, MODE ? c1[f] : 0.0) : SGPU_INF;
}

template <int NS, int MINB>
__global__ void __launch_bounds__(TS_BD, MINB) stack_tsort_kernel(const StackParams p) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	float *tile = (float *) smem_raw;	// [N][TS_BD]
	const int N = p.N, layer = blockIdx.y;
	const size_t npix = (size_t) p.H * p.W, pix = (size_t) blockIdx.x * TS_BD + threadIdx.x;
	const bool in = pix < npix;

	float v[NS];
	{
		const int mode = p.normalize == SGPU_NO_NORM ? 0 : ((p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING) ? 1 : 2);
		if (p.is_u16) {
			if (mode == 0) ts_load<u16, NS, 0>(p, v, layer, pix, npix, in);
			else if (mode == 1) ts_load<u16, NS, 1>(p, v, layer, pix, npix, in);
			else ts_load<u16, NS, 2>(p, v, layer, pix, npix, in);
		} else {
			if (mode == 0) ts_load<float, NS, 0>(p, v, layer, pix, npix, in);
			else if (mode == 1) ts_load<float, NS, 1>(p, v, layer, pix, npix, in);
			else ts_load<float, NS, 2>(p, v, layer, pix, npix, in);
		}
	}
	// nulls -> +inf (mean mode), non-finite samples flagged, sums about the first sample
	int n = 0, weird = 0;
	Sums S; S.K = 0.0; S.S1 = 0.0; S.S2 = 0.0;
	{
		const float k0 = v[0];
		const double K = (fabsf(k0) < SGPU_INF) ? (double) k0 : 0.0;
		double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
#pragma unroll
		for (int f = 0; f < NS; f++) {
			float x = v[f];
			bool ok = f < N;
			if (ok && !(fabsf(x) < SGPU_INF)) { weird = 1; ok = false; }
			if (ok && p.is_mean && x == 0.f) ok = false;
			const double y = ok ? (double) x - K : 0.0;
			if (f & 1) { b1 += y; b2 = fma(y, y, b2); } else { a1 += y; a2 = fma(y, y, a2); }
			n += ok;
			v[f] = ok ? x : SGPU_INF;
		}
		S.K = K; S.S1 = a1 + b1; S.S2 = a2 + b2;
	}
	SortNet<NS>::run(v);
	ColStack st{ tile + threadIdx.x };
#pragma unroll
	for (int i = 0; i < NS; i++) if (i < N) st.set(i, v[i]);

	int rej[2] = { 0, 0 };
	if (in) {
		int status = weird ? FAST_FALLBACK : FAST_DONE;
		double result = 0.0;
		if (status == FAST_DONE) {
			if (p.is_mean) {
				const int x = (int) (pix % p.W);
				const double *w = p.weights ? p.weights + (size_t) layer * N : nullptr;
				const Sums *S0 = (p.need_sums && n > 0) ? &S : nullptr;
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
		if ((threadIdx.x & 31) == 0 && (lo | hi)) {
			atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
		}
	}
}

}  // namespace sgpu

template <int NS, int MINB>
int sgpu_launch_tsort_variant(sgpu_ctx *ctx, const sgpu::StackParams &p, dim3 grid, size_t shm);

#define SGPU_TSORT_VARIANT(NS, MINB) \
	template <> int sgpu_launch_tsort_variant<NS, MINB>(sgpu_ctx *ctx, const sgpu::StackParams &p, dim3 grid, size_t shm) { \
		auto kern = sgpu::stack_tsort_kernel<NS, MINB>; \
		if (shm > 48 * 1024) SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm)); \
		kern<<<grid, sgpu::TS_BD, shm, ctx->stream>>>(p); \
		ctx->launches++; \
		SGPU_CUDA(ctx, cudaGetLastError()); \
		return SGPU_OK; \
	}
