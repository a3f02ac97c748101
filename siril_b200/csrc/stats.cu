// stats.cu — per-image statistics that feed normalization and noise weighting (K3):
// statistics_internal_float (src/algos/statistics_float.c:281-481) / statistics_internal_ushort
// (src/algos/statistics.c:233-443) with option STATS_NORM or STATS_LITENORM, i.e.
//   min/max, ngoodpix/mean/sigma (FnMeanSigma, src/algos/quantize.c:207-317),
//   bgnoise (FnNoise1, quantize.c:1202-1488), histogram median (src/rt/rt_algo.cc:38-171 for
//   float, src/algos/sorting.c:577-642 for WORD), MAD, and IKSSlite location/scale
//   (statistics_float.c:199-229, bwmv :103-127).
// Every pass is a memory-bound sweep of one plane; integer histograms make the medians
// independent of the summation order, so they are bit-identical to the CPU; the double sums
// (mean, sigma, bwmv, row noise) are reduced in a fixed tree order (deterministic run to run)
// and agree with the sequential CPU sums to ~1e-15 relative.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include "common.cuh"

namespace sgpu {

typedef uint16_t u16;
static constexpr int NB = 592;		// CTAs per sweep: 4 per SM on 148 SMs
static constexpr int NT = 256;

__device__ __forceinline__ float cvf(float raw, float) { return raw; }
__device__ __forceinline__ float cvf(u16 raw, float inv) { return (float) raw * inv; }

// Grid-stride sweep of a plane with 16-byte loads, four in flight per thread (a scalar loop keeps
// < 1 MB in flight on the whole GPU and runs at ~1 TB/s): f(raw) is called once per sample; the
// unaligned head and the tail go through scalar loads.
template <typename T, typename F>
__device__ __forceinline__ void sweep(const T *d, size_t n, F f) {
	constexpr int V = 16 / sizeof(T);
	union Vec { uint4 u; T e[V]; };
	const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t) gridDim.x * blockDim.x;
	size_t head = ((16 - ((uintptr_t) d & 15)) & 15) / sizeof(T);
	if (head > n) head = n;
	const size_t nv = (n - head) / V;
	const uint4 *dv = (const uint4 *) (d + head);
	size_t i = tid;
	for (; i + 3 * nth < nv; i += 4 * nth) {
		Vec a, b, c, e;
		a.u = dv[i]; b.u = dv[i + nth]; c.u = dv[i + 2 * nth]; e.u = dv[i + 3 * nth];
		#pragma unroll
		for (int k = 0; k < V; k++) f(a.e[k]);
		#pragma unroll
		for (int k = 0; k < V; k++) f(b.e[k]);
		#pragma unroll
		for (int k = 0; k < V; k++) f(c.e[k]);
		#pragma unroll
		for (int k = 0; k < V; k++) f(e.e[k]);
	}
	for (; i < nv; i += nth) {
		Vec a; a.u = dv[i];
		#pragma unroll
		for (int k = 0; k < V; k++) f(a.e[k]);
	}
	for (size_t k = tid; k < head; k += nth) f(d[k]);
	for (size_t k = head + nv * V + tid; k < n; k += nth) f(d[k]);
}

// block-wide reduction in a fixed order (xor tree inside each warp, then warps 0..n-1 in turn);
// every thread returns the result
template <typename V, typename Op>
__device__ __forceinline__ V block_reduce(V v, Op op, V *sh) {
	for (int o = 16; o; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
	if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
	__syncthreads();
	V acc = sh[0];
	for (int k = 1; k < (int) (blockDim.x >> 5); k++) acc = op(acc, sh[k]);
	__syncthreads();
	return acc;
}

struct BasicPartial {
	double sum, sum2;
	unsigned long long ngood, isum, isum2;
	float mn, mx, nzmn, nzmx;
};

// pass A: min/max of all (non-NaN) samples, count/sum/sum2 and min/max of the non-null ones
template <typename T>
__global__ void basic_kernel(const T *d, size_t n, BasicPartial *out) {
	__shared__ double shd[32]; __shared__ float shf[32]; __shared__ unsigned long long shu[32];
	double sum = 0, sum2 = 0; unsigned long long ng = 0, is = 0, is2 = 0;
	float mn = FLT_MAX, mx = -FLT_MAX, nzmn = FLT_MAX, nzmx = -FLT_MAX;
	sweep(d, n, [&](T raw) {
		const float v = (float) raw;
		if (sizeof(T) == 4 && v != v) return;
		mn = fminf(mn, v); mx = fmaxf(mx, v);
		if (raw != (T) 0) {
			ng++; nzmn = fminf(nzmn, v); nzmx = fmaxf(nzmx, v);
			if (sizeof(T) == 4) { const double x = (double) v; sum += x; sum2 += x * x; }
			else { const unsigned long long x = (unsigned long long) raw; is += x; is2 += x * x; }
		}
	});
	auto addd = [](double a, double b) { return a + b; };
	auto addu = [](unsigned long long a, unsigned long long b) { return a + b; };
	auto fmn = [](float a, float b) { return fminf(a, b); };
	auto fmx = [](float a, float b) { return fmaxf(a, b); };
	sum = block_reduce(sum, addd, shd); sum2 = block_reduce(sum2, addd, shd);
	ng = block_reduce(ng, addu, shu); is = block_reduce(is, addu, shu); is2 = block_reduce(is2, addu, shu);
	mn = block_reduce(mn, fmn, shf); mx = block_reduce(mx, fmx, shf);
	nzmn = block_reduce(nzmn, fmn, shf); nzmx = block_reduce(nzmx, fmx, shf);
	if (threadIdx.x == 0) out[blockIdx.x] = BasicPartial{ sum, sum2, ng, is, is2, mn, mx, nzmn, nzmx };
}

// selection predicate + transform shared by the min/max, histogram and bwmv sweeps:
// sample is used when non-null (and inside [lo, hi] when ranged); value = x or |x - c|
struct Sel { int ranged; float lo, hi; int absdev; float c; float inv; };

template <typename T>
__device__ __forceinline__ bool sel_value(T raw, const Sel &s, float &val, float &x) {
	if (raw == (T) 0) return false;
	x = cvf(raw, s.inv);
	if (x != x) return false;
	if (s.ranged && !(x >= s.lo && x <= s.hi)) return false;
	val = s.absdev ? fabsf(x - s.c) : x;
	return true;
}

struct MMPartial { float mn, mx; unsigned long long cnt; };

template <typename T>
__global__ void minmax_sel_kernel(const T *d, size_t n, Sel s, MMPartial *out) {
	__shared__ float shf[32]; __shared__ unsigned long long shu[32];
	float mn = FLT_MAX, mx = -FLT_MAX; unsigned long long cnt = 0;
	sweep(d, n, [&](T raw) {
		float v, x;
		if (!sel_value(raw, s, v, x)) return;
		mn = fminf(mn, v); mx = fmaxf(mx, v); cnt++;
	});
	mn = block_reduce(mn, [](float a, float b) { return fminf(a, b); }, shf);
	mx = block_reduce(mx, [](float a, float b) { return fmaxf(a, b); }, shf);
	cnt = block_reduce(cnt, [](unsigned long long a, unsigned long long b) { return a + b; }, shu);
	if (threadIdx.x == 0) out[blockIdx.x] = MMPartial{ mn, mx, cnt };
}

// histogram of (uint16_t)(scale * (val - mn)) — findMinMaxPercentile's binning (rt_algo.cc:99)
template <typename T>
__global__ void hist_float_kernel(const T *d, size_t n, Sel s, float mn, float scale, unsigned *hist) {
	sweep(d, n, [&](T raw) {
		float v, x;
		if (!sel_value(raw, s, v, x)) return;
		const unsigned bin = (unsigned) (u16) (scale * (v - mn));
		atomicAdd(&hist[bin], 1u);
	});
}

// The same binning with the counters of bins [w0, w0 + HWIN) in shared memory (no global
// atomics in the sweep): samples below the window are only counted, samples above it ignored.
// The host places the window where the median is expected and falls back to the full
// histogram when the median is not inside.
static constexpr unsigned HWIN = 32768;	// bins per window: 128 KB of shared memory, one CTA per SM
static constexpr int HNB = 148, HNT = 1024;

__device__ __forceinline__ void win_count(unsigned bin, unsigned w0, unsigned *sh, unsigned &below) {
	const unsigned b = bin - w0;
	if (bin < w0) below++;
	else if (b < HWIN) atomicAdd(&sh[b], 1u);
}

__device__ __forceinline__ void win_flush(const unsigned *sh, unsigned below, unsigned *hist, unsigned long long *d_below) {
	__syncthreads();
	for (unsigned i = threadIdx.x; i < HWIN; i += blockDim.x) { const unsigned c = sh[i]; if (c) atomicAdd(&hist[i], c); }
	below = __reduce_add_sync(0xffffffffu, below);
	if ((threadIdx.x & 31) == 0 && below) atomicAdd(d_below, (unsigned long long) below);
}

template <typename T>
__global__ void __launch_bounds__(HNT, 1) hist_float_win_kernel(const T *d, size_t n, Sel s, float mn, float scale, unsigned w0,
		unsigned *hist, unsigned long long *d_below) {
	extern __shared__ unsigned sh[];
	for (unsigned i = threadIdx.x; i < HWIN; i += blockDim.x) sh[i] = 0;
	__syncthreads();
	unsigned below = 0;
	sweep(d, n, [&](T raw) {
		float v, x;
		if (!sel_value(raw, s, v, x)) return;
		win_count((unsigned) (u16) (scale * (v - mn)), w0, sh, below);
	});
	win_flush(sh, below, hist, d_below);
}

// exact histogram of WORD samples, or of |x - med| (siril_stats_ushort_mad, statistics.c:133-154)
__device__ __forceinline__ bool u16_bin(u16 raw, int absdev, int med, unsigned &bin) {
	if (raw == 0) return false;
	int v = raw;
	if (absdev) { v -= med; v = v < 0 ? -v : v; v &= 0xffff; }
	bin = (unsigned) v;
	return true;
}

__global__ void hist_u16_kernel(const u16 *d, size_t n, int absdev, int med, unsigned *hist) {
	sweep(d, n, [&](u16 raw) {
		unsigned bin;
		if (u16_bin(raw, absdev, med, bin)) atomicAdd(&hist[bin], 1u);
	});
}

__global__ void __launch_bounds__(HNT, 1) hist_u16_win_kernel(const u16 *d, size_t n, int absdev, int med, unsigned w0,
		unsigned *hist, unsigned long long *d_below) {
	extern __shared__ unsigned sh[];
	for (unsigned i = threadIdx.x; i < HWIN; i += blockDim.x) sh[i] = 0;
	__syncthreads();
	unsigned below = 0;
	sweep(d, n, [&](u16 raw) {
		unsigned bin;
		if (u16_bin(raw, absdev, med, bin)) win_count(bin, w0, sh, below);
	});
	win_flush(sh, below, hist, d_below);
}

struct BwPartial { double up, down; };
// siril_stats_float_bwmv (statistics_float.c:103-127)
template <typename T>
__global__ void bwmv_kernel(const T *d, size_t n, Sel s, float median, float factor, BwPartial *out) {
	__shared__ double shd[32];
	double up = 0, down = 0;
	sweep(d, n, [&](T raw) {
		float v, x;
		if (!sel_value(raw, s, v, x)) return;
		const float im = x - median;
		const float yi = im * factor;
		const float yi2 = fabsf(yi) < 1.f ? yi * yi : 1.f;
		const float q = 1 - yi2;
		const float t = im * (q * q);
		up += (double) (t * t);
		down += (double) (q * (1 - 5 * yi2));
	});
	auto addd = [](double a, double b) { return a + b; };
	up = block_reduce(up, addd, shd); down = block_reduce(down, addd, shd);
	if (threadIdx.x == 0) out[blockIdx.x] = BwPartial{ up, down };
}

// FnNoise1_float / FnNoise1_ushort (quantize.c:1343-1488 / :1202-1341), one CTA per row:
// first differences of consecutive valid pixels, then <= 3 passes of 5-sigma clipping of the
// differences around their mean; row value = sigma, or -1 when the row yields < 2 differences.
// Each warp compacts a contiguous segment of the row with coalesced loads (ballot + popc keep the
// pixel order); clipped differences are overwritten with a marker.
struct NoiseAcc { double s, s2; unsigned long long k; };
static constexpr unsigned NOISE_CLIPPED = 0xffc0c11bu;	// NaN payload no arithmetic produces: marks a clipped difference

template <typename T>
__global__ void __launch_bounds__(NT) noise_rows_kernel(const T *d, long nx, long ny, double *rowsd) {
	extern __shared__ unsigned char sm[];
	float *vals = (float *) sm;				// nx floats: valid pixels in order
	float *dif = vals + nx;					// nx floats: the staged row, then the first differences
	__shared__ int wc[NT / 32];
	__shared__ NoiseAcc acc[NT / 32];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const long seg = (((nx + NT / 32 - 1) / (NT / 32)) + 31) & ~31L;
	const long s0 = warp * seg, s1 = s0 + seg < nx ? s0 + seg : nx;
	for (long row = blockIdx.x; row < ny; row += gridDim.x) {
		const T *rp = d + row * nx;
		// stage the row (independent coalesced loads), then compact from shared memory
		#pragma unroll 8
		for (long i = threadIdx.x; i < nx; i += NT) dif[i] = (float) rp[i];
		__syncthreads();
		int c = 0;
		for (long i0 = s0; i0 < s1; i0 += 32) {
			const long i = i0 + lane;
			bool ok = false;
			if (i < s1) { const float v = dif[i]; ok = v != 0.f && v == v; }
			c += __popc(__ballot_sync(0xffffffffu, ok));
		}
		if (lane == 0) wc[warp] = c;
		__syncthreads();
		int o = 0, M = 0;
		for (int k = 0; k < NT / 32; k++) { if (k < warp) o += wc[k]; M += wc[k]; }
		for (long i0 = s0; i0 < s1; i0 += 32) {
			const long i = i0 + lane;
			bool ok = false; float v = 0.f;
			if (i < s1) { v = dif[i]; ok = v != 0.f && v == v; }
			const unsigned bal = __ballot_sync(0xffffffffu, ok);
			if (ok) vals[o + __popc(bal & ((1u << lane) - 1u))] = v;
			o += __popc(bal);
		}
		__syncthreads();
		const int nd = M - 1;					// number of differences
		if (nd < 2) { if (threadIdx.x == 0) rowsd[row] = -1.0; __syncthreads(); continue; }
		for (int i = threadIdx.x; i < nd; i += NT) dif[i] = vals[i] - vals[i + 1];
		__syncthreads();
		double mean = 0, stdev = 0;
		unsigned long long nv = nd;
		for (int iter = 0; iter <= 3; iter++) {
			double s = 0, s2 = 0; unsigned long long kk = 0;
			for (int i = threadIdx.x; i < nd; i += NT) {
				const float dv = dif[i];
				if (__float_as_uint(dv) == NOISE_CLIPPED) continue;
				if (iter > 0) {
					const double dev = (sizeof(T) == 4) ? fabs((double) (dv - (float) mean)) : fabs((double) dv - mean);
					if (!(dev < 5. * stdev)) { dif[i] = __uint_as_float(NOISE_CLIPPED); continue; }
				}
				const double x = (double) dv; s += x; s2 += x * x; kk++;
			}
			for (int off = 16; off; off >>= 1) {
				s += __shfl_xor_sync(0xffffffffu, s, off); s2 += __shfl_xor_sync(0xffffffffu, s2, off);
				kk += __shfl_xor_sync(0xffffffffu, kk, off);
			}
			if (lane == 0) acc[warp] = NoiseAcc{ s, s2, kk };
			__syncthreads();
			s = acc[0].s; s2 = acc[0].s2; kk = acc[0].k;
			for (int k = 1; k < NT / 32; k++) { s += acc[k].s; s2 += acc[k].s2; kk += acc[k].k; }
			__syncthreads();
			if (iter > 0 && kk == nv) break;		// nothing clipped: keep previous mean/stdev
			nv = kk;
			if (nv > 1) { mean = s / (double) nv; stdev = sqrt((s2 / (double) nv) - (mean * mean)); }
			else if (nv == 1) { mean = s; stdev = 0.0; }
			else { mean = 0.; stdev = 0.; }
			if (iter == 0 && !(stdev > 0.)) break;
		}
		if (threadIdx.x == 0) rowsd[row] = stdev;
		__syncthreads();
	}
}

// overlap normalization (normalization.c:549-565): the two images' samples over their common rectangle, kept where
// BOTH are non-zero (zero elsewhere, which the statistics sweeps skip like the reference's compaction does), WORD data
// converted with (float) v * invnorm
template <typename T>
__global__ void overlap_pair_kernel(const T *pi, const T *pj, long pitch_i, long pitch_j, long w, size_t n, float inv, float *oi, float *oj) {
	for (size_t k = (size_t) blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (size_t) gridDim.x * blockDim.x) {
		const long r = (long) (k / (size_t) w), c = (long) (k - (size_t) r * w);
		const T a = pi[(size_t) r * pitch_i + c], b = pj[(size_t) r * pitch_j + c];
		const bool keep = !(a == (T) 0) && !(b == (T) 0);	// `!x` of a NaN is false: NaN pairs are kept, as in the reference
		oi[k] = keep ? cvf(a, inv) : 0.f;
		oj[k] = keep ? cvf(b, inv) : 0.f;
	}
}

}  // namespace sgpu


using namespace sgpu;

namespace {

struct Sweep {
	sgpu_ctx *ctx; const void *d; int dtype; size_t n;
	unsigned *d_hist; void *d_part; void *h_part;
	// row-sharded frames (one process per GPU holds rows [row0, row0 + ny) of every frame): every
	// reduction over the plane passes through the caller's all-reduce before it is used, so each
	// rank walks identical histograms and sums (integer counts make the medians
	// bit-identical to the single-GPU ones)
	sgpu_reduce_fn reduce = nullptr; void *reduce_user = nullptr;
	int allreduce(void *buf, size_t count, int dtype_, int op) const {
		if (!reduce) return SGPU_OK;
		const int rv = reduce(reduce_user, buf, count, dtype_, op);
		return rv ? sgpu_fail(ctx, SGPU_GENERIC_ERROR, "the all-reduce callback failed (%d)", rv) : SGPU_OK;
	}
};

// rt_algo.cc:112-138 on a finished histogram (minPrct = maxPrct = 0.5, outputs aliased).
// `histo` holds bins [k0, k1) and `count0` samples lie below bin k0; returns false when the
// walk would stop outside [k0, k1) (the caller then builds the full histogram)
bool finish_hist_median_float(const unsigned *histo, size_t k0, size_t k1, size_t count0, size_t n, float mn, float mx, float scale, float &median) {
	size_t k = k0, count = count0;
	const float thresh = 0.5f * n;
	if (k0 > 0 && !(count < thresh)) return false;
	while (count < thresh) { if (k >= k1) return false; count += histo[k++ - k0]; }
	float v;
	if (k > 0) {
		const size_t count_ = count - histo[k - 1 - k0];
		const float c0 = count - thresh;
		const float c1 = thresh - count_;
		v = (c1 * k + c0 * (k - 1)) / (c0 + c1);
	} else v = k;
	v /= scale; v += mn;
	{ const float t = v < mx ? v : mx; v = mn < t ? t : mn; }
	median = v;
	return true;
}

// sorting.c:624-641 on a finished histogram, same window convention
bool finish_hist_median_u16(const unsigned *h, size_t k0, size_t k1, size_t count0, size_t n, double &median) {
	size_t i = k0, j = k0;
	const size_t k = n / 2;
	size_t sum = count0;
	if (k0 > 0 && !(sum <= (n % 2 == 0 ? k - 1 : k))) return false;
	if (n % 2 == 0) { for (; sum <= k - 1; j++) { if (j >= k1) return false; sum += h[j - k0]; } i = j; }
	for (; sum <= k; i++) { if (i >= k1) return false; sum += h[i - k0]; }
	median = (n % 2 == 0) ? (double) (i + j - 2) / 2.0 : (double) (i - 1);
	return true;
}

template <typename T>
int minmax_sel(Sweep &w, const Sel &s, float &mn, float &mx, size_t &cnt) {
	sgpu_ctx *ctx = w.ctx;
	minmax_sel_kernel<T><<<NB, NT, 0, ctx->stream>>>((const T *) w.d, w.n, s, (MMPartial *) w.d_part);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	SGPU_CUDA(ctx, cudaMemcpyAsync(w.h_part, w.d_part, NB * sizeof(MMPartial), cudaMemcpyDeviceToHost, ctx->stream));
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	const MMPartial *p = (const MMPartial *) w.h_part;
	mn = FLT_MAX; mx = -FLT_MAX; cnt = 0;
	for (int i = 0; i < NB; i++) { mn = std::min(mn, p[i].mn); mx = std::max(mx, p[i].mx); cnt += p[i].cnt; }
	if (w.reduce) {
		unsigned long long c = cnt;
		int rv = w.allreduce(&mn, 1, SGPU_RED_F32, SGPU_RED_MIN); if (rv) return rv;
		rv = w.allreduce(&mx, 1, SGPU_RED_F32, SGPU_RED_MAX); if (rv) return rv;
		rv = w.allreduce(&c, 1, SGPU_RED_U64, SGPU_RED_SUM); if (rv) return rv;
		cnt = (size_t) c;
	}
	return SGPU_OK;
}

// launch-side of the windowed histograms: zero [HWIN bins | below], sweep, read back
template <typename Launch>
int run_window(Sweep &w, Launch launch, const unsigned *&h, size_t &below) {
	sgpu_ctx *ctx = w.ctx;
	SGPU_CUDA(ctx, cudaMemsetAsync(w.d_hist, 0, (HWIN + 2) * sizeof(unsigned), ctx->stream));
	launch(w.d_hist, (unsigned long long *) (w.d_hist + HWIN));
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	unsigned *hh = (unsigned *) ctx->h_params;	// pinned: the read-back is a 5 us DMA, not a staged copy
	SGPU_CUDA(ctx, cudaMemcpyAsync(hh, w.d_hist, (HWIN + 2) * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	{ int rv = w.allreduce(hh, HWIN, SGPU_RED_U32, SGPU_RED_SUM); if (rv) return rv; rv = w.allreduce(hh + HWIN, 1, SGPU_RED_U64, SGPU_RED_SUM); if (rv) return rv; }
	unsigned long long b; memcpy(&b, hh + HWIN, sizeof b);
	below = (size_t) b;
	h = hh;
	return SGPU_OK;
}

// full 65536-bin histogram read back into the pinned arena
inline int read_full_hist(Sweep &w, const unsigned *&h) {
	sgpu_ctx *ctx = w.ctx;
	unsigned *hh = (unsigned *) ctx->h_params;
	SGPU_CUDA(ctx, cudaMemcpyAsync(hh, w.d_hist, 65536 * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	{ int rv = w.allreduce(hh, 65536, SGPU_RED_U32, SGPU_RED_SUM); if (rv) return rv; }
	h = hh;
	return SGPU_OK;
}

inline unsigned window_start(double center_bin) {
	const double lo = center_bin - HWIN / 2;
	return lo <= 0 ? 0u : (lo >= 65536 - HWIN ? 65536u - HWIN : (unsigned) lo);
}

// histogram_median_float of the selected / transformed samples (sorting.c:644-649); `guess` is
// where the median is expected (only the speed depends on it)
template <typename T>
int hist_median_float(Sweep &w, const Sel &s, float mn, float mx, size_t cnt, float guess, float &median) {
	sgpu_ctx *ctx = w.ctx;
	if (fabsf(mx - mn) == 0.f) { median = mn; return SGPU_OK; }		// rt_algo.cc:78-81
	const unsigned hs = (unsigned) std::min<size_t>(65536, cnt);
	const float scale = (hs - 1) / (mx - mn);
	const unsigned *h = nullptr;
	if (ctx->smem_optin >= HWIN * sizeof(unsigned) && !ctx->stats_full_hist) {
		const unsigned w0 = window_start((double) scale * ((double) guess - (double) mn));
		size_t below;
		SGPU_CUDA(ctx, cudaFuncSetAttribute(hist_float_win_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (HWIN * sizeof(unsigned))));
		int rv = run_window(w, [&](unsigned *hist, unsigned long long *d_below) {
			hist_float_win_kernel<T><<<HNB, HNT, HWIN * sizeof(unsigned), ctx->stream>>>((const T *) w.d, w.n, s, mn, scale, w0, hist, d_below);
		}, h, below);
		if (rv) return rv;
		if (finish_hist_median_float(h, w0, w0 + HWIN, below, cnt, mn, mx, scale, median)) return SGPU_OK;
		ctx->stats_window_misses++;
	}
	SGPU_CUDA(ctx, cudaMemsetAsync(w.d_hist, 0, 65536 * sizeof(unsigned), ctx->stream));
	hist_float_kernel<T><<<NB, NT, 0, ctx->stream>>>((const T *) w.d, w.n, s, mn, scale, w.d_hist);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	{ int rv = read_full_hist(w, h); if (rv) return rv; }
	if (!finish_hist_median_float(h, 0, 65536, 0, cnt, mn, mx, scale, median))
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "histogram median: counts do not reach half of the samples");
	return SGPU_OK;
}

// IKSSlite (statistics_float.c:199-229) on the non-null samples of the plane (converted with
// `inv` for WORD data, statistics.c:416-426)
template <typename T>
int ikss(Sweep &w, float inv, float median, float mad, double &location, double &scale) {
	sgpu_ctx *ctx = w.ctx;
	const float xlow = median - 6.0 * mad, xhigh = median + 6.0 * mad;
	Sel s{ 1, xlow, xhigh, 0, 0.f, inv };
	float mn, mx; size_t kept;
	int rv = minmax_sel<T>(w, s, mn, mx, kept); if (rv) return rv;
	if (kept == 0) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "IKSS: no sample within 6 MAD of the median");
	float loc;
	rv = hist_median_float<T>(w, s, mn, mx, kept, median, loc); if (rv) return rv;
	location = loc;
	Sel sa{ 1, xlow, xhigh, 1, (float) location, inv };
	float amn, amx; size_t c2;
	rv = minmax_sel<T>(w, sa, amn, amx, c2); if (rv) return rv;
	float mad2;
	rv = hist_median_float<T>(w, sa, amn, amx, kept, mad, mad2); if (rv) return rv;
	if (mad2 == 0.0f) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "MAD is null. Statistics cannot be computed.");
	double bw = 0.0;
	if (mad2 > 0.f) {
		const float factor = 1.f / (9.f * mad2);
		bwmv_kernel<T><<<NB, NT, 0, ctx->stream>>>((const T *) w.d, w.n, s, (float) location, factor, (BwPartial *) w.d_part);
		ctx->launches++;
		SGPU_CUDA(ctx, cudaGetLastError());
		SGPU_CUDA(ctx, cudaMemcpyAsync(w.h_part, w.d_part, NB * sizeof(BwPartial), cudaMemcpyDeviceToHost, ctx->stream));
		SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		const BwPartial *p = (const BwPartial *) w.h_part;
		double up = 0, down = 0;
		for (int i = 0; i < NB; i++) { up += p[i].up; down += p[i].down; }
		{ double ud[2] = { up, down }; rv = w.allreduce(ud, 2, SGPU_RED_F64, SGPU_RED_SUM); if (rv) return rv; up = ud[0]; down = ud[1]; }
		bw = down ? kept * (up / (down * down)) : 0.0;
	}
	scale = sqrt(bw) * .991;
	return SGPU_OK;
}

template <typename T>
int image_stats_impl(sgpu_ctx *ctx, const T *d, long nx, long ny, int lite, sgpu_imstats *out,
		sgpu_reduce_fn reduce = nullptr, void *reduce_user = nullptr, long row0 = 0, long full_ny = 0) {
	const size_t n = (size_t) nx * ny;
	const size_t part_bytes = NB * 64;
	const size_t need = 65536 * sizeof(unsigned) + part_bytes + (size_t) ny * sizeof(double) + 256;
	int rv = sgpu_ensure_scratch(ctx, need); if (rv) return rv;
	Sweep w{ ctx, d, sizeof(T) == 4 ? SGPU_F32 : SGPU_U16, n, (unsigned *) ctx->d_scratch,
		 (char *) ctx->d_scratch + 65536 * sizeof(unsigned), nullptr };
	double *d_rows = (double *) ((char *) w.d_part + part_bytes);
	w.h_part = (char *) ctx->h_params + 512 * 1024;	// pinned arena: [0, 256 KB) histograms, [512 KB, ...) partials and row noise
	w.reduce = reduce; w.reduce_user = reduce_user;
	if (!reduce) { row0 = 0; full_ny = ny; }
	if (row0 < 0 || row0 + ny > full_ny) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "row shard [%ld, %ld) outside the frame of %ld rows", row0, row0 + ny, full_ny);
	memset(out, 0, sizeof(*out));

	// pass A
	basic_kernel<T><<<NB, NT, 0, ctx->stream>>>(d, n, (BasicPartial *) w.d_part);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	SGPU_CUDA(ctx, cudaMemcpyAsync(w.h_part, w.d_part, NB * sizeof(BasicPartial), cudaMemcpyDeviceToHost, ctx->stream));
	// pass B (bgnoise) queued behind it
	if (nx >= 3 && ny > 0) {
		const size_t shm = (size_t) nx * 8 + 16;
		if (shm + 4096 > ctx->smem_optin) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "rows of %ld pixels do not fit in shared memory", nx);	// + the kernel's static arrays
		if (shm + 4096 > 48 * 1024) SGPU_CUDA(ctx, cudaFuncSetAttribute(noise_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm));
		noise_rows_kernel<T><<<(unsigned) std::min<long>(ny, 148 * 8), NT, shm, ctx->stream>>>(d, nx, ny, d_rows);
		ctx->launches++;
		SGPU_CUDA(ctx, cudaGetLastError());
	}
	std::vector<double> rows_big;
	const double *rows = nullptr;
	long nrows = nx >= 3 ? ny : 0;
	if (nrows) {
		double *dst = (double *) ((char *) ctx->h_params + 576 * 1024);
		if ((size_t) nrows * sizeof(double) > ctx->d_params_cap - 576 * 1024) { rows_big.resize(nrows); dst = rows_big.data(); }
		SGPU_CUDA(ctx, cudaMemcpyAsync(dst, d_rows, (size_t) nrows * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
		rows = dst;
	}
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	std::vector<double> rows_all;
	if (reduce && nx >= 3) {	// every rank's row values, in image order: a sum over zero-padded arrays is an all-gather
		rows_all.assign((size_t) full_ny, 0.0);
		for (long i = 0; i < nrows; i++) rows_all[(size_t) (row0 + i)] = rows[i];
		rv = w.allreduce(rows_all.data(), (size_t) full_ny, SGPU_RED_F64, SGPU_RED_SUM); if (rv) return rv;
		rows = rows_all.data(); nrows = full_ny;
	}

	const BasicPartial *bp = (const BasicPartial *) w.h_part;
	double sum = 0, sum2 = 0; unsigned long long ngood = 0, isum = 0, isum2 = 0;
	float mn = FLT_MAX, mx = -FLT_MAX, nzmn = FLT_MAX, nzmx = -FLT_MAX;
	for (int i = 0; i < NB; i++) {
		sum += bp[i].sum; sum2 += bp[i].sum2; ngood += bp[i].ngood; isum += bp[i].isum; isum2 += bp[i].isum2;
		mn = std::min(mn, bp[i].mn); mx = std::max(mx, bp[i].mx); nzmn = std::min(nzmn, bp[i].nzmn); nzmx = std::max(nzmx, bp[i].nzmx);
	}
	unsigned long long total = n;
	if (reduce) {
		double sd[2] = { sum, sum2 }; unsigned long long cu[4] = { ngood, isum, isum2, total }; float lo[2] = { mn, nzmn }, hi[2] = { mx, nzmx };
		rv = w.allreduce(sd, 2, SGPU_RED_F64, SGPU_RED_SUM); if (rv) return rv;
		rv = w.allreduce(cu, 4, SGPU_RED_U64, SGPU_RED_SUM); if (rv) return rv;
		rv = w.allreduce(lo, 2, SGPU_RED_F32, SGPU_RED_MIN); if (rv) return rv;
		rv = w.allreduce(hi, 2, SGPU_RED_F32, SGPU_RED_MAX); if (rv) return rv;
		sum = sd[0]; sum2 = sd[1]; ngood = cu[0]; isum = cu[1]; isum2 = cu[2]; total = cu[3];
		mn = lo[0]; nzmn = lo[1]; mx = hi[0]; nzmx = hi[1];
	}
	if (sizeof(T) == 2) { sum = (double) isum; sum2 = (double) isum2; }
	const double normValue = sizeof(T) == 4 ? 1.0 : 65535.0;
	double mean, sigma;							// quantize.c:293-315
	if (ngood > 1) { mean = sum / ngood; sigma = sqrt((sum2 / ngood) - (mean * mean)); }
	else if (ngood == 1) { mean = sum; sigma = 0.0; }
	else { mean = 0.; sigma = 0.; }
	double bgnoise = 0.0;							// quantize.c:1470-1484
	{
		std::vector<double> v;
		for (long i = 0; i < nrows; i++) if (rows[i] >= 0.0) v.push_back(rows[i]);
		if (v.size() == 1) bgnoise = v[0];
		else if (v.size() > 1) { std::sort(v.begin(), v.end()); bgnoise = (v[(v.size() - 1) / 2] + v[v.size() / 2]) / 2.; }
		bgnoise *= .70710678;
	}
	out->total = (double) total; out->ngoodpix = (double) ngood; out->mean = mean; out->sigma = sigma;
	out->min = mn; out->max = mx; out->normValue = normValue; out->bgnoise = bgnoise;
	if (ngood == 0) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "image has no non-null pixel");	// statistics_float.c:399-403
	if (lite == SGPU_STATS_NOISE_ONLY) return SGPU_OK;	// STATS_BASIC | STATS_BGNOISE: what the noise estimate of a stacked result needs

	double median, mad, location = 0, scale = 0;
	if (sizeof(T) == 4) {
		Sel s{ 0, 0.f, 0.f, 0, 0.f, 1.f };
		float med;
		rv = hist_median_float<T>(w, s, nzmn, nzmx, ngood, (float) mean, med); if (rv) return rv;
		median = med;
		Sel sa{ 0, 0.f, 0.f, 1, med, 1.f };
		float amn, amx, madf; size_t c;
		rv = minmax_sel<T>(w, sa, amn, amx, c); if (rv) return rv;
		rv = hist_median_float<T>(w, sa, amn, amx, ngood, (float) (0.6745 * sigma), madf); if (rv) return rv;
		mad = madf;
		if (!lite) { rv = ikss<T>(w, 1.f, med, madf, location, scale); if (rv) return rv; }
	} else {
		const unsigned *h = nullptr;
		for (int pass = 0; pass < 2; pass++) {
			const int imed = pass ? (int) (median + (median >= 0.0 ? 0.5 : -0.5)) : 0;		// round_to_int
			double m = 0.0;
			bool done = false;
			if (ctx->smem_optin >= HWIN * sizeof(unsigned) && !ctx->stats_full_hist) {
				const unsigned w0 = window_start(pass ? 0.6745 * sigma : mean);
				size_t below;
				SGPU_CUDA(ctx, cudaFuncSetAttribute(hist_u16_win_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (HWIN * sizeof(unsigned))));
				rv = run_window(w, [&](unsigned *hist, unsigned long long *d_below) {
					hist_u16_win_kernel<<<HNB, HNT, HWIN * sizeof(unsigned), ctx->stream>>>((const u16 *) d, n, pass, imed, w0, hist, d_below);
				}, h, below);
				if (rv) return rv;
				done = finish_hist_median_u16(h, w0, w0 + HWIN, below, ngood, m);
				if (!done) ctx->stats_window_misses++;
			}
			if (!done) {
				SGPU_CUDA(ctx, cudaMemsetAsync(w.d_hist, 0, 65536 * sizeof(unsigned), ctx->stream));
				hist_u16_kernel<<<NB, NT, 0, ctx->stream>>>((const u16 *) d, n, pass, imed, w.d_hist);
				ctx->launches++;
				SGPU_CUDA(ctx, cudaGetLastError());
				rv = read_full_hist(w, h); if (rv) return rv;
				if (!finish_hist_median_u16(h, 0, 65536, 0, ngood, m))
					return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "histogram median: counts do not reach half of the samples");
			}
			if (pass == 0) median = m; else mad = (double) (float) m;
		}
		if (!lite) {									// statistics.c:409-438
			const float inv = (float) (1.0 / normValue);
			const float medf = (float) median * inv, madf = (float) mad * inv;
			rv = ikss<T>(w, inv, medf, madf, location, scale); if (rv) return rv;
			location *= normValue; scale *= normValue;
		}
	}
	out->median = median; out->mad = mad; out->location = location; out->scale = scale;
	return SGPU_OK;
}

// one side of a pair (normalization.c:566-586): median, MAD and — unless lite — IKSSlite location / scale of the `cnt`
// non-zero samples of d[0..n); IKSSlite's failures are ignored by the reference, which then keeps location 0 / scale 1
// (or the location it had reached)
int overlap_side(sgpu_ctx *ctx, const float *d, size_t n, int lite, size_t &cnt, float &med, float &mad, float &loc, float &sca) {
	const size_t part_bytes = NB * 64;
	Sweep w{ ctx, d, SGPU_F32, n, (unsigned *) ctx->d_scratch, (char *) ctx->d_scratch + 65536 * sizeof(unsigned), nullptr };
	w.h_part = (char *) ctx->h_params + 512 * 1024;
	(void) part_bytes;
	basic_kernel<float><<<NB, NT, 0, ctx->stream>>>(d, n, (BasicPartial *) w.d_part);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	SGPU_CUDA(ctx, cudaMemcpyAsync(w.h_part, w.d_part, NB * sizeof(BasicPartial), cudaMemcpyDeviceToHost, ctx->stream));
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	const BasicPartial *bp = (const BasicPartial *) w.h_part;
	double sum = 0, sum2 = 0; unsigned long long ngood = 0; float nzmn = FLT_MAX, nzmx = -FLT_MAX;
	for (int i = 0; i < NB; i++) { sum += bp[i].sum; sum2 += bp[i].sum2; ngood += bp[i].ngood; nzmn = std::min(nzmn, bp[i].nzmn); nzmx = std::max(nzmx, bp[i].nzmx); }
	cnt = (size_t) ngood;
	med = mad = loc = sca = 0.f;
	if (cnt <= 3) return SGPU_OK;	// "we want at least 3 pixels with non-zero data to compute stats" (:565)
	const double mean = sum / ngood, sigma = sqrt(std::max(0.0, sum2 / ngood - mean * mean));
	Sel s{ 0, 0.f, 0.f, 0, 0.f, 1.f };
	int rv = hist_median_float<float>(w, s, nzmn, nzmx, cnt, (float) mean, med); if (rv) return rv;
	Sel sa{ 0, 0.f, 0.f, 1, med, 1.f };
	float amn, amx; size_t c;
	rv = minmax_sel<float>(w, sa, amn, amx, c); if (rv) return rv;
	rv = hist_median_float<float>(w, sa, amn, amx, cnt, (float) (0.6745 * sigma), mad); if (rv) return rv;
	if (!lite) {
		double location = 0., scale = 1.;	// `double li = 0., lj = 0., si = 1., sj = 1.;` (:574)
		if (ikss<float>(w, 1.f, med, mad, location, scale)) ctx->err.clear();
		loc = (float) location; sca = (float) scale;
	}
	return SGPU_OK;
}

}  // namespace

/* compute_overlap (normalization.c:420-456): dx* / dy* are translation_from_H of the two frames' homographies */
extern "C" size_t sgpu_compute_overlap(int rxi, int ryi, int rxj, int ryj, double dxi, double dyi, double dxj, double dyj,
		sgpu_rect *areai, sgpu_rect *areaj) {
	auto rnd = [](double x) { x = x > 2147483646.5 ? 2147483646.5 : x; x = x < -2147483647.5 ? -2147483647.5 : x; return (int) (x + (x >= 0.0 ? 0.5 : -0.5)); };	// round_to_int, proto.h:208-213
	int dx = rnd(dxj - dxi), dy = rnd(dyi - dyj);
	const int x_tli = std::max(0, dx), y_tli = std::max(0, dy);
	const int x_bri = std::min(rxi, dx + rxj), y_bri = std::min(ryi, dy + ryj);
	const int x_tlj = std::max(0, -dx), y_tlj = std::max(0, -dy);
	const int x_brj = std::min(rxj, -dx + rxi), y_brj = std::min(ryj, -dy + ryi);
	if (x_tli < x_bri && y_tli < y_bri) {
		if (areai) *areai = sgpu_rect{ x_tli, y_tli, x_bri - x_tli, y_bri - y_tli };
		if (areaj) *areaj = sgpu_rect{ x_tlj, y_tlj, x_brj - x_tlj, y_brj - y_tlj };
		return (size_t) ((x_bri - x_tli) * (y_bri - y_tli));	// an int product, as in the reference
	}
	return 0;
}

/* the statistics of one pair of frames on one layer (_compute_estimators_for_images, normalization.c:549-593) */
extern "C" int sgpu_overlap_stats_pair(sgpu_ctx *ctx, const void *plane_i, int rxi, int ryi, const void *plane_j, int rxj, int ryj,
		int on_device, int dtype, const sgpu_rect *areai, const sgpu_rect *areaj, int lite, sgpu_overlap_stats *out) {
	if (!ctx || !plane_i || !plane_j || !areai || !areaj || !out) return SGPU_GENERIC_ERROR;
	if (dtype != SGPU_F32 && dtype != SGPU_U16) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "bad dtype");
	memset(out, 0, sizeof *out);
	const long w = areai->w, h = areai->h;
	if (w != areaj->w || h != areaj->h || w < 1 || h < 1 || areai->x < 0 || areai->y < 0 || areaj->x < 0 || areaj->y < 0 ||
	    areai->x + w > rxi || areai->y + h > ryi || areaj->x + w > rxj || areaj->y + h > ryj)
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "overlap rectangles do not fit the frames");
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const size_t es = dtype == SGPU_F32 ? 4 : 2, n = (size_t) w * h;
	void *staged[2] = { nullptr, nullptr };
	const void *src[2] = { plane_i, plane_j };
	const int rx[2] = { rxi, rxj }, ry[2] = { ryi, ryj };
	const sgpu_rect *ar[2] = { areai, areaj };
	const char *first[2];	// the rectangle's first sample in memory order: rows are bottom-up, area.y counts from the top
	long pitch[2];
	for (int k = 0; k < 2; k++) {
		const long row0 = ry[k] - ar[k]->y - h;
		if (on_device) { first[k] = (const char *) src[k] + ((size_t) row0 * rx[k] + ar[k]->x) * es; pitch[k] = rx[k]; continue; }
		// host planes: only the rows of the rectangle travel (readfits_partial_all_layers reads just that region too)
		cudaError_t e = cudaMalloc(&staged[k], (size_t) h * rx[k] * es);
		if (e == cudaSuccess) e = cudaMemcpyAsync(staged[k], (const char *) src[k] + (size_t) row0 * rx[k] * es, (size_t) h * rx[k] * es, cudaMemcpyHostToDevice, ctx->stream);
		if (e != cudaSuccess) { cudaFree(staged[0]); cudaFree(staged[1]); return sgpu_cuda_fail(ctx, e, "upload overlap rows"); }
		first[k] = (const char *) staged[k] + (size_t) ar[k]->x * es; pitch[k] = rx[k];
	}
	// scratch: [histogram | partials | pair arrays]
	const size_t head = 65536 * sizeof(unsigned) + NB * 64 + 256;
	int rv = sgpu_ensure_scratch(ctx, head + 2 * n * sizeof(float));
	if (!rv) {
		float *oi = (float *) ((char *) ctx->d_scratch + head), *oj = oi + n;
		const float inv = 1. / 65535.0;	// `float invnorm = 1. / USHRT_MAX;` (:471)
		const unsigned grid = (unsigned) std::min<size_t>((n + 255) / 256, (size_t) ctx->sm_count * 16);
		if (dtype == SGPU_F32) overlap_pair_kernel<float><<<grid, 256, 0, ctx->stream>>>((const float *) first[0], (const float *) first[1], pitch[0], pitch[1], w, n, 1.f, oi, oj);
		else overlap_pair_kernel<u16><<<grid, 256, 0, ctx->stream>>>((const u16 *) first[0], (const u16 *) first[1], pitch[0], pitch[1], w, n, inv, oi, oj);
		ctx->launches++;
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) rv = sgpu_cuda_fail(ctx, e, "overlap_pair_kernel");
		size_t ci = 0, cj = 0;
		if (!rv) rv = overlap_side(ctx, oi, n, lite, ci, out->medij, out->madij, out->locij, out->scaij);
		if (!rv) rv = overlap_side(ctx, oj, n, lite, cj, out->medji, out->madji, out->locji, out->scaji);
		if (!rv) out->Nij = ci > 3 ? ci : 0;	// seq->ostats[n][ijth].Nij is only set when there are more than 3 pairs
	}
	if (staged[0] || staged[1]) { cudaStreamSynchronize(ctx->stream); cudaFree(staged[0]); cudaFree(staged[1]); }
	return rv;
}

static int image_stats_any(sgpu_ctx *ctx, const void *plane, int on_device, int dtype, long nx, long ny, int lite, sgpu_imstats *out,
		sgpu_reduce_fn reduce, void *user, long row0, long full_ny);

extern "C" int sgpu_image_stats(sgpu_ctx *ctx, const void *plane, int on_device, int dtype, long nx, long ny,
		int lite, sgpu_imstats *out) {
	return image_stats_any(ctx, plane, on_device, dtype, nx, ny, lite, out, nullptr, nullptr, 0, ny);
}

extern "C" int sgpu_image_stats_sharded(sgpu_ctx *ctx, const void *shard, int on_device, int dtype, long nx, long ny,
		long row0, long full_ny, int lite, sgpu_reduce_fn reduce, void *user, sgpu_imstats *out) {
	if (!reduce) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "sgpu_image_stats_sharded needs an all-reduce callback");
	return image_stats_any(ctx, shard, on_device, dtype, nx, ny, lite, out, reduce, user, row0, full_ny);
}

static int image_stats_any(sgpu_ctx *ctx, const void *plane, int on_device, int dtype, long nx, long ny, int lite, sgpu_imstats *out,
		sgpu_reduce_fn reduce, void *user, long row0, long full_ny) {
	if (!ctx || !plane || !out || nx < 1 || ny < 1) return SGPU_GENERIC_ERROR;
	if (dtype != SGPU_F32 && dtype != SGPU_U16) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "bad dtype");
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const size_t bytes = (size_t) nx * ny * (dtype == SGPU_F32 ? 4 : 2);
	void *staged = nullptr;
	const void *d = plane;
	if (!on_device) {
		SGPU_CUDA(ctx, cudaMalloc(&staged, bytes));
		cudaError_t e = cudaMemcpyAsync(staged, plane, bytes, cudaMemcpyHostToDevice, ctx->stream);
		if (e != cudaSuccess) { cudaFree(staged); return sgpu_cuda_fail(ctx, e, "upload plane"); }
		d = staged;
	}
	int rv = dtype == SGPU_F32 ? image_stats_impl<float>(ctx, (const float *) d, nx, ny, lite, out, reduce, user, row0, full_ny)
				   : image_stats_impl<uint16_t>(ctx, (const uint16_t *) d, nx, ny, lite, out, reduce, user, row0, full_ny);
	if (staged) { cudaStreamSynchronize(ctx->stream); cudaFree(staged); }
	return rv;
}
