// exact.cuh — thread-per-pixel engine: one CUDA thread owns one pixel stack, held as a strided
// column (shared memory when it fits, global scratch otherwise), and walks it in exactly the
// element order the reference's per-pixel code does, so that every order-dependent behaviour —
// the permutation quickselect leaves behind, the `N - r <= 4` guard evaluated in stack order,
// the stack state the "kept == 0" fallback sees — comes out identical.  It covers every
// rejection type and both sample types, and it is what the warp-cooperative engine falls back
// to for pixels whose result would depend on stack order.
//
// Reference: src/stacking/rejection_float.c:100-354 (float), src/stacking/median_and_mean.c:
// 590-954 (WORD), :956-1101 (combine), src/algos/sorting.c, src/algos/statistics.h:82-107,
// src/algos/statistics.c:100-154, src/rt/rt_algo.cc:38-171, src/stacking/siril_fit_linear.c.
// Compiled with -fmad=false: the reference's float expressions are not contracted.
#pragma once
#include <cstdint>
#include <cfloat>
#include <cuda_runtime.h>

#define SGPU_HD __host__ __device__

namespace sgpu {

typedef uint16_t u16;

// strided view of one thread's column
template <typename T>
struct Col {
	T *p; int s;
	SGPU_HD __forceinline__ T get(int i) const { return p[(unsigned) i * (unsigned) s]; }
	SGPU_HD __forceinline__ void set(int i, T v) const { p[(unsigned) i * (unsigned) s] = v; }
};

struct RejCfg {
	int type; float s0, s1; const float *crit; int nb_frames; float m_x, m_dx2;
	const float *xf;	// optional table 1.f / (i + 1) (median_and_mean.c:1493) for the linear-fit running means; nullptr: computed in place
	const double *rcp;	// optional table of correctly rounded 1.0 / i, i < SGPU_RCP_N (device memory); nullptr: computed in place
};
#define SGPU_RCP_N 1282

// ---- rounding helpers (src/core/proto.h:208-237, 324-346) ----
SGPU_HD __forceinline__ u16 round_to_word(double x) {
	x += 0.5; x = x > 65535.0 ? 65535.0 : x; x = x < 0.0 ? 0.0 : x; return (u16) x;
}
SGPU_HD __forceinline__ u16 roundf_to_word(float f) {
	f += 0.5f; f = f > 65535.f ? 65535.f : f; f = f < 0.f ? 0.f : f; return (u16) f;
}
SGPU_HD __forceinline__ int round_to_int_d(double x) {
	x = x > 2147483646.5 ? 2147483646.5 : x; x = x < -2147483647.5 ? -2147483647.5 : x;
	return (int) (x + (x >= 0.0 ? 0.5 : -0.5));
}
SGPU_HD __forceinline__ u16 truncate_to_word(int x) { x = x < 0 ? 0 : x; x = x > 65535 ? 65535 : x; return (u16) x; }

SGPU_HD __forceinline__ float to_f(float v) { return v; }
SGPU_HD __forceinline__ float to_f(u16 v) { return (float) v; }

// ---- sorting / selection ----
// ascending in-place sort; the result of a sort is unique so the scheme is free: insertion for
// short columns, heapsort beyond (sorting.c:110-186 quicksort_f/_s)
template <typename T>
SGPU_HD void sort_col(Col<T> a, int n) {
	if (n <= 16) {
		for (int i = 1; i < n; i++) {
			T v = a.get(i); int j = i - 1;
			while (j >= 0 && a.get(j) > v) { a.set(j + 1, a.get(j)); --j; }
			a.set(j + 1, v);
		}
		return;
	}
	for (int start = n / 2 - 1; start >= 0; --start) {	// heapify
		int root = start; T v = a.get(root);
		for (;;) {
			int child = 2 * root + 1; if (child >= n) break;
			T c = a.get(child);
			if (child + 1 < n) { T c2 = a.get(child + 1); if (c2 > c) { c = c2; child++; } }
			if (!(c > v)) break;
			a.set(root, c); root = child;
		}
		a.set(root, v);
	}
	for (int end = n - 1; end > 0; --end) {
		T v = a.get(end); a.set(end, a.get(0));
		int root = 0;
		for (;;) {
			int child = 2 * root + 1; if (child >= end) break;
			T c = a.get(child);
			if (child + 1 < end) { T c2 = a.get(child + 1); if (c2 > c) { c = c2; child++; } }
			if (!(c > v)) break;
			a.set(root, c); root = child;
		}
		a.set(root, v);
	}
}

// quickmedian_float / quickmedian (sorting.c:240-273, :195-230) and their n < 9 sorting-network
// front ends (:366-415, :468-513), reproduced including the permutation left in `a`
template <typename T> struct MedSmall;
template <> struct MedSmall<float> {	// float network averages the middles IN FLOAT
	static SGPU_HD __forceinline__ double even(float x, float y) { return (x + y) / 2.0; }
	static SGPU_HD __forceinline__ double evenq(float x, float y) { return ((double) x + y) / 2.0; }
};
template <> struct MedSmall<u16> {
	static SGPU_HD __forceinline__ double even(u16 x, u16 y) { return ((int) x + (int) y) / 2.0; }
	static SGPU_HD __forceinline__ double evenq(u16 x, u16 y) { return ((double) x + (double) y) / 2.0; }
};

template <typename T>
SGPU_HD double quickmedian_col(Col<T> a, int n) {
	if (n <= 0) return 0.0;
	const int k = n / 2;
	if (n < 9) {
		if (n == 1) return (double) a.get(0);
		sort_col(a, n);
		return (n % 2 == 0) ? MedSmall<T>::even(a.get(k - 1), a.get(k)) : (double) a.get(k);
	}
	int lo = 0, hi = n - 1;
	while (lo < hi) {
		const int mid = (lo + hi) / 2;
		const T pv = a.get(mid);
		a.set(mid, a.get(hi)); a.set(hi, pv);
		int store = lo;
		for (int i = lo; i < hi; i++) {
			const T v = a.get(i);
			if (v < pv) { const T t = a.get(store); a.set(store, v); a.set(i, t); store++; }
		}
		a.set(hi, a.get(store)); a.set(store, pv);
		if (store < k) lo = store + 1; else hi = store;
	}
	return (n % 2 == 0) ? MedSmall<T>::evenq(a.get(k - 1), a.get(k)) : (double) a.get(k);
}

// ---- per-stack statistics ----
// siril_stats_float_sd (statistics.h:82-107); `cl`/`cu` apply the accumulated winsorization
// clamp on the fly (the reference's w_stack, rejection_float.c:228-236)
SGPU_HD __forceinline__ float clampw(float v, float lo, float hi) { float t = lo > v ? lo : v; return hi < t ? hi : t; }
SGPU_HD inline float sd_f32(Col<float> d, int N, float *m, float cl = -FLT_MAX, float cu = FLT_MAX, int off = 0) {
	double sum = 0.0, vsum = 0.0;
	for (int i = 0; i < N; i++) sum += (double) clampw(d.get(off + i), cl, cu);
	const float mean = (float) (sum / N);
	for (int i = 0; i < N; i++) { const float e = clampw(d.get(off + i), cl, cu) - mean; vsum += (double) (e * e); }
	if (m) *m = mean;
	return sqrtf((float) (vsum / (N - 1)));
}
// siril_stats_ushort_sd_32 / _64 (statistics.c:100-127)
SGPU_HD inline float sd_u16(Col<u16> d, int N, u16 cl = 0, u16 cu = 65535) {
	unsigned long long isum = 0;
	for (int i = 0; i < N; i++) { u16 v = d.get(i); v = v < cl ? cl : v; v = v > cu ? cu : v; isum += v; }
	const float mean = (float) (((double) isum) / ((double) N));
	double acc = 0.0;
	for (int i = 0; i < N; i++) {
		u16 v = d.get(i); v = v < cl ? cl : v; v = v > cu ? cu : v;
		const float px = (float) v; acc += (px - mean) * (px - mean);
	}
	return sqrtf((float) (acc / (N - 1)));
}
// the GESDT-only WORD variant (median_and_mean.c:649-662) on the window [off, off+N)
SGPU_HD inline float sd_u16_gesdt(Col<u16> d, int off, int N, float *m) {
	double acc = 0.0;
	for (int i = 0; i < N; i++) acc += d.get(off + i);
	const float mean = (float) (acc / N);
	acc = 0.0;
	for (int i = 0; i < N; i++) { const float e = d.get(off + i) - mean; acc += e * e; }
	*m = mean;
	return sqrtf((float) (acc / (N - 1)));
}

// siril_stats_float_mad -> histogram_median_float -> findMinMaxPercentile (statistics_float.c:
// 79-101, sorting.c:644-649, rt_algo.cc:38-171): approximate median of |x - median| from an
// n-bin histogram with in-bin interpolation.  `aux` receives the bin index of every sample and
// is then sorted, which yields the same cumulative counts as the reference's histogram.
SGPU_HD inline float mad_f32(Col<float> d, int n, double m, Col<u16> aux) {
	const float med = (float) m;
	float mn = fabsf(d.get(0) - med), mx = mn;
	for (int i = 1; i < n; i++) { const float v = fabsf(d.get(i) - med); mn = v < mn ? v : mn; mx = mx < v ? v : mx; }
	if (fabsf(mx - mn) == 0.f) return mn;
	const unsigned hs = n < 65536 ? (unsigned) n : 65536u;
	const float scale = (hs - 1) / (mx - mn);
	for (int i = 0; i < n; i++) aux.set(i, (u16) (scale * (fabsf(d.get(i) - med) - mn)));
	sort_col(aux, n);
	const float thresh = 0.5f * n;
	// smallest k with count(bin < k) >= thresh; sorted bins: count(bin < k) = lower_bound(k)
	int idx = 0; unsigned kbin = 0, count = 0, last = 0;
	while ((float) count < thresh) {		// bins kbin = 0,1,2,... as `count += histo[k++]`
		unsigned c = 0;
		while (idx < n && aux.get(idx) == kbin) { idx++; c++; }
		count += c; last = c; kbin++;
	}
	float v;
	if (kbin > 0) {
		const unsigned count_ = count - last;
		const float c0 = (float) count - thresh;
		const float c1 = thresh - (float) count_;
		v = (c1 * (float) kbin + c0 * (float) (kbin - 1)) / (c0 + c1);
	} else v = 0.f;
	v /= scale; v += mn;
	{ const float t = v < mx ? v : mx; v = mn < t ? t : mn; }
	return v;
}
// siril_stats_ushort_mad (statistics.c:133-154): exact median of |x - round(median)| (the
// 65536-bin histogram median of sorting.c:577-642 is the exact sorted median)
SGPU_HD inline float mad_u16(Col<u16> d, int n, double m, Col<u16> aux) {
	const int med = round_to_int_d(m);
	for (int i = 0; i < n; i++) { int v = (int) d.get(i) - med; aux.set(i, (u16) (v < 0 ? -v : v)); }
	sort_col(aux, n);
	const int k = n / 2;
	const double r = (n % 2 == 0) ? ((int) aux.get(k - 1) + (int) aux.get(k)) / 2.0 : (double) aux.get(k);
	return (float) r;
}

// siril_fit_linear (siril_fit_linear.c:25-48) with xf[i] = 1.f / (i + 1)
template <typename T>
SGPU_HD void fit_linear(Col<T> y, float m_x, float m_dx2, int n, float *c0, float *c1) {
	float m_y = to_f(y.get(0));
	for (int i = 1; i < n; i++) m_y += (to_f(y.get(i)) - m_y) * (1.f / (i + 1));
	float m_dxdy = 0.f, dx = -m_x;
	for (int i = 0; i < n; i++, dx += 1.f) {
		const float dy = to_f(y.get(i)) - m_y;
		m_dxdy += (dx * dy - m_dxdy) * (1.f / (i + 1));
	}
	const float b = m_dxdy * m_dx2;
	*c0 = m_y - m_x * b; *c1 = b;
}

// sigma_clipping_float (rejection_float.c:49-60) / sigma_clipping (median_and_mean.c:608-621)
SGPU_HD __forceinline__ int sclip(float px, float sigma, float lo, float hi, float med, int rej[2]) {
	if (med - px > sigma * lo) { rej[0]++; return -1; }
	else if (px - med > sigma * hi) { rej[1]++; return 1; }
	return 0;
}
SGPU_HD __forceinline__ int sclip(u16 px, float sigma, float lo, float hi, float med, int rej[2]) {
	if (med - px > lo * sigma) { rej[0]++; return -1; }
	else if (px - med > hi * sigma) { rej[1]++; return 1; }
	return 0;
}

// the shared "clip with the N - r <= 4 guard, compact in place" pass of SIGMA / MAD / WINSORIZED
template <typename T>
SGPU_HD __forceinline__ int clip_pass(Col<T> st, int N, int &r, float var, float lo, float hi, float med, int rej[2]) {
	int o = 0;
	for (int f = 0; f < N; f++) {
		const T v = st.get(f);
		int flag = 0;
		if (N - r > 4) { flag = sclip(v, var, lo, hi, med, rej); if (flag) r++; }
		if (!flag) { if (f != o) st.set(o, v); o++; }
	}
	return o;
}

template <typename T> struct Pre { };
template <> struct Pre<float> { static SGPU_HD __forceinline__ bool need(int t) { return t == 1 || t == 2 || t == 3; } };	// rejection_float.c:147-157
template <> struct Pre<u16> { static SGPU_HD __forceinline__ bool need(int t) { return t >= 1 && t <= 5; } };		// median_and_mean.c:747-759

// Per-sample weights of the weighted combine (median_and_mean.c:1043-1070).  The combine takes any type with
// `bool(w)` and `w[f]`: a plain `const double *` is the per-frame weight alone (args->weights + layer * N);
// PixWeights adds the feather mask and the drizzle weight of the sample (data->mstack / data->dstack, gathered at
// the pixel's shifted column like the sample itself, :1687-1692): n = 1 * d * m * w in the reference's order.
struct PixWeights {
	const double *w;	// frame weights of this layer, or null (NO_WEIGHT)
	const float *m, *d;	// frame 0's mask / drizzle sample of this pixel, or null
	size_t ms, ds;		// frame strides of the two plane sets, in floats
	const int *shiftx;	// per-frame x shift, or null
	SGPU_HD explicit operator bool() const { return w || m || d; }
	SGPU_HD float dval(int f) const { return d[(size_t) f * ds - (shiftx ? shiftx[f] : 0)]; }
	SGPU_HD double operator[](int f) const {	// only asked for samples inside their frame (non-zero ones)
		double n = 1.;
		if (d) n *= dval(f);
		if (m) n *= m[(size_t) f * ms - (shiftx ? shiftx[f] : 0)];
		if (w) n *= w[f];
		return n;
	}
};
// "remove null pixels (or null drizzle weights or both)" (rejection_float.c:116-126, median_and_mean.c:716-726)
SGPU_HD __forceinline__ bool drizzle_null(const double *, int) { return false; }
SGPU_HD __forceinline__ bool drizzle_null(const PixWeights &w, int f) { return w.d && w.dval(f) == 0.f; }

// apply_rejection_float / apply_rejection_ushort.  Returns the kept count; `st` holds the kept
// samples in front exactly as the reference leaves them.
template <typename T, typename W = const double *>
SGPU_HD int apply_rejection(Col<T> st, int nb, const RejCfg &c, Col<u16> aux, int rej[2], const W &w = W()) {
	int N = nb, kept = 0, r = 0, first = 1, changed;
	for (int f = 0; f < N; f++) { const T v = st.get(f); if (v != (T) 0 && !drizzle_null(w, f)) { if (f != kept) st.set(kept, v); kept++; } }
	if (kept <= 1) return kept;
	const int removed = N - kept;
	N = kept;
	double median = 0.0;	// the WORD path keeps it as float (median_and_mean.c:705): fmed below
	float fmed = 0.f;
	if (Pre<T>::need(c.type)) {
		median = quickmedian_col(st, N);
		fmed = (float) median;
		if (sizeof(T) == 4 ? (median == 0.0) : (fmed == 0.f)) return 0;
	}
	switch (c.type) {
	case 1: {	// PERCENTILE (rejection_float.c:31-44,160-173 / median_and_mean.c:590-603,762-776)
		int o = 0;
		for (int f = 0; f < N; f++) {
			const T v = st.get(f); int flag = 0;
			if (sizeof(T) == 4) {
				if (fmed - to_f(v) > fmed * c.s0) { rej[0]++; flag = -1; }
				else if (to_f(v) - fmed > fmed * c.s1) { rej[1]++; flag = 1; }
			} else {
				if ((fmed - to_f(v)) / fmed > c.s0) { rej[0]++; flag = -1; }
				else if ((to_f(v) - fmed) / fmed > c.s1) { rej[1]++; flag = 1; }
			}
			if (!flag) { if (f != o) st.set(o, v); o++; }
		}
		N = o;
		break; }
	case 2: case 3:	// SIGMA, MAD (rejection_float.c:174-209 / median_and_mean.c:777-813)
		do {
			float var;
			if (sizeof(T) == 4) {
				Col<float> sf = { (float *) st.p, st.s };
				var = (c.type == 2) ? sd_f32(sf, N, nullptr) : mad_f32(sf, N, median, aux);
			} else {
				Col<u16> sw = { (u16 *) st.p, st.s };
				var = (c.type == 2) ? sd_u16(sw, N) : mad_u16(sw, N, sizeof(T) == 4 ? median : (double) fmed, aux);
			}
			if (!first) { median = quickmedian_col(st, N); fmed = (float) median; } else first = 0;
			const int o = clip_pass(st, N, r, var, c.s0, c.s1, fmed, rej);
			changed = N != o; N = o;
		} while (changed && N > 3);
		break;
	case 4: {	// SIGMEDIAN (rejection_float.c:210-222 / median_and_mean.c:814-830)
		int n;
		do {
			float sigma;
			if (sizeof(T) == 4) { Col<float> sf = { (float *) st.p, st.s }; sigma = sd_f32(sf, N, nullptr); fmed = (float) quickmedian_col(st, N); }
			else {
				Col<u16> sw = { (u16 *) st.p, st.s }; sigma = sd_u16(sw, N);
				if (!first) fmed = (float) quickmedian_col(st, N); else first = 0;
			}
			n = 0;
			for (int f = 0; f < N; f++)
				if (sclip(st.get(f), sigma, c.s0, c.s1, fmed, rej)) { st.set(f, (T) fmed); n++; }
		} while (n > 0);
		break; }
	case 5:		// WINSORIZED (rejection_float.c:223-259 / median_and_mean.c:831-867)
		do {
			float sigma, sigma0;
			if (sizeof(T) == 4) {
				Col<float> sf = { (float *) st.p, st.s };
				sigma = sd_f32(sf, N, nullptr);
				fmed = (float) quickmedian_col(st, N);
				float cl = -FLT_MAX, cu = FLT_MAX;	// composition of the successive clamps on w_stack
				do {
					const float m0 = fmed - 1.5f * sigma, m1 = fmed + 1.5f * sigma;
					cl = clampw(cl, m0, m1); cu = clampw(cu, m0, m1);
					sigma0 = sigma;
					sigma = 1.134f * sd_f32(sf, N, nullptr, cl, cu);
#ifdef SGPU_DEBUG_WINSOR
					printf("winsor N=%d med=%.9g m0=%.9g m1=%.9g cl=%.9g cu=%.9g sigma0=%.9g sigma=%.9g\n", N, fmed, m0, m1, cl, cu, sigma0, sigma);
#endif
				} while (fabsf(sigma - sigma0) > sigma0 * 0.0005f);
			} else {
				Col<u16> sw = { (u16 *) st.p, st.s };
				sigma = sd_u16(sw, N);
				if (!first) fmed = (float) quickmedian_col(st, N); else first = 0;
				u16 cl = 0, cu = 65535;
				do {
					const u16 m0 = roundf_to_word(fmed - 1.5f * sigma), m1 = roundf_to_word(fmed + 1.5f * sigma);
					cl = cl < m0 ? m0 : cl; cl = cl > m1 ? m1 : cl;
					cu = cu < m0 ? m0 : cu; cu = cu > m1 ? m1 : cu;
					sigma0 = sigma;
					sigma = 1.134f * sd_u16(sw, N, cl, cu);
				} while (fabs((double) (sigma - sigma0)) > (double) (sigma0 * 0.0005f));
			}
			const int o = clip_pass(st, N, r, sigma, c.s0, c.s1, fmed, rej);
			changed = N != o; N = o;
		} while (changed && N > 3);
		break;
	case 6:		// LINEARFIT (rejection_float.c:260-300 / median_and_mean.c:868-902)
		do {
			sort_col(st, N);
			float a, b;
			fit_linear(st, c.m_x, c.m_dx2, N, &b, &a);
			float sigma = 0.f;
			for (int f = 0; f < N; f++) sigma += fabsf(to_f(st.get(f)) - (a * f + b));
			sigma /= (float) N;
			int o = 0;
			for (int f = 0; f < N; f++) {
				const T v = st.get(f); int flag = 0;
				if (N - r > 4) {
					if (a * f + b - to_f(v) > sigma * c.s0) { rej[0]++; flag = -1; r++; }
					else if (to_f(v) - a * f - b > sigma * c.s1) { rej[1]++; flag = 1; r++; }
				}
				if (!flag) { if (f != o) st.set(o, v); o++; }
			}
			changed = N != o; N = o;
		} while (changed && N > 3);
		break;
	case 7: {	// GESDT (rejection_float.c:301-348 / median_and_mean.c:903-948)
		sort_col(st, N);
		if (N % 2) median = (double) st.get(N / 2);
		else median = MedSmall<T>::even(st.get((N - 1) / 2), st.get(N / 2));	// gsl median_from_sorted_data
		if (sizeof(T) == 2) median = (double) (float) median;
		int max_out = (int) ((float) nb * c.s0);
		if (removed >= max_out) return kept;
		max_out -= removed;
		// pass 1: run the removal sequence (w_stack is the window [lo, lo+size) of the sorted
		// stack), find the last iteration confirm_outliers (median_and_mean.c:685-701) accepts
		int last = -1;
		{
			int lo = 0;
			for (int it = 0, size = N; it < max_out; it++, size--) {
				float avg, sd;
				if (sizeof(T) == 4) { Col<float> sf = { (float *) st.p, st.s }; sd = sd_f32(sf, size, &avg, -FLT_MAX, FLT_MAX, lo); }
				else { Col<u16> sw = { (u16 *) st.p, st.s }; sd = sd_u16_gesdt(sw, lo, size, &avg); }
				float dev = avg - to_f(st.get(lo));
				const float md2 = to_f(st.get(lo + size - 1)) - avg;
				int low_end = 1;
				if (md2 > dev) { dev = md2; low_end = 0; }
				const float G = dev / sd;
				if (it > 1 && G > c.crit[it + removed]) last = it;
				if (low_end) lo++;
			}
			if (last < 0) last = max_out - 1 < 1 ? max_out - 1 : 1;
		}
		// pass 2: replay iterations 0..last, flag positions as the reference indexes them
		// (low removals: cold++; high removals: index inside the shrunk window, :333)
		for (int i = 0; i < N; i++) aux.set(i, 0);
		{
			int lo = 0, cold = 0;
			for (int it = 0, size = N; it <= last; it++, size--) {
				float avg;
				if (sizeof(T) == 4) { Col<float> sf = { (float *) st.p, st.s }; sd_f32(sf, size, &avg, -FLT_MAX, FLT_MAX, lo); }
				else { Col<u16> sw = { (u16 *) st.p, st.s }; float sd = sd_u16_gesdt(sw, lo, size, &avg); (void) sd; }
				const float dev = avg - to_f(st.get(lo));
				const float md2 = to_f(st.get(lo + size - 1)) - avg;
				float x; int idx;
				if (md2 > dev) { x = to_f(st.get(lo + size - 1)); idx = size - 1; }
				else { x = to_f(st.get(lo)); idx = cold++; lo++; }
				if ((double) x >= median) rej[1]++; else rej[0]++;
				aux.set(idx, 1);
			}
		}
		int o = 0;
		for (int f = 0; f < N; f++) if (!aux.get(f)) { if (f != o) st.set(o, st.get(f)); o++; }
		N = o;
		break; }
	default: break;
	}
	return N;
}

// mean_and_reject (median_and_mean.c:956-1101).  `orig(f)` re-reads the normalized sample of
// frame f (the reference's o_stack copy); w == nullptr is NO_WEIGHT.
template <typename T, typename OrigFn, typename W>
SGPU_HD double mean_and_reject(Col<T> st, int nb, const RejCfg &c, const W &w, Col<u16> aux,
		OrigFn orig, int rej[2]) {
	const int kept = apply_rejection(st, nb, c, aux, rej, w);
	if (kept == 0) return quickmedian_col(st, nb);
	if (w) {
		if (sizeof(T) == 4) {
			float pmin = FLT_MAX, pmax = -FLT_MAX;
			for (int f = 0; f < kept; f++) { const float v = to_f(st.get(f)); if (pmin > v) pmin = v; if (pmax < v) pmax = v; }
			double sum = 0.0, norm = 0.0;
			for (int f = 0; f < nb; f++) {
				const float v = to_f(orig(f));
				if (v >= pmin && v <= pmax && v != 0.f) { const double n = w[f]; sum += (double) v * n; norm += n; }
			}
			if (norm == 0. || sum == 0.) {
				sum = 0.;
				for (int f = 0; f < nb; f++) { const float v = to_f(orig(f)); if (v >= pmin && v <= pmax && v > 0) sum += (double) v; }
				return sum / (double) kept;
			}
			return sum / norm;
		} else {
			unsigned pmin = 65535, pmax = 0;
			for (int f = 0; f < kept; f++) { const unsigned v = (unsigned) st.get(f); if (pmin > v) pmin = v; if (pmax < v) pmax = v; }
			double sum = 0.0, norm = 0.0;
			for (int f = 0; f < nb; f++) {
				const unsigned v = (unsigned) orig(f);
				if (v >= pmin && v <= pmax && v > 0) { const double n = w[f]; sum += (double) v * n; norm += n; }
			}
			if (norm == 0. || sum == 0.) {
				sum = 0.;
				for (int f = 0; f < nb; f++) { const unsigned v = (unsigned) orig(f); if (v >= pmin && v <= pmax && v > 0) sum += (double) v; }
				return sum / (double) kept;
			}
			return sum / norm;
		}
	}
	if (sizeof(T) == 4) {
		double sum = 0.0;
		for (int f = 0; f < kept; f++) sum += (double) to_f(st.get(f));
		return sum / (double) kept;
	} else {
		long long sum = 0;
		for (int f = 0; f < kept; f++) sum += (long long) st.get(f);
		return sum / (double) kept;
	}
}

}  // namespace sgpu
