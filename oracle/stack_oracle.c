/* TEST INFRASTRUCTURE ONLY — CPU restatement (plain C) of Siril's rejection-stacking hot path.
 *
 * This file is the ORACLE the CUDA path is checked against.  It is a from-scratch restatement
 * of the reference algorithm (lock042/siril @ a5869d93); every function cites the reference
 * file:line it follows.  It is pinned (tests/test_oracle_*.py) against
 *   (1) the reference's own golden vectors: src/tests/rejection_test.c (set1/set2),
 *       src/tests/sorting.c (median property), src/tests/stacking_blocks_test.c (planner), and
 *   (2) the reference's own code compiled here from /root/reference (oracle/_ref, see
 *       build_ref.py) on randomized stacks, images and planner inputs.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product path (siril_b200/, libsgpu.so) never links or calls it.
 *
 * Third-party arithmetic: GSL (un-vendored, unpinned system dependency of the reference,
 * meson.build:284).  gsl_stats_*_median_from_sorted_data is restated from its published
 * definition; gsl_cdf_tdist_Pinv is NOT restated — the GESDT critical values are an input
 * array here exactly as they are in struct stacking_args (stacking.h:96), computed by the
 * harness with scipy on the float-rounded probability (median_and_mean.c:1482).
 *
 * Build: gcc -O2 -fopenmp -fno-math-errno -ffp-contract=off -shared -fPIC (oracle/Makefile).
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <limits.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef uint16_t u16;

/* enum values: src/core/settings.h:34-53, src/stacking/stacking.h:46-52 */
enum { R_NONE = 0, R_PERCENTILE, R_SIGMA, R_MAD, R_SIGMEDIAN, R_WINSORIZED, R_LINEARFIT, R_GESDT };
enum { NRM_NONE = 0, NRM_ADD, NRM_MUL, NRM_ADDSCALE, NRM_MULSCALE };

/* ------------------------------------------------------------------ rounding helpers */
/* src/core/proto.h:232-237 */
static u16 rnd_word(double x) {
	x += 0.5;
	if (x > 65535.0) x = 65535.0;
	if (x < 0.0) x = 0.0;
	return (u16) x;
}
/* src/core/proto.h:341-346 */
static u16 rndf_word(float f) {
	f += 0.5f;
	if (f > 65535.f) f = 65535.f;
	if (f < 0.f) f = 0.f;
	return (u16) f;
}
/* src/core/proto.h:208-213 */
static int rnd_int(double x) {
	if (x > (double) INT_MAX - 0.5) x = (double) INT_MAX - 0.5;
	if (x < (double) INT_MIN + 0.5) x = (double) INT_MIN + 0.5;
	return (int) (x + (x >= 0.0 ? 0.5 : -0.5));
}
/* src/core/proto.h:324-328 */
static u16 trunc_word(int x) { return (u16) (x < 0 ? 0 : (x > 65535 ? 65535 : x)); }
/* src/core/proto.h:384-388 */
static float clampf(float v, float lo, float hi) { if (v < lo) v = lo; if (v > hi) v = hi; return v; }

/* ------------------------------------------------------------------ selection / sorting */
static int cmp_f32(const void *a, const void *b) {
	float x = *(const float *) a, y = *(const float *) b; return (x > y) - (x < y);
}
static int cmp_u16(const void *a, const void *b) {
	return (int) *(const u16 *) a - (int) *(const u16 *) b;
}
static int cmp_f64(const void *a, const void *b) {
	double x = *(const double *) a, y = *(const double *) b; return (x > y) - (x < y);
}
/* src/algos/sorting.c:110-135 (quicksort_f) / :161-186 (quicksort_s): plain ascending sort;
 * the sorted result is unique, so the partition scheme does not matter. */
static void sort_f32(float *a, size_t n) { qsort(a, n, sizeof(float), cmp_f32); }
static void sort_u16(u16 *a, size_t n) { qsort(a, n, sizeof(u16), cmp_u16); }

/* src/algos/sorting.c:240-273 quickmedian_float.  n < 9: sorting networks (:468-513) leave the
 * array fully sorted and average the middles IN FLOAT before the /2.0; n >= 9: quickselect with
 * the pivot taken from the middle, moved to the right end, one forward sweep moving smaller
 * elements to the front — the permutation it leaves matters to the callers (order-dependent
 * `N - r <= 4` guard), so it is reproduced step for step. */
double orc_quickmedian_float(float *a, size_t n) {
	if (n == 0) return 0.0;
	size_t k = n / 2;
	if (n < 9) {
		if (n == 1) return a[0];
		sort_f32(a, n);
		return (n % 2 == 0) ? (a[k - 1] + a[k]) / 2.0 : a[k];
	}
	size_t lo = 0, hi = n - 1;
	while (lo < hi) {
		size_t mid = (lo + hi) / 2;
		float pv = a[mid];
		a[mid] = a[hi]; a[hi] = pv;
		size_t store = lo;
		for (size_t i = lo; i < hi; i++) {
			if (a[i] < pv) { float t = a[store]; a[store] = a[i]; a[i] = t; store++; }
		}
		a[hi] = a[store]; a[store] = pv;
		if (store < k) lo = store + 1; else hi = store;
	}
	return (n % 2 == 0) ? ((double) a[k - 1] + a[k]) / 2.0 : (double) a[k];
}
/* src/algos/sorting.c:195-230 quickmedian (WORD) and :366-415 sortnet_median */
double orc_quickmedian(u16 *a, size_t n) {
	if (n == 0) return 0.0;
	size_t k = n / 2;
	if (n < 9) {
		if (n == 1) return a[0];
		sort_u16(a, n);
		return (n % 2 == 0) ? (a[k - 1] + a[k]) / 2.0 : a[k];
	}
	size_t lo = 0, hi = n - 1;
	while (lo < hi) {
		size_t mid = (lo + hi) / 2;
		u16 pv = a[mid];
		a[mid] = a[hi]; a[hi] = pv;
		size_t store = lo;
		for (size_t i = lo; i < hi; i++) {
			if (a[i] < pv) { u16 t = a[store]; a[store] = a[i]; a[i] = t; store++; }
		}
		a[hi] = a[store]; a[store] = pv;
		if (store < k) lo = store + 1; else hi = store;
	}
	return (n % 2 == 0) ? ((double) a[k - 1] + (double) a[k]) / 2.0 : (double) a[k];
}

/* src/algos/sorting.c:577-642 histogram_median: exact median of u16 data by counting */
double orc_histogram_median(const u16 *a, size_t n) {
	if (n < 10) {
		u16 tmp[10];
		memcpy(tmp, a, n * sizeof(u16));
		return orc_quickmedian(tmp, n);	/* n < 9 -> sorted median; n == 9 handled by sortnet too */
	}
	unsigned int *h = calloc(65536, sizeof(unsigned int));
	for (size_t i = 0; i < n; i++) h[a[i]]++;
	unsigned int i = 0, j = 0, k = (unsigned int) (n / 2), sum = 0;
	if (n % 2 == 0) {
		for (; sum <= k - 1; j++) sum += h[j];
		i = j;
	}
	for (; sum <= k; i++) sum += h[i];
	free(h);
	return (n % 2 == 0) ? (double) (i + j - 2) / 2.0 : (double) (i - 1);
}

/* src/rt/rt_algo.cc:38-171 findMinMaxPercentile with minPrct = maxPrct = 0.5 and both outputs
 * aliased (src/algos/sorting.c:644-649 histogram_median_float): an APPROXIMATE median from a
 * min(65536, n)-bin histogram with linear interpolation inside the bin, all in float. */
float orc_histogram_median_float(const float *data, size_t n) {
	if (n == 0) return 0.f;	/* reference leaves `median` uninitialised; callers never pass 0 */
	float mn = data[0], mx = data[0];
	for (size_t i = 1; i < n; i++) { mn = data[i] < mn ? data[i] : mn; mx = mx < data[i] ? data[i] : mx; }
	if (fabsf(mx - mn) == 0.f) return mn;
	const unsigned int hs = n < 65536 ? (unsigned int) n : 65536u;
	const float scale = (hs - 1) / (mx - mn);
	uint32_t *histo = calloc(hs, sizeof(uint32_t));
	for (size_t i = 0; i < n; i++) histo[(uint16_t) (scale * (data[i] - mn))]++;
	size_t k = 0, count = 0;
	const float thresh = 0.5f * n;
	while (count < thresh) count += histo[k++];
	float v;
	if (k > 0) {
		const size_t count_ = count - histo[k - 1];
		const float c0 = count - thresh;
		const float c1 = thresh - count_;
		v = (c1 * k + c0 * (k - 1)) / (c0 + c1);
	} else v = k;
	v /= scale;
	v += mn;
	{ float t = v < mx ? v : mx; v = mn < t ? t : mn; }	/* LIM(v, mn, mx) = max(mn, min(v, mx)) */
	free(histo);
	return v;
}

/* ------------------------------------------------------------------ per-stack statistics */
/* src/algos/statistics.h:82-107 */
float orc_float_sd(const float *d, int N, float *m) {
	double sum = 0.0, vsum = 0.0;
	for (int i = 0; i < N; i++) sum += (double) d[i];
	float mean = (float) (sum / N);
	for (int i = 0; i < N; i++) { float e = d[i] - mean; vsum += (double) (e * e); }
	if (m) *m = mean;
	return sqrtf((float) (vsum / (N - 1)));
}
/* src/algos/statistics.c:115-127 (sd_32; sd_64 :100-112 differs only in the width of the
 * integer sum, identical below 65536 frames) */
static float sd_u16(const u16 *d, int N) {
	uint64_t isum = 0;
	for (int i = 0; i < N; i++) isum += d[i];
	float mean = (float) (((double) isum) / ((double) N));
	double acc = 0.0;
	for (int i = 0; i < N; i++) { float px = (float) d[i]; acc += (px - mean) * (px - mean); }
	return sqrtf((float) (acc / (N - 1)));
}
/* src/stacking/median_and_mean.c:649-662 (the GESDT-only u16 variant: double sum, float mean) */
static float sd_u16_gesdt(const u16 *d, int N, float *m) {
	double acc = 0.0;
	for (int i = 0; i < N; i++) acc += d[i];
	float mean = (float) (acc / N);
	acc = 0.0;
	for (int i = 0; i < N; i++) acc += (d[i] - mean) * (d[i] - mean);
	if (m) *m = mean;
	return sqrtf((float) (acc / (N - 1)));
}
/* src/algos/statistics_float.c:79-101 */
double orc_float_mad(const float *d, size_t n, double m) {
	const float med = (float) m;
	float *tmp = malloc((n ? n : 1) * sizeof(float));
	for (size_t i = 0; i < n; i++) tmp[i] = fabsf(d[i] - med);
	double mad = orc_histogram_median_float(tmp, n);
	free(tmp);
	return mad;
}
/* src/algos/statistics.c:133-154 */
static float mad_u16(const u16 *d, size_t n, double m) {
	int med = rnd_int(m);
	u16 *tmp = malloc((n ? n : 1) * sizeof(u16));
	for (size_t i = 0; i < n; i++) tmp[i] = (u16) abs(d[i] - med);
	float mad = (float) orc_histogram_median(tmp, n);
	free(tmp);
	return mad;
}
/* src/stacking/siril_fit_linear.c:25-48 with x[i] = 1/(i+1) (median_and_mean.c:1495) */
void orc_fit_linear(const float *y, float m_x, float m_dx2, size_t n, float *c0, float *c1) {
	float m_y = y[0];
	for (size_t i = 1; i < n; i++) m_y += (y[i] - m_y) * (1.f / (i + 1));
	float m_dxdy = 0.f, dx = -m_x;
	for (size_t i = 0; i < n; i++, dx += 1.f) {
		const float dy = y[i] - m_y;
		m_dxdy += (dx * dy - m_dxdy) * (1.f / (i + 1));
	}
	const float b = m_dxdy * m_dx2;
	*c0 = m_y - m_x * b;
	*c1 = b;
}
/* src/stacking/median_and_mean.c:1491-1499: constants of the fit, computed for the FULL frame
 * count and reused unchanged when the stack shrinks */
void orc_linfit_consts(int nb_frames, float *m_x, float *m_dx2) {
	float mx = (nb_frames - 1) * 0.5f, acc = 0.f;
	for (int j = 0; j < nb_frames; ++j) {
		const float dx = j - mx;
		acc += (dx * dx - acc) * (1.f / (j + 1));
	}
	*m_x = mx; *m_dx2 = 1.f / acc;
}

/* ------------------------------------------------------------------ rejection, float */
typedef struct {
	int type; float sig[2]; const float *crit; int nb_frames;	/* nb_frames: full stack size */
	float m_x, m_dx2;
} rej_cfg;

/* compacts the entries whose flag is 0; returns the new length */
static int compact_f32(float *s, const int *flag, int n) {
	int o = 0;
	for (int i = 0; i < n; i++) if (!flag[i]) { if (i != o) s[o] = s[i]; o++; }
	return o;
}
static int compact_u16(u16 *s, const int *flag, int n) {
	int o = 0;
	for (int i = 0; i < n; i++) if (!flag[i]) { if (i != o) s[o] = s[i]; o++; }
	return o;
}

/* src/stacking/rejection_float.c:49-60 */
static int sclip_f(float px, float sigma, float lo, float hi, float med, int rej[2]) {
	if (med - px > sigma * lo) { rej[0]++; return -1; }
	else if (px - med > sigma * hi) { rej[1]++; return 1; }
	return 0;
}

/* src/stacking/median_and_mean.c:685-701 */
static void confirm(float *ox, int *oi, int *oo, int n, double median, int *flag, int rej[2]) {
	int i = n - 1;
	while (i > 1 && !oo[i]) i--;
	for (int j = i; j >= 0; j--) {
		oo[j] = 1;
		if (ox[j] >= median) { flag[oi[j]] = 1; rej[1]++; }
		else { flag[oi[j]] = -1; rej[0]++; }
	}
}

/* src/stacking/rejection_float.c:100-354 apply_rejection_float.  `stack` is modified in place
 * exactly as the reference leaves it (callers use that state). */
static int reject_f32(float *stack, int nb, const rej_cfg *c, int *flag, int rej[2], const float *ds) {
	int N = nb, kept = 0, r = 0, first = 1, changed;
	double median = 0.0;
	/* null pixels, and with drizzle the samples of null drizzle weight as well (:116-135) */
	for (int f = 0; f < N; f++) if (stack[f] != 0.f && (!ds || ds[f] != 0.f)) { if (f != kept) stack[kept] = stack[f]; kept++; }
	if (kept <= 1) return kept;								/* :140 */
	const int removed = N - kept;
	N = kept;
	if (c->type == R_PERCENTILE || c->type == R_SIGMA || c->type == R_MAD) {		/* :147-157 */
		median = orc_quickmedian_float(stack, N);
		if (median == 0.0) return 0;
	}
	switch (c->type) {
	case R_PERCENTILE: {									/* :160-173, :31-44 */
		const float med = (float) median;
		for (int f = 0; f < N; f++) {
			if (med - stack[f] > med * c->sig[0]) { rej[0]++; flag[f] = -1; }
			else if (stack[f] - med > med * c->sig[1]) { rej[1]++; flag[f] = 1; }
			else flag[f] = 0;
		}
		N = compact_f32(stack, flag, N);
		break; }
	case R_SIGMA:
	case R_MAD:										/* :174-209 */
		do {
			float var = (c->type == R_SIGMA) ? orc_float_sd(stack, N, NULL)
							 : (float) orc_float_mad(stack, N, median);
			if (!first) median = orc_quickmedian_float(stack, N); else first = 0;
			for (int f = 0; f < N; f++) {
				if (N - r <= 4) flag[f] = 0;
				else { flag[f] = sclip_f(stack[f], var, c->sig[0], c->sig[1], (float) median, rej); if (flag[f]) r++; }
			}
			int o = compact_f32(stack, flag, N);
			changed = N != o; N = o;
		} while (changed && N > 3);
		break;
	case R_SIGMEDIAN: {									/* :210-222 */
		int n;
		do {
			const float sigma = orc_float_sd(stack, N, NULL);
			const float medf = (float) orc_quickmedian_float(stack, N);
			n = 0;
			for (int f = 0; f < N; f++)
				if (sclip_f(stack[f], sigma, c->sig[0], c->sig[1], medf, rej)) { stack[f] = medf; n++; }
		} while (n > 0);
		break; }
	case R_WINSORIZED: {									/* :223-259 */
		float *w = malloc(N * sizeof(float));
		do {
			float sigma0, sigma = orc_float_sd(stack, N, NULL);
			const float medf = (float) orc_quickmedian_float(stack, N);
			memcpy(w, stack, N * sizeof(float));
			do {
				const float m0 = medf - 1.5f * sigma, m1 = medf + 1.5f * sigma;
				for (int j = 0; j < N; j++) { float t = m0 > w[j] ? m0 : w[j]; w[j] = m1 < t ? m1 : t; }
				sigma0 = sigma;
				sigma = 1.134f * orc_float_sd(w, N, NULL);
			} while (fabsf(sigma - sigma0) > sigma0 * 0.0005f);
			for (int f = 0; f < N; f++) {
				if (N - r <= 4) flag[f] = 0;
				else { flag[f] = sclip_f(stack[f], sigma, c->sig[0], c->sig[1], medf, rej); if (flag[f]) r++; }
			}
			int o = compact_f32(stack, flag, N);
			changed = N != o; N = o;
		} while (changed && N > 3);
		free(w);
		break; }
	case R_LINEARFIT:									/* :260-300 */
		do {
			sort_f32(stack, N);
			float a, b;
			orc_fit_linear(stack, c->m_x, c->m_dx2, N, &b, &a);
			float sigma = 0.f;
			for (int f = 0; f < N; f++) sigma += fabsf(stack[f] - (a * f + b));
			sigma /= (float) N;
			for (int f = 0; f < N; f++) {
				if (N - r <= 4) { flag[f] = 0; continue; }
				if (a * f + b - stack[f] > sigma * c->sig[0]) { rej[0]++; flag[f] = -1; r++; }	/* :62-75 */
				else if (stack[f] - a * f - b > sigma * c->sig[1]) { rej[1]++; flag[f] = 1; r++; }
				else flag[f] = 0;
			}
			int o = compact_f32(stack, flag, N);
			changed = N != o; N = o;
		} while (changed && N > 3);
		break;
	case R_GESDT: {										/* :301-348 */
		sort_f32(stack, N);
		median = (N % 2) ? stack[N / 2] : (stack[(N - 1) / 2] + stack[N / 2]) / 2.0;	/* gsl median, float add */
		int max_out = (int) nb * c->sig[0];
		if (removed >= max_out) return kept;
		max_out -= removed;
		float *ox = malloc(max_out * sizeof(float)); int *oi = malloc(max_out * sizeof(int)), *oo = malloc(max_out * sizeof(int));
		float *w = malloc(N * sizeof(float));
		memcpy(w, stack, N * sizeof(float));
		memset(flag, 0, N * sizeof(int));
		int cold = 0;
		for (int it = 0, size = N; it < max_out; it++, size--) {
			float avg, sd = orc_float_sd(w, size, &avg);				/* :82-98 grubbs_stat */
			float dev = avg - w[0], md2 = w[size - 1] - avg;
			int idx = 0;
			if (md2 > dev) { dev = md2; idx = size - 1; }
			float G = dev / sd;
			oo[it] = G > c->crit[it + removed];
			ox[it] = w[idx];
			oi[it] = (idx == 0) ? cold++ : idx;
			for (int i = idx; i < size - 1; i++) w[i] = w[i + 1];			/* remove_element */
		}
		confirm(ox, oi, oo, max_out, median, flag, rej);
		free(ox); free(oi); free(oo); free(w);
		N = compact_f32(stack, flag, N);
		break; }
	default: break;
	}
	return N;
}

/* ------------------------------------------------------------------ rejection, u16 */
/* src/stacking/median_and_mean.c:608-621 */
static int sclip_w(u16 px, float lo, float hi, float sigma, float med, int rej[2]) {
	if (med - px > lo * sigma) { rej[0]++; return -1; }
	else if (px - med > hi * sigma) { rej[1]++; return 1; }
	return 0;
}

/* src/stacking/median_and_mean.c:703-954 apply_rejection_ushort */
static int reject_u16(u16 *stack, int nb, const rej_cfg *c, int *flag, int rej[2], const float *ds) {
	int N = nb, kept = 0, r = 0, first = 1, changed;
	float median = 0.f;
	/* :716-735 (null pixels; with drizzle also the samples of null drizzle weight) */
	for (int f = 0; f < N; f++) if (stack[f] != 0 && (!ds || ds[f] != 0.f)) { if (f != kept) stack[kept] = stack[f]; kept++; }
	if (kept <= 1) return kept;
	const int removed = N - kept;
	N = kept;
	if (c->type >= R_PERCENTILE && c->type <= R_WINSORIZED) {				/* :747-759 */
		median = orc_quickmedian(stack, N);
		if (median == 0.f) return 0;
	}
	switch (c->type) {
	case R_PERCENTILE:									/* :762-776, :590-603 */
		for (int f = 0; f < N; f++) {
			if ((median - (float) stack[f]) / median > c->sig[0]) { rej[0]++; flag[f] = -1; }
			else if (((float) stack[f] - median) / median > c->sig[1]) { rej[1]++; flag[f] = 1; }
			else flag[f] = 0;
		}
		N = compact_u16(stack, flag, N);
		break;
	case R_SIGMA:
	case R_MAD:										/* :777-813 */
		do {
			float var = (c->type == R_SIGMA) ? sd_u16(stack, N) : mad_u16(stack, N, median);
			if (!first) median = orc_quickmedian(stack, N); else first = 0;
			for (int f = 0; f < N; f++) {
				if (N - r <= 4) flag[f] = 0;
				else { flag[f] = sclip_w(stack[f], c->sig[0], c->sig[1], var, median, rej); if (flag[f]) r++; }
			}
			int o = compact_u16(stack, flag, N);
			changed = N != o; N = o;
		} while (changed && N > 3);
		break;
	case R_SIGMEDIAN: {									/* :814-830 */
		int n;
		do {
			const float sigma = sd_u16(stack, N);
			if (!first) median = orc_quickmedian(stack, N); else first = 0;
			n = 0;
			for (int f = 0; f < N; f++)
				if (sclip_w(stack[f], c->sig[0], c->sig[1], sigma, median, rej)) { stack[f] = median; n++; }
		} while (n > 0);
		break; }
	case R_WINSORIZED: {									/* :831-867 */
		u16 *w = malloc(N * sizeof(u16));
		do {
			float sigma0, sigma = sd_u16(stack, N);
			if (!first) median = orc_quickmedian(stack, N); else first = 0;
			memcpy(w, stack, N * sizeof(u16));
			do {
				const u16 m0 = rndf_word(median - 1.5f * sigma), m1 = rndf_word(median + 1.5f * sigma);
				for (int j = 0; j < N; j++) { w[j] = w[j] < m0 ? m0 : w[j]; w[j] = w[j] > m1 ? m1 : w[j]; }	/* :623-628 */
				sigma0 = sigma;
				sigma = 1.134f * sd_u16(w, N);
			} while (fabs(sigma - sigma0) > sigma0 * 0.0005f);
			for (int f = 0; f < N; f++) {
				if (N - r <= 4) flag[f] = 0;
				else { flag[f] = sclip_w(stack[f], c->sig[0], c->sig[1], sigma, median, rej); if (flag[f]) r++; }
			}
			int o = compact_u16(stack, flag, N);
			changed = N != o; N = o;
		} while (changed && N > 3);
		free(w);
		break; }
	case R_LINEARFIT: {									/* :868-902 */
		float *yf = malloc(N * sizeof(float));
		do {
			sort_u16(stack, N);
			for (int f = 0; f < N; f++) yf[f] = (float) stack[f];
			float a, b;
			orc_fit_linear(yf, c->m_x, c->m_dx2, N, &b, &a);
			float sigma = 0.f;
			for (int f = 0; f < N; f++) sigma += fabsf(stack[f] - (a * f + b));
			sigma /= (float) N;
			for (int f = 0; f < N; f++) {
				if (N - r <= 4) { flag[f] = 0; continue; }
				if (a * f + b - stack[f] > sigma * c->sig[0]) { rej[0]++; flag[f] = -1; r++; }	/* :630-642 */
				else if (stack[f] - a * f - b > sigma * c->sig[1]) { rej[1]++; flag[f] = 1; r++; }
				else flag[f] = 0;
			}
			int o = compact_u16(stack, flag, N);
			changed = N != o; N = o;
		} while (changed && N > 3);
		free(yf);
		break; }
	case R_GESDT: {										/* :903-948 */
		sort_u16(stack, N);
		median = (N % 2) ? stack[N / 2] : (stack[(N - 1) / 2] + stack[N / 2]) / 2.0;	/* -> float */
		int max_out = (int) nb * c->sig[0];
		if (removed >= max_out) return kept;
		max_out -= removed;
		float *ox = malloc(max_out * sizeof(float)); int *oi = malloc(max_out * sizeof(int)), *oo = malloc(max_out * sizeof(int));
		u16 *w = malloc(N * sizeof(u16));
		memcpy(w, stack, N * sizeof(u16));
		memset(flag, 0, N * sizeof(int));
		int cold = 0;
		for (int it = 0, size = N; it < max_out; it++, size--) {
			float avg, sd = sd_u16_gesdt(w, size, &avg);				/* :664-679 */
			float dev = avg - w[0], md2 = w[size - 1] - avg;
			int idx = 0;
			if (md2 > dev) { dev = md2; idx = size - 1; }
			float G = dev / sd;
			oo[it] = G > c->crit[it + removed];
			ox[it] = w[idx];
			oi[it] = (idx == 0) ? cold++ : idx;
			for (int i = idx; i < size - 1; i++) w[i] = w[i + 1];
		}
		confirm(ox, oi, oo, max_out, median, flag, rej);
		free(ox); free(oi); free(oo); free(w);
		N = compact_u16(stack, flag, N);
		break; }
	default: break;
	}
	return N;
}

/* ------------------------------------------------------------------ combine */
/* the weight of one sample in the weighted combine: n = 1 * drizzle * mask * frame weight, in this
 * order (median_and_mean.c:998-1004, :1060-1066); ms / ds: the pixel's gathered mask / drizzle stacks */
static double sample_weight(const double *w, const float *ms, const float *ds, int f) {
	double n = 1.;
	if (ds) n *= ds[f];
	if (ms) n *= ms[f];
	if (w) n *= w[f];
	return n;
}
/* src/stacking/median_and_mean.c:956-1101 mean_and_reject.  `orig` is the untouched copy of the
 * stack (o_stack, rejection_float.c:114); weights == NULL means NO_WEIGHT. */
static double combine_f32(float *stack, const float *orig, int nb, const rej_cfg *c,
		const double *w, int *flag, int rej[2], const float *ms, const float *ds) {
	int kept = reject_f32(stack, nb, c, flag, rej, ds);
	if (kept == 0) return orc_quickmedian_float(stack, nb);					/* :1040-1041 */
	if (w || ms || ds) {									/* :1043-1082 */
		float pmin = FLT_MAX, pmax = -FLT_MAX;
		for (int f = 0; f < kept; f++) { if (pmin > stack[f]) pmin = stack[f]; if (pmax < stack[f]) pmax = stack[f]; }
		double sum = 0.0, norm = 0.0;
		for (int f = 0; f < nb; f++) {
			float v = orig[f];
			if (v >= pmin && v <= pmax && v != 0.f) { const double n = sample_weight(w, ms, ds, f); sum += (double) v * n; norm += n; }
		}
		if (norm == 0. || sum == 0.) {
			sum = 0.;
			for (int f = 0; f < nb; f++) { float v = orig[f]; if (v >= pmin && v <= pmax && v > 0) sum += (double) v; }
			return sum / (double) kept;
		}
		return sum / norm;
	}
	double sum = 0.0;									/* :1084-1096 */
	for (int f = 0; f < kept; f++) sum += (double) stack[f];
	return sum / (double) kept;
}
static double combine_u16(u16 *stack, const u16 *orig, int nb, const rej_cfg *c,
		const double *w, int *flag, int rej[2], const float *ms, const float *ds) {
	int kept = reject_u16(stack, nb, c, flag, rej, ds);
	if (kept == 0) return orc_quickmedian(stack, nb);					/* :964-965 */
	if (w || ms || ds) {									/* :967-1021 */
		u16 pmin = 65535, pmax = 0;
		for (int f = 0; f < kept; f++) { if (pmin > stack[f]) pmin = stack[f]; if (pmax < stack[f]) pmax = stack[f]; }
		double sum = 0.0, norm = 0.0;
		for (int f = 0; f < nb; f++) {
			u16 v = orig[f];
			if (v >= pmin && v <= pmax && v > 0) { const double n = sample_weight(w, ms, ds, f); sum += (double) v * n; norm += n; }
		}
		if (norm == 0. || sum == 0.) {
			sum = 0.;
			for (int f = 0; f < nb; f++) { u16 v = orig[f]; if (v >= pmin && v <= pmax && v > 0) sum += (double) v; }
			return sum / (double) kept;
		}
		return sum / norm;
	}
	int64_t sum = 0;									/* :1023-1035 */
	for (int f = 0; f < kept; f++) sum += stack[f];
	return sum / (double) kept;
}

/* one pixel stack -> (mean, rej[2]); same signature as ref_mean_and_reject in ref_driver.c */
double orc_mean_and_reject(int is_float, const void *stack, int N, int rejtype, float s0, float s1,
		const float *crit, const double *weights, int layer, int rej[2]) {
	rej_cfg c = { rejtype, { s0, s1 }, crit, N, 0.f, 0.f };
	if (rejtype == R_LINEARFIT) orc_linfit_consts(N, &c.m_x, &c.m_dx2);
	const double *w = weights ? weights + (size_t) layer * N : NULL;
	int *flag = calloc(N + 1, sizeof(int));
	size_t es = is_float ? 4 : 2;
	void *s = malloc((N + 1) * es), *o = malloc((N + 1) * es);
	memcpy(s, stack, N * es); memcpy(o, stack, N * es);
	rej[0] = rej[1] = 0;
	double m = is_float ? combine_f32(s, o, N, &c, w, flag, rej, NULL, NULL) : combine_u16(s, o, N, &c, w, flag, rej, NULL, NULL);
	free(s); free(o); free(flag);
	return m;
}

int orc_apply_rejection_float(float *stack, int N, int rejtype, float s0, float s1, const float *crit, int rej[2]) {
	rej_cfg c = { rejtype, { s0, s1 }, crit, N, 0.f, 0.f };
	if (rejtype == R_LINEARFIT) orc_linfit_consts(N, &c.m_x, &c.m_dx2);
	int *flag = calloc(N + 1, sizeof(int));
	rej[0] = rej[1] = 0;
	int kept = reject_f32(stack, N, &c, flag, rej, NULL);
	free(flag);
	return kept;
}

/* ------------------------------------------------------------------ whole-stack driver */
typedef struct {
	int is_float, N, C, H, W;
	int is_mean, rejtype;
	float sig[2];
	const float *crit;
	int normalize;
	const double *scale, *offset, *mul;	/* [C*N] layer-major (coeff.pscale[layer][frame]) */
	const double *weights;			/* [C*N] or NULL */
	const int *shiftx, *shifty;		/* [N] or NULL */
	int output_norm, use_32bit_output;
	int create_rejmaps, merge_rejmaps;
	int nthreads;
	long max_rows;				/* unused by the restatement (kept for ABI parity with ref_stack) */
	int source_bitpix;			/* seq->bitpix; 8 (BYTE_IMG): normalize_to16bit */
	const float *mask;			/* [N][H][W] up-scaled distance masks (data->mask[frame] before :513-521) or NULL */
	const float *drizz;			/* [N][C][H][W] drizzle weights (data->drizz[frame]) or NULL */
	int feather_dist, drizzle;		/* args->feather_dist (> 0: masking), args->drizzle */
} orc_stack_params;

/* the mask normalization of stack_read_block_data (median_and_mean.c:513-521) with the ramp table of
 * blending.c:34-50 (r^3 (6 r^2 - 15 r + 10) at 1001 points, looked up by truncation) */
static float orc_mask_ramp(float v, float distf, float invdistf) {
	if (!v) return v;
	if (v > distf) return 1.f;
	const int index = (int) ((v * invdistf) * 1000);
	const float norm = 1.f / (float) 1000;
	const float r = (float) index * norm;
	return r * r * r * (6.f * r * r - 15.f * r + 10.f);
}

/* src/stacking/median_and_mean.c:557-582; NB the min/max scan starts at index 1 (:566) */
void orc_norm_to_0_1(float *d, long W, long H, long C) {
	long n = W * H * C;
	float mini = FLT_MAX, maxi = -FLT_MAX;
	for (long i = 1; i < n; i++) {
		float t = d[i];
		if (t == 0.f) continue;
		if (t < mini) mini = t;
		if (t > maxi) maxi = t;
	}
	for (long i = 0; i < n; i++) d[i] = (d[i] == 0.f) ? 0.f : (d[i] - mini) / (maxi - mini);
}

/* Per-pixel gather + normalization + combine + output conversion:
 * src/stacking/median_and_mean.c:1612-1736; frames [N][C][H][W] and out [C][H][W] in memory
 * (bottom-up FITS) row order; out row r reads frame row r - shifty (read-time y shift,
 * :407-448) and column x - shiftx (:1617-1636); outside the frame the sample is 0. */
int orc_stack(const orc_stack_params *p, const void *frames, void *out,
		u16 *rej_low, u16 *rej_high, uint64_t irej[6]) {
	const int N = p->N, C = p->C;
	const long H = p->H, W = p->W;
	const size_t plane = (size_t) H * W;
	rej_cfg c = { p->rejtype, { p->sig[0], p->sig[1] }, p->crit, N, 0.f, 0.f };
	if (p->rejtype == R_LINEARFIT) orc_linfit_consts(N, &c.m_x, &c.m_dx2);
	uint64_t tot[6] = { 0, 0, 0, 0, 0, 0 };
	int nthreads = p->nthreads > 0 ? p->nthreads : 1;
#pragma omp parallel num_threads(nthreads)
	{
		size_t es = p->is_float ? 4 : 2;
		void *stack = malloc((N + 1) * es), *orig = malloc((N + 1) * es);
		int *flag = calloc(N + 1, sizeof(int));
		const int masking = p->is_mean && p->feather_dist > 0 && p->mask, drizzle = p->is_mean && p->drizzle && p->drizz;
		float *ms = masking ? calloc(N + 1, sizeof(float)) : NULL, *ds = drizzle ? calloc(N + 1, sizeof(float)) : NULL;
		const float distf = (float) p->feather_dist, invdistf = 1.f / distf;
		uint64_t loc[6] = { 0, 0, 0, 0, 0, 0 };
#pragma omp for schedule(dynamic, 4) collapse(2)
		for (int layer = 0; layer < C; layer++) for (long r = 0; r < H; r++) {
			for (long x = 0; x < W; x++) {
				for (int f = 0; f < N; f++) {
					long sx = x, sr = r; int inside = 1;
					if (p->shiftx) { sx = x - p->shiftx[f]; if (sx < 0 || sx >= W) inside = 0; }
					if (p->shifty) { sr = r - p->shifty[f]; if (sr < 0 || sr >= H) inside = 0; }
					u16 px = 0; float fpx = 0.f;
					if (inside) {
						size_t idx = ((size_t) f * C + layer) * plane + (size_t) sr * W + sx;
						if (p->is_float) fpx = ((const float *) frames)[idx]; else px = ((const u16 *) frames)[idx];
						/* :1687-1692: the mask / drizzle samples are gathered at the same (shifted) place; outside the
						 * frame the reference leaves stale values, which nothing reads (the sample is null) */
						if (ms) ms[f] = orc_mask_ramp(p->mask[(size_t) f * plane + (size_t) sr * W + sx], distf, invdistf);
						if (ds) ds[f] = p->drizz[idx];
					}
					const double sc = p->scale ? p->scale[layer * N + f] : 1.0;
					double tmp;
					switch (p->normalize) {
					default: case NRM_NONE:							/* :1646-1653 */
						if (p->is_float) ((float *) stack)[f] = fpx; else ((u16 *) stack)[f] = px;
						break;
					case NRM_ADD: case NRM_ADDSCALE:					/* :1654-1673 */
						if (p->is_float) {
							if (fpx != 0.f) { tmp = fpx * sc; ((float *) stack)[f] = (float) (tmp - p->offset[layer * N + f]); }
							else ((float *) stack)[f] = 0.f;
						} else {
							if (px > 0) { tmp = (double) px * sc; ((u16 *) stack)[f] = rnd_word(tmp - p->offset[layer * N + f]); }
							else ((u16 *) stack)[f] = 0;
						}
						break;
					case NRM_MUL: case NRM_MULSCALE:					/* :1674-1685 */
						if (p->is_float) { tmp = fpx * sc; ((float *) stack)[f] = (float) (tmp * p->mul[layer * N + f]); }
						else { tmp = (double) px * sc; ((u16 *) stack)[f] = rnd_word(tmp * p->mul[layer * N + f]); }
						break;
					}
				}
				double result;
				const size_t o = (size_t) layer * plane + (size_t) r * W + x;
				if (p->is_mean) {								/* :1696-1710 */
					int rej[2] = { 0, 0 };
					memcpy(orig, stack, N * es);
					const double *w = p->weights ? p->weights + (size_t) layer * N : NULL;
					result = p->is_float ? combine_f32(stack, orig, N, &c, w, flag, rej, ms, ds)
							     : combine_u16(stack, orig, N, &c, w, flag, rej, ms, ds);
					loc[layer * 2] += rej[0]; loc[layer * 2 + 1] += rej[1];
					if (p->create_rejmaps) {
						if (p->merge_rejmaps) rej_low[o] = trunc_word(rej[0] + rej[1]);
						else { rej_low[o] = trunc_word(rej[0]); rej_high[o] = trunc_word(rej[1]); }
					}
				} else {									/* :1711-1715 */
					result = p->is_float ? orc_quickmedian_float(stack, N) : orc_quickmedian(stack, N);
				}
				if (p->use_32bit_output) {							/* :1717-1727 */
					float v = p->is_float ? (float) result : (float) result * .000015259022f;
					if (!p->output_norm) v = clampf(v, 0.f, 1.f);
					((float *) out)[o] = v;
				} else {
					if (p->output_norm && p->source_bitpix == 8) result *= (65535.0 / 255.0);	/* normalize_to16bit :547-555, :1729-1732 */
					((u16 *) out)[o] = rnd_word(result);					/* :1733 */
				}
			}
		}
#pragma omp critical
		for (int i = 0; i < 6; i++) tot[i] += loc[i];
		free(stack); free(orig); free(flag); free(ms); free(ds);
	}
	for (int i = 0; i < 6; i++) irej[i] = (p->is_mean && p->rejtype != R_NONE) ? tot[i] : 0;	/* :1748 */
	if (p->use_32bit_output && p->output_norm) orc_norm_to_0_1(out, W, H, C);			/* :1774-1775 */
	return 0;
}

/* ------------------------------------------------------------------ per-image statistics */
/* src/algos/quantize.c:376-422 / :327-373 mean & rms of first differences */
static void diff_mean_sigma(const double *sum_, const double *sum2_, long n, double *mean, double *sigma) {
	double sum = *sum_, sum2 = *sum2_;
	if (n > 1) { double m = sum / n; *mean = m; *sigma = sqrt((sum2 / n) - (m * m)); }
	else if (n == 1) { *mean = sum; *sigma = 0.0; }
	else { *mean = 0.; *sigma = 0.; }
}

/* src/algos/quantize.c:1343-1488 FnNoise1_float / :1202-1341 FnNoise1_ushort: per row, sigma of
 * the first differences of consecutive valid pixels after <= 3 passes of 5-sigma clipping;
 * result 0.70710678 x median over rows. */
static double bgnoise_plane(int is_float, const void *plane, long nx, long ny) {
	if (nx < 3) return 0.0;
	double *rows = calloc(ny, sizeof(double));
	long nrows = 0;
	double *diff = malloc(nx * sizeof(double));	/* differences are exactly representable either way */
	float *fdiff = malloc(nx * sizeof(float));
	for (long jj = 0; jj < ny; jj++) {
		long nvals = 0, ii = 0;
		if (is_float) {
			const float *row = (const float *) plane + jj * nx;
			while (ii < nx && (row[ii] == 0.f || isnan(row[ii]))) ii++;
			if (ii == nx) continue;
			float v1 = row[ii];
			for (ii++; ii < nx; ii++) {
				while (ii < nx && (row[ii] == 0.f || isnan(row[ii]))) ii++;
				if (ii == nx) break;
				fdiff[nvals++] = v1 - row[ii];
				v1 = row[ii];
			}
			for (long i = 0; i < nvals; i++) diff[i] = (double) fdiff[i];
		} else {
			const u16 *row = (const u16 *) plane + jj * nx;
			while (ii < nx && row[ii] == 0) ii++;
			if (ii == nx) continue;
			u16 v1 = row[ii];
			for (ii++; ii < nx; ii++) {
				while (ii < nx && row[ii] == 0) ii++;
				if (ii == nx) break;
				diff[nvals++] = (double) ((int) v1 - (int) row[ii]);
				v1 = row[ii];
			}
		}
		if (nvals < 2) continue;
		double mean, stdev, s = 0., s2 = 0.;
		for (long i = 0; i < nvals; i++) { s += diff[i]; s2 += diff[i] * diff[i]; }
		diff_mean_sigma(&s, &s2, nvals, &mean, &stdev);
		if (stdev > 0.) {
			for (int iter = 0; iter < 3; iter++) {
				long kk = 0;
				for (long i = 0; i < nvals; i++) {
					/* float: fabs(d - (float)mean) with d float -> float subtraction (:1447);
					 * u16: fabs(d - mean) in double (:1303) */
					double dev = is_float ? fabs((double) ((float) diff[i] - (float) mean)) : fabs(diff[i] - mean);
					if (dev < 5. * stdev) { if (kk < i) diff[kk] = diff[i]; kk++; }
				}
				if (kk == nvals) break;
				nvals = kk;
				s = 0.; s2 = 0.;
				for (long i = 0; i < nvals; i++) { s += diff[i]; s2 += diff[i] * diff[i]; }
				diff_mean_sigma(&s, &s2, nvals, &mean, &stdev);
			}
		}
		rows[nrows++] = stdev;
	}
	double xnoise;
	if (nrows == 0) xnoise = 0;
	else if (nrows == 1) xnoise = rows[0];
	else { qsort(rows, nrows, sizeof(double), cmp_f64); xnoise = (rows[(nrows - 1) / 2] + rows[nrows / 2]) / 2.; }
	free(rows); free(diff); free(fdiff);
	return .70710678 * xnoise;
}

/* src/algos/statistics_float.c:103-127 */
static double bwmv_f32(const float *d, size_t n, float mad, float median) {
	double up = 0.0, down = 0.0;
	if (!(mad > 0.f)) return 0.0;
	const float factor = 1.f / (9.f * mad);
	for (size_t i = 0; i < n; i++) {
		const float im = d[i] - median;
		const float yi = im * factor;
		const float yi2 = fabsf(yi) < 1.f ? yi * yi : 1.f;
		const float t = im * ((1 - yi2) * (1 - yi2));
		up += t * t;
		down += (1 - yi2) * (1 - 5 * yi2);
	}
	return down ? n * (up / (down * down)) : 0.0;
}

/* src/algos/statistics_float.c:199-229 IKSSlite (data is filtered in place) */
static int ikss_lite(float *d, size_t n, float median, float mad, double *location, double *scale) {
	float xlow = median - 6.0 * mad, xhigh = median + 6.0 * mad;
	size_t kept = 0;
	for (size_t i = 0; i < n; i++) if (d[i] >= xlow && d[i] <= xhigh) { if (i != kept) d[kept] = d[i]; kept++; }
	if (kept == 0) return 1;
	*location = orc_histogram_median_float(d, kept);
	mad = orc_float_mad(d, kept, *location);
	if (mad == 0.0f) return 1;
	*scale = sqrt(bwmv_f32(d, kept, mad, *location)) * .991;
	return 0;
}

/* out[14] in imstats order (core/siril.h:407-413): total, ngoodpix, mean, median, sigma, avgDev,
 * mad, sqrtbwmv, location, scale, min, max, normValue, bgnoise.  Sequencing:
 * src/algos/statistics_float.c:362-477 and src/algos/statistics.c:311-438 with
 * option = STATS_NORM (lite: STATS_LITENORM), statistics.h:35-36. */
int orc_image_stats(int is_float, const void *plane, long nx, long ny, int lite, double out[14]) {
	const size_t total = (size_t) nx * ny;
	for (int i = 0; i < 14; i++) out[i] = 0;
	long ngood = 0;
	double sum = 0., sum2 = 0., mean, sigma, median, mad, location = 0, scale = 0, minv, maxv, normValue;
	if (is_float) {
		const float *d = plane;
		normValue = 1.0;
		float mn = d[0], mx = d[0];								/* :254-275 */
		for (size_t i = 0; i < total; i++) if (!isnan(d[i])) { mn = d[i] < mn ? d[i] : mn; mx = d[i] > mx ? d[i] : mx; }
		minv = mn; maxv = mx;
		for (size_t i = 0; i < total; i++) if (d[i] != 0.f && !isnan(d[i])) {		/* quantize.c:264-317 */
			ngood++; double x = d[i]; sum += x; sum2 += x * x;
		}
	} else {
		const u16 *d = plane;
		normValue = 65535.0;
		u16 mn = d[0], mx = d[0];
		for (size_t i = 0; i < total; i++) { mn = d[i] < mn ? d[i] : mn; mx = d[i] > mx ? d[i] : mx; }
		minv = mn; maxv = mx;
		for (size_t i = 0; i < total; i++) if (d[i] != 0) { ngood++; double x = d[i]; sum += x; sum2 += x * x; }	/* quantize.c:207-261 */
	}
	if (ngood > 1) { mean = sum / ngood; sigma = sqrt((sum2 / ngood) - (mean * mean)); }
	else if (ngood == 1) { mean = sum; sigma = 0.0; }
	else { mean = 0.; sigma = 0.; }
	double bgnoise = bgnoise_plane(is_float, plane, nx, ny);
	if (ngood == 0) return 1;
	if (is_float) {
		const float *d = plane;
		float *nz = malloc(ngood * sizeof(float));						/* :231-252 */
		size_t j = 0;
		for (size_t i = 0; i < total; i++) if (d[i] != 0.f && !isnan(d[i])) nz[j++] = d[i];
		median = orc_histogram_median_float(nz, ngood);
		mad = orc_float_mad(nz, ngood, median);
		if (!lite && ikss_lite(nz, ngood, median, mad, &location, &scale)) { free(nz); return 2; }
		free(nz);
	} else {
		const u16 *d = plane;
		u16 *nz = malloc(ngood * sizeof(u16));
		size_t j = 0;
		for (size_t i = 0; i < total; i++) if (d[i] != 0) nz[j++] = d[i];
		median = orc_histogram_median(nz, ngood);
		mad = mad_u16(nz, ngood, median);
		if (!lite) {											/* statistics.c:409-438 */
			float *nd = malloc(ngood * sizeof(float));
			float inv = (float) (1.0 / normValue);
			for (long i = 0; i < ngood; i++) nd[i] = (float) nz[i] * inv;
			float med = (float) median * inv, madf = (float) mad * inv;
			int rv = ikss_lite(nd, ngood, med, madf, &location, &scale);
			free(nd);
			if (rv) { free(nz); return 2; }
			location *= normValue; scale *= normValue;
		}
		free(nz);
	}
	out[0] = (double) total; out[1] = (double) ngood; out[2] = mean; out[3] = median; out[4] = sigma;
	out[6] = mad; out[8] = location; out[9] = scale; out[10] = minv; out[11] = maxv;
	out[12] = normValue; out[13] = bgnoise;
	return 0;
}

/* src/stacking/normalization.c:144-181: arrays [C*N] hold the estimators on entry, factors on exit */
void orc_compute_factors(int normalize, int nb_layers, int N, int ref_index, int reglayer,
		int equalizeRGB, double *offset, double *mul, double *scale) {
	double o0[3], m0[3], s0[3];
	for (int l = 0; l < nb_layers; l++) { o0[l] = offset[l * N + ref_index]; m0[l] = mul[l * N + ref_index]; s0[l] = scale[l * N + ref_index]; }
	int rl = (reglayer > -1) ? reglayer : 1;
	for (int l = 0; l < nb_layers; l++) {
		int ref = equalizeRGB ? rl : l;
		for (int i = 0; i < N; i++) {
			double *ps = &scale[l * N + i], *po = &offset[l * N + i], *pm = &mul[l * N + i];
			switch (normalize) {
			default:
			case NRM_ADDSCALE: *ps = (*ps == 0) ? 1 : s0[ref] / *ps;	/* fallthrough */
			case NRM_ADD: *po = *ps * *po - o0[ref]; break;
			case NRM_MULSCALE: *ps = (*ps == 0) ? 1 : s0[ref] / *ps;	/* fallthrough */
			case NRM_MUL: *pm = (*pm == 0) ? 1 : m0[ref] / *pm; break;
			}
		}
	}
}

/* src/stacking/median_and_mean.c:1111-1135 */
void orc_noise_weights(int nb_layers, int N, const double *scale, const double *bgnoise, double *weights) {
	for (int l = 0; l < nb_layers; l++) {
		double norm = 0.0;
		for (int i = 0; i < N; i++) {
			double s = scale[l * N + i], b = bgnoise[l * N + i];
			weights[l * N + i] = 1.f / (s * s * b * b);
			norm += weights[l * N + i];
		}
		norm /= (double) N;
		for (int i = 0; i < N; i++) weights[l * N + i] /= norm;
	}
}

/* src/stacking/median_and_mean.c:261-356 block planner (refine_blocks_candidate +
 * stack_compute_parallel_blocks); blocks_out: 4 longs per block */
static int ceil_mult(int x, int f) { int r = x % f; return r ? x + f - r : x; }	/* proto.h:295-299 */
int orc_planner(long max_rows, long W, long H, long C, int nb_threads, long *blocks_out, int max_blocks,
		long *largest, int *nb_blocks) {
	(void) W;
	if (nb_threads < 1 || max_rows < 1) return -1;
	int cand = nb_threads;
	while ((max_rows * cand) / nb_threads < H * C) cand++;
	int nch = (C == 3) ? 3 : 1, factor = nch;
	if (nb_threads < 4) {
		if (factor != 1 && nb_threads % factor == 0) factor = nb_threads; else factor *= nb_threads;
		cand = ceil_mult(cand, factor);
	} else {
		int slack = nb_threads < 8 ? 1 : 3;
		cand = ceil_mult(cand, factor);
		for (;;) { int rem = cand % nb_threads; if (rem == 0 || rem >= nb_threads - slack) break; cand += factor; }
	}
	*nb_blocks = cand;
	long hb = H * C / cand;
	int remainder = H % (cand / C);
	*largest = 0;
	long ch = 0, row = 0, j = 0;
	do {
		if (j >= cand) return -1;
		long end = row + hb - 1;
		if (remainder > 0) { end++; remainder--; }
		long start = row, channel = ch;
		if (end >= H - 1 || (H - end < hb / 10)) {
			end = H - 1; row = 0; ch++;
			remainder = H - (cand / C * hb);
		} else row = end + 1;
		long h = end - start + 1;
		if (j < max_blocks) { blocks_out[4 * j] = channel; blocks_out[4 * j + 1] = start; blocks_out[4 * j + 2] = end; blocks_out[4 * j + 3] = h; }
		if (*largest < h) *largest = h;
		j++;
	} while (ch < C);
	return 0;
}

/* ---- sum / min / max stackers (src/stacking/sum.c, src/stacking/stackminmax.c) ------------------
 * Frames [N][C][H][W], output [C][H][W]; shiftx/shifty are the integer registration shifts the
 * reference rounds from the homography (sum.c:160-173): output pixel (x, y) takes source pixel
 * (x - shiftx, y - shifty) of each frame when that lies inside the frame (sum.c:186-193,
 * stackminmax.c:160-167). */

/* sum.c:193-205 (accumulate: double for float input, 64-bit integer for WORD), :256-278 (maximum
 * over all layers; an all-zero sum is an error), :307-343 (output: float = sum / max; WORD = sum,
 * or sum * 65535 / max when the maximum does not fit) */
int orc_stack_sum(int is_float, const void *frames, int N, int C, int H, int W,
		const int *shiftx, const int *shifty, int output_32bits, void *out) {
	const size_t np = (size_t) H * W, nb = np * C;
	double *fsum = is_float ? calloc(nb, sizeof(double)) : NULL;
	uint64_t *sum = is_float ? NULL : calloc(nb, sizeof(uint64_t));
	for (int f = 0; f < N; f++) {
		const int sx = shiftx ? shiftx[f] : 0, sy = shifty ? shifty[f] : 0;
		for (int y = 0; y < H; y++) {
			const int ny = y - sy;
			if (ny < 0 || ny >= H) continue;
			for (int x = 0; x < W; x++) {
				const int nx = x - sx;
				if (nx < 0 || nx >= W) continue;
				for (int l = 0; l < C; l++) {
					const size_t src = ((size_t) f * C + l) * np + (size_t) ny * W + nx, dst = l * np + (size_t) y * W + x;
					if (is_float) fsum[dst] += (double) ((const float *) frames)[src];
					else sum[dst] += ((const u16 *) frames)[src];
				}
			}
		}
	}
	int rv = 0;
	if (is_float) {
		double fmax = 0.0;
		for (size_t i = 0; i < nb; i++) if (fsum[i] > fmax) fmax = fsum[i];
		if (!fmax) rv = 1;
		else if (output_32bits) {
			const double ratio = 1.0 / fmax;
			for (size_t i = 0; i < nb; i++) ((float *) out)[i] = (float) (fsum[i] * ratio);
		} else rv = 1;	/* sum.c has no float -> WORD branch (a float sequence always stacks to 32 bits) */
	} else {
		uint64_t max = 0;
		for (size_t i = 0; i < nb; i++) if (sum[i] > max) max = sum[i];
		if (!max) rv = 1;
		else if (output_32bits) {
			const double ratio = 1.0 / (double) max;
			for (size_t i = 0; i < nb; i++) ((float *) out)[i] = (float) ((double) sum[i] * ratio);
		} else {
			double ratio = 1.0;
			if (max > 65535) ratio = 65535.0 / (double) max;
			for (size_t i = 0; i < nb; i++)
				((u16 *) out)[i] = ratio == 1.0 ? rnd_word((double) sum[i]) : rnd_word((double) sum[i] * ratio);
		}
	}
	free(fsum); free(sum);
	return rv;
}

/* stackminmax.c:49-84 (start from 0 / 1.0f / USHRT_MAX), :168-181 (strict comparisons) */
int orc_stack_minmax(int is_float, const void *frames, int N, int C, int H, int W,
		const int *shiftx, const int *shifty, int ismax, void *out) {
	const size_t np = (size_t) H * W, nb = np * C;
	if (is_float) for (size_t i = 0; i < nb; i++) ((float *) out)[i] = ismax ? 0.0f : 1.0f;
	else for (size_t i = 0; i < nb; i++) ((u16 *) out)[i] = ismax ? 0 : 65535;
	for (int f = 0; f < N; f++) {
		const int sx = shiftx ? shiftx[f] : 0, sy = shifty ? shifty[f] : 0;
		for (int y = 0; y < H; y++) {
			const int ny = y - sy;
			if (ny < 0 || ny >= H) continue;
			for (int x = 0; x < W; x++) {
				const int nx = x - sx;
				if (nx < 0 || nx >= W) continue;
				for (int l = 0; l < C; l++) {
					const size_t src = ((size_t) f * C + l) * np + (size_t) ny * W + nx, dst = l * np + (size_t) y * W + x;
					if (is_float) {
						const float cur = ((const float *) frames)[src], acc = ((float *) out)[dst];
						if ((ismax && cur > acc) || (!ismax && cur < acc)) ((float *) out)[dst] = cur;
					} else {
						const u16 cur = ((const u16 *) frames)[src], acc = ((u16 *) out)[dst];
						if ((ismax && cur > acc) || (!ismax && cur < acc)) ((u16 *) out)[dst] = cur;
					}
				}
			}
		}
	}
	return 0;
}
