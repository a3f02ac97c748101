// sorted.cuh — rejection and combine on a pixel stack that is ALREADY SORTED ascending (zeros
// removed in mean mode).  This is "phase B" of the sorted engine (stack_sorted.cu): phase A sorts
// every pixel stack cooperatively in registers (bitonic network across the lanes of a warp) and
// parks it in shared memory; phase B then runs one thread per pixel over its sorted stack.
//
// What the order buys.  With the stack sorted, every order statistic the reference needs is a
// lookup: the median (quickmedian_float, sorting.c:240-273) is the middle element(s); the kept
// set of percentile / sigma / winsorized clipping is a contiguous rank range [a, b) because the
// reference's float predicates (rejection_float.c:31-60) are monotone in the sample value, so a
// clip pass is a walk inwards from both ends; linear-fit clipping and GESDT sort the stack
// themselves (rejection_float.c:262, :311).
//
// How identical decisions are kept without the reference's per-element passes.
//  - clip predicates are evaluated with the reference's own float operations on the boundary
//    elements, so they are exact given the sigma they are fed;
//  - the standard deviation (siril_stats_float_sd, statistics.h:82-107: double sum, float mean,
//    float deviations squared in float, double accumulation, float sqrt) is NOT re-accumulated
//    element by element: running sums sum(x-K), sum((x-K)^2) are kept in double and down-dated as
//    samples leave, which gives the real-number sigma to ~1e-15.  The reference's float sigma
//    differs from it by at most ~3e-7 relative (analysis in DESIGN.md §4.3), so each decision is
//    taken for BOTH ends of the interval sigma*(1 -+ delta) (delta = 1e-6, plus the float-mean
//    rounding term); if the two ends disagree on any sample — it sits within 1e-6 relative of a
//    clip threshold — the pixel is NOT decided here: it is reported as FAST_FALLBACK and redone
//    by the exact engine (exact.cuh), which replays the reference operation by operation.
//  - the same fallback takes every pixel whose result depends on the order quickselect leaves
//    the stack in (the `N - r <= 4` guard regime), the `median == 0` / `kept == 0`
//    paths whose output is the median of a partially compacted stack, and non-finite samples.
//  - linear-fit clipping is replayed in float exactly as the reference does it (the sorted
//    order is canonical there), GESDT uses the down-dated sums with the same interval test on
//    every Grubbs decision that can change the outcome.
//
// Everything is __host__ __device__ so that tests/helpers/sorted_host.cu can run it on the CPU
// against the oracle.
#pragma once
#include "exact.cuh"
#include "../../include/sgpu.h"

namespace sgpu {

enum { FAST_DONE = 0, FAST_FALLBACK = 1 };

#ifdef SGPU_FAST_DEBUG	// host-side test builds only: why a stack was handed to the exact engine
static int g_fast_reason = 0;
SGPU_HD inline void fast_why(int x) {
#ifndef __CUDA_ARCH__
	g_fast_reason = x;
#endif
}
#define FAST_WHY(x) fast_why(x)
#else
#define FAST_WHY(x) ((void) 0)
#endif

#define SGPU_DELTA_SD 1.0e-6f	// half-width of the interval put around a float sd
#define SGPU_DELTA_WS 2.0e-6f	// same for the winsorized sigma (error feeds back through the clamp bounds)

struct FastRes {
	int status;	// FAST_DONE | FAST_FALLBACK
	int kept;	// number of samples kept
	double sum;	// their sum
	float pmin, pmax;	// their range (the weighted combine re-selects by value, median_and_mean.c:1045-1070)
};

// sum(x - K) and sum((x - K)^2) in double; K is a sample near the middle of the stack so that the
// variance below has no cancellation to speak of
// correctly rounded 1 / n (a division on the host, MUFU.RCP64H + Newton steps on the device)
SGPU_HD __forceinline__ double rcp_d(int n) {
#ifdef __CUDA_ARCH__
	return __drcp_rn((double) n);
#else
	return 1.0 / (double) n;
#endif
}
// a / n correctly rounded without the division routine (Markstein: y = RN(1/n), q = RN(a y),
// r = a - q n exactly by FMA, q' = RN(q + r y)); tests/test_sorted_host.py checks it against `/`
SGPU_HD __forceinline__ double div_by_int(double a, int n) {
	const double y = rcp_d(n), b = (double) n;
	const double q = a * y;
	const double r = fma(-q, b, a);
	return fma(r, y, q);
}

// 1 / n from the table of correctly rounded reciprocals when there is one (a load instead of MUFU.RCP64H and its
// Newton steps), else rcp_d: the same value either way
SGPU_HD __forceinline__ double rcp_t(const double *tab, int n) {
#ifdef __CUDA_ARCH__
	if (tab && n < SGPU_RCP_N) return __ldg(tab + n);
#endif
	(void) tab;
	return rcp_d(n);
}

struct Sums {
	double K, S1, S2;
	SGPU_HD __forceinline__ void init(float k) { K = (double) k; S1 = 0.0; S2 = 0.0; }
	SGPU_HD __forceinline__ void add(float x) { const double y = (double) x - K; S1 += y; S2 = fma(y, y, S2); }
	SGPU_HD __forceinline__ void sub(float x) { const double y = (double) x - K; S1 -= y; S2 = fma(-y, y, S2); }
	SGPU_HD __forceinline__ double sum(int n) const { return S1 + (double) n * K; }
	SGPU_HD __forceinline__ double mean(int n) const { return K + S1 * rcp_d(n); }
	SGPU_HD __forceinline__ double var(int n) const { return (S2 - S1 * S1 * rcp_d(n)) * rcp_d(n - 1); }
};

// the pre-accumulated sums of the whole stack (phase A leaves them beside the sorted stack), or,
// where there are none (host tests), a pass over the stack
template <class Acc>
SGPU_HD __forceinline__ Sums sums_of(const Acc &s, int n, const Sums *S0) {
	if (S0) return *S0;
	Sums S; S.init(s(n >> 1));
	for (int i = 0; i < n; i++) S.add(s(i));
	return S;
}

// interval [lo, hi] that contains the reference's float sd of the n samples summarised by S;
// `scale` is the 1.134f of the winsorized estimate (1 otherwise).  false: degenerate, fall back.
SGPU_HD __forceinline__ bool sd_interval(double var, double mean, int n, float delta, float scale, float *lo, float *hi) {
	if (!(var >= 0.0)) return false;
	// the reference subtracts a mean rounded to float: sum (x - m~)^2 = sum (x - m)^2 + n (m - m~)^2
	const double e = fabs(mean) * 5.9604645e-08;	// >= half an ulp of the float mean
	const double var_hi = var + e * e * (double) n / (double) (n - 1);
	const float s_lo = scale * sqrtf((float) var), s_hi = scale * sqrtf((float) var_hi);
	*lo = s_lo * (1.f - delta); *hi = s_hi * (1.f + delta);
	return *hi < FLT_MAX;
}

// quickmedian_float / quickmedian of the rank range [a, a+n) (sorting.c:195-273; n < 9 goes
// through the sorting networks, whose float version averages the middles in float, :366-415)
template <bool U16, class Acc>
SGPU_HD __forceinline__ double median_sorted(const Acc &s, int a, int n) {
	const int k = n >> 1;
	if (n & 1) return (double) s(a + k);
	const float x = s(a + k - 1), y = s(a + k);
	if (!U16 && n < 9) return (x + y) / 2.0;
	return ((double) x + (double) y) / 2.0;
}

template <class Acc>
SGPU_HD __forceinline__ void finish_range(const Acc &s, int a, int b, const Sums &S, FastRes &o) {
	o.status = FAST_DONE; o.kept = b - a; o.sum = S.sum(b - a); o.pmin = s(a); o.pmax = s(b - 1);
}

// ---- NO_REJEC / kept <= 1 -----------------------------------------------------------------------
template <class Acc>
SGPU_HD void fast_all(const Acc &s, int n, const Sums *S0, FastRes &o) {
	double sum = 0.0;
	if (S0) sum = S0->sum(n);
	else for (int i = 0; i < n; i++) sum += (double) s(i);
	o.status = FAST_DONE; o.kept = n; o.sum = sum;
	o.pmin = n ? s(0) : 0.f; o.pmax = n ? s(n - 1) : 0.f;
}

// ---- PERCENTILE (rejection_float.c:31-44,160-173 / median_and_mean.c:590-603,762-776) ----------
template <bool U16, class Acc>
SGPU_HD void fast_percentile(const Acc &s, int n, const RejCfg &c, const Sums *S0, int rej[2], FastRes &o) {
	const double median = median_sorted<U16>(s, 0, n);
	const float medf = (float) median;
	if (U16 ? (medf == 0.f) : (median == 0.0)) { o.status = FAST_FALLBACK; return; }
	int a = 0, b = n;
	if (!U16) {
		while (a < b && medf - s(a) > medf * c.s0) a++;
		while (b > a && s(b - 1) - medf > medf * c.s1) b--;
	} else {
		while (a < b && (medf - s(a)) / medf > c.s0) a++;
		while (b > a && (s(b - 1) - medf) / medf > c.s1) b--;
	}
	if (a >= b) { o.status = FAST_FALLBACK; return; }	// kept == 0: median of the compacted stack
	rej[0] += a; rej[1] += n - b;
	double sum = 0.0;
	if (S0) {	// the whole-stack sum less what was rejected
		sum = S0->sum(n);
		for (int i = 0; i < a; i++) sum -= (double) s(i);
		for (int i = b; i < n; i++) sum -= (double) s(i);
	} else for (int i = a; i < b; i++) sum += (double) s(i);
	o.status = FAST_DONE; o.kept = b - a; o.sum = sum; o.pmin = s(a); o.pmax = s(b - 1);
}

// one sigma-clip pass over the kept range [a, b) with the sd known to lie in [slo, shi]:
// 0 = decided (a2, b2 set), 1 = undecidable here
template <class Acc>
SGPU_HD __forceinline__ int clip_range(const Acc &s, int a, int b, float medf, float slo, float shi, float s0, float s1,
		int *a2o, int *b2o) {
	int a2 = a, b2 = b;
	const float tl = shi * s0, th = shi * s1;	// sigma_clipping_float: `sigma * sigmalow`
	while (a2 < b && medf - s(a2) > tl) a2++;
	while (b2 > a2 && s(b2 - 1) - medf > th) b2--;
	if (a2 >= b2) return 1;
	if (medf - s(a2) > slo * s0 || s(b2 - 1) - medf > slo * s1) return 1;	// the low end of the interval rejects more
	*a2o = a2; *b2o = b2;
	return 0;
}

// ---- SIGMA (rejection_float.c:174-209 / median_and_mean.c:777-813) ------------------------------
template <bool U16, class Acc>
SGPU_HD void fast_sigma(const Acc &s, int n, const RejCfg &c, const Sums *S0, int rej[2], FastRes &o) {
	o.status = FAST_FALLBACK;
	{
		const double median = median_sorted<U16>(s, 0, n);
		if (U16 ? ((float) median == 0.f) : (median == 0.0)) return;
	}
	Sums S = sums_of(s, n, S0);
	int a = 0, b = n, r = 0;
	for (;;) {
		const int nn = b - a;
		if (nn - r <= 4) break;		// the guard refuses every rejection of this pass
		float slo, shi;
		if (!sd_interval(S.var(nn), S.mean(nn), nn, SGPU_DELTA_SD, 1.f, &slo, &shi)) return;
		const float medf = (float) median_sorted<U16>(s, a, nn);
		int a2, b2;
		if (clip_range(s, a, b, medf, slo, shi, c.s0, c.s1, &a2, &b2)) return;
		const int K = (a2 - a) + (b - b2);
		if (K == 0) break;
		if (nn - r - K < 4) return;	// the guard would cut this pass short: order dependent
		for (int i = a; i < a2; i++) S.sub(s(i));
		for (int i = b2; i < b; i++) S.sub(s(i));
		rej[0] += a2 - a; rej[1] += b - b2; r += K; a = a2; b = b2;
		if (b - a <= 3) break;
	}
	finish_range(s, a, b, S, o);
}

// ---- MAD (rejection_float.c:174-209 with siril_stats_float_mad, statistics_float.c:79-101 -> findMinMaxPercentile,
// rt_algo.cc:38-171; median_and_mean.c:777-813 with siril_stats_ushort_mad, statistics.c:133-154) --------------------
// On the sorted stack the absolute deviations from the median come out in ascending order from a two-pointer walk
// outwards from where the median sits, so neither the reference's scratch copy nor its histogram is needed: the
// histogram median of the float version depends only on how many deviations fall into each of its n bins, and the
// bin of a deviation grows with it — the walk meets the bins in order.  The result is the reference's value exactly
// (same float operations per sample, integer counts), not an interval.
// first rank in [a, a + n) whose sample is >= x
template <class Acc>
SGPU_HD __forceinline__ int lower_rank(const Acc &s, int a, int n, float x) {
	int lo = a, hi = a + n;
	while (lo < hi) { const int mid = (lo + hi) >> 1; if (s(mid) < x) lo = mid + 1; else hi = mid; }
	return lo;
}

template <class Acc>
SGPU_HD float mad_sorted_f32(const Acc &s, int a, int n, double m) {
	const float med = (float) m;
	const int p = lower_rank(s, a, n, med);
	int i = p - 1, j = p;
	const int e = a + n;
	float vi = i >= a ? fabsf(s(i) - med) : __builtin_huge_valf(), vj = j < e ? fabsf(s(j) - med) : __builtin_huge_valf();
	const float mn = vi < vj ? vi : vj;
	const float dl = fabsf(s(a) - med), dh = fabsf(s(e - 1) - med);
	const float mx = dl < dh ? dh : dl;
	if (fabsf(mx - mn) == 0.f) return mn;
	const unsigned hs = n < 65536 ? (unsigned) n : 65536u;
	const float scale = (hs - 1) / (mx - mn);
	const float thresh = 0.5f * n;
	unsigned cum = 0, cur = 0, bin = 0;	// samples in the bins below `bin`, and in it
	for (int k = 0; k < n; k++) {
		const bool low = vi <= vj;
		const float v = low ? vi : vj;
		const unsigned bk = (unsigned) (u16) (scale * (v - mn));
		if (bk != bin) {
			if (!((float) (cum + cur) < thresh)) break;	// `while (count < thresh) count += histo[k++]` stopped on the bin before
			cum += cur; cur = 0; bin = bk;
		}
		cur++;
		if (low) { i--; vi = i >= a ? fabsf(s(i) - med) : __builtin_huge_valf(); }
		else { j++; vj = j < e ? fabsf(s(j) - med) : __builtin_huge_valf(); }
	}
	const unsigned count = cum + cur, kbin = bin + 1;
	const float c0 = (float) count - thresh, c1 = thresh - (float) cum;
	float v = (c1 * (float) kbin + c0 * (float) (kbin - 1)) / (c0 + c1);
	v /= scale; v += mn;
	{ const float t = v < mx ? v : mx; v = mn < t ? t : mn; }
	return v;
}

// WORD samples: the exact median of |x - round(m)| (histogram_median, sorting.c:577-642)
template <class Acc>
SGPU_HD float mad_sorted_u16(const Acc &s, int a, int n, double m) {
	const int med = round_to_int_d(m);
	const int p = lower_rank(s, a, n, (float) med);
	int i = p - 1, j = p;
	const int e = a + n, k = n / 2;
	int prev = 0, cur = 0;
	for (int q = 0; q <= k; q++) {	// order statistics k - 1 and k of the deviations
		const int vi = i >= a ? med - (int) s(i) : 0x7fffffff, vj = j < e ? (int) s(j) - med : 0x7fffffff;
		prev = cur;
		if (vi <= vj) { cur = vi; i--; } else { cur = vj; j++; }
	}
	const double r = (n % 2 == 0) ? (prev + cur) / 2.0 : (double) cur;
	return (float) r;
}

template <bool U16, class Acc>
SGPU_HD void fast_mad(const Acc &s, int n, const RejCfg &c, const Sums *S0, int rej[2], FastRes &o) {
	o.status = FAST_FALLBACK;
	double median = median_sorted<U16>(s, 0, n);
	if (U16) median = (double) (float) median;	// the WORD path keeps the median as a float (median_and_mean.c:705)
	if (median == 0.0) return;
	Sums S = sums_of(s, n, S0);
	int a = 0, b = n, r = 0;
	bool first = true;
	for (;;) {
		const int nn = b - a;
		if (nn - r <= 4) break;		// the guard refuses every rejection of this pass
		// the MAD is taken about the median of the PASS BEFORE; the median is renewed after it (:183-186)
		const float var = U16 ? mad_sorted_u16(s, a, nn, median) : mad_sorted_f32(s, a, nn, median);
		if (!first) { median = median_sorted<U16>(s, a, nn); if (U16) median = (double) (float) median; }
		first = false;
		int a2, b2;
		if (clip_range(s, a, b, (float) median, var, var, c.s0, c.s1, &a2, &b2)) return;
		const int K = (a2 - a) + (b - b2);
		if (K == 0) break;
		if (nn - r - K < 4) return;	// the guard would cut this pass short: order dependent
		for (int i = a; i < a2; i++) S.sub(s(i));
		for (int i = b2; i < b; i++) S.sub(s(i));
		rej[0] += a2 - a; rej[1] += b - b2; r += K; a = a2; b = b2;
		if (b - a <= 3) break;
	}
	finish_range(s, a, b, S, o);
}

// ---- SIGMEDIAN (rejection_float.c:210-222 / median_and_mean.c:814-830) -----------------------------
// Every sample beyond median -+ k sigma is REPLACED by the median and the pass is repeated until none is; nothing is
// removed.  On the sorted stack the replaced samples are the two ends of the originals' range [a, b); the copies
// that took their place are kept as a short list of (value, count) groups — each value a median, so inside the
// range — and sums, medians and extremes are those of originals + groups.
template <bool U16, class Acc>
SGPU_HD void fast_sigmedian(const Acc &s, int n, const RejCfg &c, const Sums *S0, int rej[2], FastRes &o) {
	o.status = FAST_FALLBACK;
	constexpr int MAXG = 6;
	float gv[MAXG]; int gc[MAXG], gpos[MAXG];	// ascending by value; gpos: originals (of all n) below the value
	int ng = 0;
	if (U16 && (float) median_sorted<true>(s, 0, n) == 0.f) return;	// median_and_mean.c:747-759: the exact engine's case
	Sums S = sums_of(s, n, S0);
	int a = 0, b = n;
	auto rank = [&](int k) -> float {	// k-th smallest of originals [a, b) + groups
		int used = 0;
		for (int g = 0; g < ng; g++) {
			int below = gpos[g] - a; below = below < 0 ? 0 : (below > b - a ? b - a : below);
			const int base = below + used;
			if (k < base) return s(a + k - used);
			if (k < base + gc[g]) return gv[g];
			used += gc[g];
		}
		return s(a + k - used);
	};
	for (;;) {
		float slo, shi;
		if (!sd_interval(S.var(n), S.mean(n), n, SGPU_DELTA_SD, 1.f, &slo, &shi)) return;
		const int k = n >> 1;
		double med;
		if (n & 1) med = (double) rank(k);
		else { const float x = rank(k - 1), y = rank(k); med = (!U16 && n < 9) ? (x + y) / 2.0 : ((double) x + (double) y) / 2.0; }
		const float medf = (float) med;
		int a2, b2;
		if (clip_range(s, a, b, medf, slo, shi, c.s0, c.s1, &a2, &b2)) return;
		for (int g = 0; g < ng; g++)	// a copy that a later, narrower pass would replace again: the exact engine's case
			if (medf - gv[g] > slo * c.s0 || gv[g] - medf > slo * c.s1) return;
		const int K = (a2 - a) + (b - b2);
		if (K == 0) break;
		if (ng == MAXG) return;
		const float rv = U16 ? (float) (u16) medf : medf;	// `stack[f] = (T) median`
		for (int i = a; i < a2; i++) S.sub(s(i));
		for (int i = b2; i < b; i++) S.sub(s(i));
		{ const double y = (double) rv - S.K; S.S1 += (double) K * y; S.S2 += (double) K * y * y; }
		rej[0] += a2 - a; rej[1] += b - b2; a = a2; b = b2;
		int g = ng;
		while (g > 0 && gv[g - 1] > rv) { gv[g] = gv[g - 1]; gc[g] = gc[g - 1]; gpos[g] = gpos[g - 1]; g--; }
		gv[g] = rv; gc[g] = K; gpos[g] = lower_rank(s, 0, n, rv); ng++;
	}
	o.status = FAST_DONE; o.kept = n; o.sum = S.sum(n);
	o.pmin = s(a); o.pmax = s(b - 1);
	if (ng) { o.pmin = gv[0] < o.pmin ? gv[0] : o.pmin; o.pmax = gv[ng - 1] > o.pmax ? gv[ng - 1] : o.pmax; }
}

// ---- WINSORIZED, float samples (rejection_float.c:223-259) --------------------------------------
template <class Acc>
SGPU_HD void fast_winsorized(const Acc &s, int n, const RejCfg &c, const Sums *S0, int rej[2], FastRes &o) {
	o.status = FAST_FALLBACK;
	Sums S = sums_of(s, n, S0);
	int a = 0, b = n, r = 0;
	for (;;) {
		const int nn = b - a;
		if (nn - r <= 4) break;
		const double rn = rcp_d(nn), rn1 = rcp_d(nn - 1);
		const double var0 = (S.S2 - S.S1 * S.S1 * rn) * rn1;
		if (!(var0 >= 0.0)) return;
		float sigma = sqrtf((float) var0), sigma0;
		const float medf = (float) median_sorted<false>(s, a, nn);
		// w_stack = stack clamped to [cl, cu]; [ia, ib) is the part the clamp leaves alone
		int ia = a, ib = b, it = 0;
		Sums In = S;
		float cl = -FLT_MAX, cu = FLT_MAX;
		double varw, meanw;
		// stopping points the reference MAY have taken: where its convergence test is too close to
		// call the iteration goes on and the sigma it would have stopped with is remembered
		float clo = FLT_MAX, chi = 0.f;
		for (;;) {
			const float m0 = medf - 1.5f * sigma, m1 = medf + 1.5f * sigma;
			const float ncl = clampw(cl, m0, m1), ncu = clampw(cu, m0, m1);
			if (ncl < cl || ncu > cu) return;	// clamp bounds only ever tighten with sane data
			cl = ncl; cu = ncu;
			while (ia < ib && s(ia) < cl) { In.sub(s(ia)); ia++; }
			while (ib > ia && s(ib - 1) > cu) { ib--; In.sub(s(ib)); }
			const double yl = (double) cl - In.K, yh = (double) cu - In.K;
			const double nl = (double) (ia - a), nh = (double) (b - ib);
			const double W1 = In.S1 + nl * yl + nh * yh;
			const double W2 = In.S2 + nl * yl * yl + nh * yh * yh;
			varw = (W2 - W1 * W1 * rn) * rn1;
			meanw = In.K + W1 * rn;
			if (!(varw >= 0.0)) return;
			sigma0 = sigma;
			sigma = 1.134f * sqrtf((float) varw);
			const float d = fabsf(sigma - sigma0), thr = sigma0 * 0.0005f;
			if (++it > 64) return;
			if (fabsf(d - thr) <= 2.f * SGPU_DELTA_WS * sigma0) {
				float l, h;
				if (!sd_interval(varw, meanw, nn, SGPU_DELTA_WS, 1.134f, &l, &h)) return;
				clo = l < clo ? l : clo; chi = h > chi ? h : chi;
				continue;
			}
			if (!(d > thr)) break;
		}
		float slo, shi;
		if (!sd_interval(varw, meanw, nn, SGPU_DELTA_WS, 1.134f, &slo, &shi)) return;
		slo = clo < slo ? clo : slo; shi = chi > shi ? chi : shi;	// every sigma the reference may have ended on
		int a2, b2;
		if (clip_range(s, a, b, medf, slo, shi, c.s0, c.s1, &a2, &b2)) return;
		const int K = (a2 - a) + (b - b2);
		if (K == 0) break;
		if (nn - r - K < 4) return;
		for (int i = a; i < a2; i++) S.sub(s(i));
		for (int i = b2; i < b; i++) S.sub(s(i));
		rej[0] += a2 - a; rej[1] += b - b2; r += K; a = a2; b = b2;
		if (b - a <= 3) break;
	}
	finish_range(s, a, b, S, o);
}

// ---- LINEARFIT (rejection_float.c:260-300 / median_and_mean.c:868-902) --------------------------
// replayed in float exactly as the reference does: its quicksort leaves the canonical ascending
// order, which is what `s` already holds (and compaction keeps it), so nothing here depends on
// anything but the values.  `xf` is the reference's table 1.f / (i + 1) (median_and_mean.c:1493).
// ranks [i0, i0 + 8) of the stack (0.f beyond n): the loads are independent of the float chains
// below, so a block is in flight while the previous one is consumed
template <class Acc>
SGPU_HD __forceinline__ void acc_load8(const Acc &s, int i0, int n, float (&x)[8]) {
#pragma unroll
	for (int k = 0; k < 8; k++) x[k] = (i0 + k < n) ? s(i0 + k) : 0.f;
}
SGPU_HD __forceinline__ void xf_load8(const float *xf, int i0, float (&r)[8]) {
#pragma unroll
	for (int k = 0; k < 8; k++) r[k] = xf ? xf[i0 + k] : 1.f / (float) (i0 + k + 1);
}

template <class Acc>
SGPU_HD void fast_linearfit(Acc &s, int n, const RejCfg &c, int rej[2], FastRes &o) {
	int N = n, r = 0, changed;
	const float *xf = c.xf;		// holds >= nb_frames rounded up to 8 entries when given
	do {
		float x[8], rc[8];
		float m_y = 0.f;
		for (int i0 = 0; i0 < N; i0 += 8) {
			acc_load8(s, i0, N, x); xf_load8(xf, i0, rc);
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int i = i0 + k;
				if (i == 0) m_y = x[0];
				else if (i < N) m_y += (x[k] - m_y) * rc[k];
			}
		}
		float m_dxdy = 0.f, dx = -c.m_x;
		for (int i0 = 0; i0 < N; i0 += 8) {
			acc_load8(s, i0, N, x); xf_load8(xf, i0, rc);
#pragma unroll
			for (int k = 0; k < 8; k++) {
				if (i0 + k < N) {
					const float dy = x[k] - m_y;
					m_dxdy += (dx * dy - m_dxdy) * rc[k];
					dx += 1.f;
				}
			}
		}
		const float a = m_dxdy * c.m_dx2;	// slope
		const float b = m_y - c.m_x * a;	// intercept
		float sigma = 0.f;
		for (int i0 = 0; i0 < N; i0 += 8) {
			acc_load8(s, i0, N, x);
#pragma unroll
			for (int k = 0; k < 8; k++) if (i0 + k < N) sigma += fabsf(x[k] - (a * (i0 + k) + b));
		}
		sigma /= (float) N;
		int out = 0;
		const float tl = sigma * c.s0, th = sigma * c.s1;
		for (int i0 = 0; i0 < N; i0 += 8) {
			acc_load8(s, i0, N, x);
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int f = i0 + k;
				if (f < N) {
					const float v = x[k]; int flag = 0;
					if (N - r > 4) {
						if (a * f + b - v > tl) { rej[0]++; flag = -1; r++; }
						else if (v - a * f - b > th) { rej[1]++; flag = 1; r++; }
					}
					if (!flag) { if (f != out) s.set(out, v); out++; }
				}
			}
		}
		changed = N != out; N = out;
	} while (changed && N > 3);
	double sum = 0.0;
	for (int i = 0; i < N; i++) sum += (double) s(i);
	o.status = FAST_DONE; o.kept = N; o.sum = sum; o.pmin = N ? s(0) : 0.f; o.pmax = N ? s(N - 1) : 0.f;
}

// ---- GESDT (rejection_float.c:301-348 / median_and_mean.c:903-948) ------------------------------
// The reference removes, max_outliers times, whichever end of the sorted window deviates most
// from the window mean (grubbs_stat, :82-98), records whether G = dev / sd beats the critical
// value, and confirm_outliers (median_and_mean.c:685-701) then rejects every iteration up to the
// last one that did (never fewer than two).  The window mean and sd come from the down-dated
// sums; the quirk that a high-end removal flags index size-1 of the SHRUNK window in the
// ORIGINAL sorted stack (:333, `out[iter].i = max_index`) is reproduced: after k low removals
// it flags the element k places below the one it removed.
template <bool U16, class Acc>
SGPU_HD void fast_gesdt(const Acc &s, int n, int nb, const RejCfg &c, const Sums *S0, int rej[2], FastRes &o) {
	o.status = FAST_FALLBACK;
	const int removed = nb - n;
	double median;
	if (n & 1) median = (double) s(n / 2);
	else if (U16) median = ((double) s((n - 1) / 2) + (double) s(n / 2)) / 2.0;
	else median = (s((n - 1) / 2) + s(n / 2)) / 2.0;	// gsl_stats_float_median_from_sorted_data: float add
	if (U16) median = (double) (float) median;
	int max_out = (int) ((float) nb * c.s0);
	if (removed >= max_out) { fast_all(s, n, S0, o); return; }
	max_out -= removed;
	if (2 * max_out >= n || n - max_out < 3) { FAST_WHY(71); return; }	// flagged positions could collide: leave it to the exact engine
	Sums S = sums_of(s, n, S0);
	const double total = S.sum(n);
	const bool exact_sums = U16 || (s(0) > 0.f && s(n - 1) <= s(0) * 65536.f && n <= 2048);
	int lo = 0, clo = 0, chi = 0, first_low = -1;
	double fsum = 0.0;
	int last = -1, amb = -1, s_clo = 0, s_chi = 0, s_first_low = -1, s_nl = 0;
	double s_fsum = 0.0;
	for (int it = 0, size = n; it < max_out; it++, size--) {
		// grubbs_stat's mean is (float) (sum / N) with the sum in double.  When every partial sum is
		// exact in double (samples of one sign within 2^16 of each other, or integers) S.sum() IS
		// the reference's sum, so the float mean is bit-identical, ties included; otherwise the
		// reference's own summation order matters whenever the quotient sits on a rounding boundary
		const double r0 = rcp_t(c.rcp, size), r1 = rcp_t(c.rcp, size - 1), b = (double) size;
		const double a = S.sum(size);
		double avg_d = a * r0;		// div_by_int(a, size) with the reciprocal from the table
		avg_d = fma(fma(-avg_d, b, a), r0, avg_d);
		const float avg = (float) avg_d;
		if (!exact_sums) {
			const double tiny = fabs(avg_d) * 1e-12;
			if ((float) (avg_d + tiny) != avg || (float) (avg_d - tiny) != avg) { FAST_WHY(72); return; }
		}
		const float xl = s(lo), xh = s(lo + size - 1);
		float dev = avg - xl;
		const float md2 = xh - avg;
		const bool low_end = !(md2 > dev);
		if (!low_end) dev = md2;
		if (it > 1) {
			// G = dev / sd against the critical value, decided for both ends of the interval [slo, shi] that holds the
			// reference's float sd (sd_interval).  First in squares, in double, with room for every rounding between here
			// and the float quotient (the sd's two float roundings and its (1 +- delta) factors, the division's one): only
			// what falls between the two bounds takes the float path below
			const double var = (S.S2 - S.S1 * S.S1 * r0) * r1;
			if (!(var >= 0.0)) { FAST_WHY(73); return; }
			const double e = fabs(avg_d) * 5.9604645e-08;
			const double var_hi = var + e * e * 2.0;	// sd_interval's term, n / (n - 1) <= 2
			const double cr = (double) c.crit[it + removed], d2 = (double) dev * (double) dev, c2 = cr * cr;
			const bool normal = var > 1e-30 && var_hi < 1e30;	// the float sd is a normal number: its roundings are relative
			if (normal && d2 > c2 * var_hi * (1.0 + 4.0e-6)) last = -2;	// certainly an outlier: snapshot below
			else if (!normal || !(d2 < c2 * var * (1.0 - 4.0e-6))) {	// too close to call in squares (or a degenerate sd): the float path
				float slo, shi;
				if (!sd_interval(var, avg_d, size, SGPU_DELTA_SD, 1.f, &slo, &shi)) { FAST_WHY(73); return; }
				const float crf = c.crit[it + removed];
				if (dev / shi > crf) last = -2;
				else if (dev / slo > crf) amb = it;
			}
		}
		float xrem, xflag;
		if (low_end) { xrem = xl; xflag = xl; lo++; if (first_low < 0) first_low = it; }
		else { xrem = xh; xflag = s(n - 1 - it); }
		S.sub(xrem);
		if ((double) xrem >= median) chi++; else clo++;
		fsum += (double) xflag;
		if (it <= 1 || last == -2) { last = it; s_clo = clo; s_chi = chi; s_fsum = fsum; s_first_low = first_low; s_nl = lo; }
	}
	if (amb > last) { FAST_WHY(74); return; }
	rej[0] += s_clo; rej[1] += s_chi;
	o.status = FAST_DONE; o.kept = n - (last + 1); o.sum = total - s_fsum;
	o.pmin = s(s_nl);
	o.pmax = (s_first_low >= 0) ? s(n - 1 - s_first_low) : s(n - last - 2);
}

// ---- combine: mean_and_reject (median_and_mean.c:956-1101) on the sorted stack ------------------
// `s` holds the n non-zero samples ascending; nb is the full stack size; orig(f) re-reads the
// normalized sample of frame f (the reference's o_stack).  status FAST_FALLBACK: result unset.
// the rejection alone: which rank range is kept, its sum and its extremes
template <bool U16, class Acc>
SGPU_HD void fast_reject(Acc &s, int n, int nb, const RejCfg &c, const Sums *S0, int rej[2], FastRes &o) {
	o.status = FAST_DONE; o.kept = n; o.sum = 0.0; o.pmin = o.pmax = 0.f;
	if (n == 0) return;	// all samples null: median of zeros
	if (n == 1 || c.type == SGPU_NO_REJEC) fast_all(s, n, S0, o);
	else switch (c.type) {
	case SGPU_PERCENTILE: fast_percentile<U16>(s, n, c, S0, rej, o); break;
	case SGPU_SIGMA: fast_sigma<U16>(s, n, c, S0, rej, o); break;
	case SGPU_WINSORIZED: if (U16) o.status = FAST_FALLBACK; else fast_winsorized(s, n, c, S0, rej, o); break;
	case SGPU_LINEARFIT: fast_linearfit(s, n, c, rej, o); break;
	case SGPU_GESDT: fast_gesdt<U16>(s, n, nb, c, S0, rej, o); break;
	case SGPU_MAD: fast_mad<U16>(s, n, c, S0, rej, o); break;
	case SGPU_SIGMEDIAN: fast_sigmedian<U16>(s, n, c, S0, rej, o); break;
	default: o.status = FAST_FALLBACK; break;
	}
	if (o.status != FAST_DONE || o.kept == 0) o.status = FAST_FALLBACK;
}

// the combine of a decided stack (o.status == FAST_DONE, n > 0)
template <typename OrigFn, typename W>
SGPU_HD double fast_combine(const FastRes &o, int nb, const W &w, OrigFn orig) {
	if (w) {	// frame weights (a plain pointer) or PixWeights with the pixel's feather-mask samples (exact.cuh)
		double sum = 0.0, norm = 0.0;
		for (int f = 0; f < nb; f++) {
			const float v = orig(f);
			if (v >= o.pmin && v <= o.pmax && v != 0.f) { const double n = w[f]; sum += (double) v * n; norm += n; }
		}
		if (norm == 0. || sum == 0.) {
			sum = 0.;
			for (int f = 0; f < nb; f++) { const float v = orig(f); if (v >= o.pmin && v <= o.pmax && v > 0) sum += (double) v; }
			return sum / (double) o.kept;
		}
		return sum / norm;
	}
	return div_by_int(o.sum, o.kept);	// the correctly rounded quotient, without the division routine
}

template <bool U16, class Acc, typename OrigFn, typename W>
SGPU_HD double fast_mean_and_reject(Acc &s, int n, int nb, const RejCfg &c, const W &w, OrigFn orig,
		const Sums *S0, int rej[2], int *status) {
	*status = FAST_DONE;
	if (n == 0) return 0.0;
	FastRes o;
	fast_reject<U16>(s, n, nb, c, S0, rej, o);
	*status = o.status;
	if (o.status != FAST_DONE) return 0.0;
	return fast_combine(o, nb, w, orig);
}

}  // namespace sgpu
