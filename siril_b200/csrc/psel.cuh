// psel.cuh — per-pixel logic of the SELECTION engine (stack_psel.cu): mean-with-rejection and
// median WITHOUT sorting the stack.
//
// The reference never needs the order of a pixel stack, only (i) the median of the kept samples
// (quickmedian_float, sorting.c:240-273), (ii) their sums (siril_stats_float_sd,
// statistics.h:82-107; the mean, median_and_mean.c:1071-1090) and (iii) which samples lie beyond
// median -+ k * sigma (rejection_float.c:49-60) — a few samples at both ends of the stack.  So
// the kernel (phase A, stack_psel.cuh) keeps the N samples of a pixel in the registers of a small
// lane group and makes three streaming passes over them:
//   sums     n, sum x, sum x^2 in double, min, max;
//   count    how many samples lie below the (trimmed) mean and how many within one sd of it:
//            the first locates the median, the second gives a scale no outlier can inflate;
//   extract  every sample inside a narrow window around the predicted median ("centre", a few
//            dozen ranks wide) or beyond a radius the clipping thresholds cannot fall inside
//            ("tails") is appended to a small per-lane list; the rest is only counted.
// Phase B (this file, one thread per pixel) sorts the three short lists by insertion into the
// first rows of the pixel's shared-memory column — they stand for ranks [0, nl), [cl, cl+cc) and
// [n-nh, n) of the stack — and then runs the very same sorted-stack rejection code as the sorted /
// lane-sort engines (sorted.cuh) through an accessor that knows which ranks exist.  A lookup of a
// rank that was not extracted sets a flag and the pixel is handed to the exact engine, exactly
// as sorted.cuh's own FAST_FALLBACK; the window parameters are chosen (scripts/psel_plan_sim.py)
// so that this happens for < 1e-5 of well-behaved pixels.
//
// Everything here is __host__ __device__: tests/helpers/psel_host.cu replays phase A on the CPU
// and runs this code against the oracle.
#pragma once
#include "sorted.cuh"

namespace sgpu {

enum { PSEL_WEIRD = 1, PSEL_TRIVIAL = 2, PSEL_HANDOVER = 4, PSEL_DONE = 8 };	// DONE (split form): the stack kernel finished the pixel itself

#define PSEL_NAN __int_as_float_hd(0x7fffffff)
SGPU_HD __forceinline__ float __int_as_float_hd(int i) {
#ifdef __CUDA_ARCH__
	return __int_as_float(i);
#else
	float f; memcpy(&f, &i, 4); return f;
#endif
}

// per pixel, written by phase A, read by phase B
struct PselHdr {
	double S1, S2, K;	// sum (x - K), sum (x - K)^2 over the n kept samples (K = 0 unless the plain sums were ill-conditioned)
	float mn, mx;		// their minimum and maximum (ranks 0 and n - 1)
	float p1, w;		// centre window: |x - p1| < w
	float blo, bhi;		// every kept sample outside the low list is >= blo, outside the high list <= bhi
	int n;			// kept samples (non-null in mean mode, all N in median mode)
	int cl;			// kept samples below the centre window: rank of the first centre entry
	int nl, cc, m;		// list entries below / inside the centre window, and all of them (the rest are above)
	int tid0;		// CTA thread of the group's first lane (its lists follow at tid0 + t)
	int flags;
};

// split form (stack_psel.cu): record of one pixel in global memory between the two kernels
#define PSEL_REC_HDR 16		// S1, S2, K (doubles), mn, mx, blo, bhi, n, cl, nl, cc, m, flags
#define PSEL_REC_ENTRIES 64	// more entries than this: hand-over
#define PSEL_REC_WORDS (PSEL_REC_HDR + PSEL_REC_ENTRIES)

#define PSEL_CAP 64	// longest centre / tail list phase B sorts by insertion (ties can make a class arbitrarily long)
#define PSEL_CAPT 48
#define PSEL_ROWS (PSEL_CAP + 2 * PSEL_CAPT + 8)	// rows of the column the sorted lists (and two stand-ins) can take

// list slots per lane for groups of T lanes (a lane stops appending when fewer than 8 are left)
SGPU_HD constexpr int psel_list_cap(int T) { return T == 1 ? 56 : (T == 2 ? 40 : (T == 4 ? 32 : (T == 8 ? 24 : 16))); }

struct PselPlan {
	float p1, w;		// centre window
	float cT, rT;		// tails: |x - cT| > rT
	float blo, bhi;
};

// planning arithmetic: approximate division / square root on the device (the PLACEMENT of the windows
// is all that depends on it, never a result), the plain operators on the host
#ifdef __CUDA_ARCH__
#define PSEL_DIV(a, b) __fdividef((a), (b))
#define PSEL_SQRT(x) ((x) * rsqrtf(x))
#else
#define PSEL_DIV(a, b) ((a) / (b))
#define PSEL_SQRT(x) sqrtf(x)
#endif

// inverse of the standard normal distribution function for 0.02425 <= p <= 0.97575 (central
// rational approximation of P. Acklam, |relative error| < 1.2e-9 in exact arithmetic; only the
// PLACEMENT of the windows depends on it, never a result)
SGPU_HD inline float psel_ndtri(float p) {
	const float q = p - 0.5f, r = q * q;
	const float num = (((((-3.969683028665376e+01f * r + 2.209460984245205e+02f) * r - 2.759285104469687e+02f) * r
			+ 1.383577518672690e+02f) * r - 3.066479806614716e+01f) * r + 2.506628277459239e+00f) * q;
	const float den = ((((-5.447609879822406e+01f * r + 1.615858368580409e+02f) * r - 1.556989798598866e+02f) * r
			+ 6.680131188771972e+01f) * r - 1.328068155288572e+01f) * r + 1.f;
	return PSEL_DIV(num, den);
}

// ---- planning (group level, after each pass) ----------------------------------------------------
// after the sums: where the count pass looks.  Mean and sd of the samples less their minimum and
// maximum (one huge outlier is the common way a plain sd gets inflated).  false: degenerate
// (too few samples, no spread, or sums too ill-conditioned to tell) — take every sample as centre.
SGPU_HD inline bool psel_plan_count(int n, double S1, double S2, double K, float mn, float mx, float *mean, float *hw) {
	if (n < 12) return false;
	const double a = (double) mn - K, b = (double) mx - K;
	const double T1 = S1 - a - b, T2 = S2 - a * a - b * b;
	const double rnt = rcp_d(n - 2);
	const double m = T1 * rnt, ss = T2 - T1 * m;		// sum of squares about the trimmed mean
	if (!(ss > 0.0) || !(ss * 1e9 > T2)) return false;
	*mean = (float) (K + m);
	*hw = PSEL_SQRT(PSEL_DIV((float) ss, (float) (n - 3)));	// one sd either side: two thirds of a clean stack, the bulk's scale to ~1 / sqrt(n)
	return *hw > 0.f && *hw < FLT_MAX;
}

// the window's half-width in units of the bulk's sd, u = hw / sigma, when the window is NOT centred on the bulk:
// its centre sits at the q-quantile (zq = ndtri(q)) and it holds the fraction f:  Phi(zq + u) - Phi(zq - u) = f.
// Newton from the centred answer; only reached when outliers have dragged the mean off the bulk (|zq| > 0.4).
SGPU_HD inline float psel_offcentre_u(float zq, float f, float u) {
	for (int it = 0; it < 4; it++) {
		const float a = zq + u, b = zq - u;
		const float g = 0.5f * (erfcf(-a * 0.70710678f) - erfcf(-b * 0.70710678f)) - f;
		const float dg = 0.39894228f * (expf(-0.5f * a * a) + expf(-0.5f * b * b));
		if (!(dg > 1e-6f)) break;
		u -= g / dg;
		if (!(u > 0.05f)) u = 0.05f;
	}
	return u;
}

// after the count pass (c_neg samples below `mean`, c_in within hw of it): the windows.
// 0: plan filled; 1: the count window swallowed nearly everything (scale still inflated) — count
// again with the smaller *hw; 2: degenerate
SGPU_HD inline int psel_plan_window(int n, float *meanp, float *hw, int c_neg, int c_in, bool tails, int mode,
		float s0, float s1, int attempt, PselPlan *pl) {
	const float fn = (float) n, mean = *meanp;
	const float rfn = PSEL_DIV(1.f, fn);
	float f = (float) c_in * rfn;
	float q = ((float) c_neg + 0.5f) * rfn;
	q = q < 0.03f ? 0.03f : (q > 0.97f ? 0.97f : q);
	const float zq = psel_ndtri(q);
	if (f >= 0.93f && attempt < 3) {	// several outliers inflated the sd: the bulk is a fraction of the window (too little of it outside
						// to tell its scale) and the mean sits off its middle: count again, about one bulk sd either side
		if (f > 0.97f) {	// scale unknown: a quarter of the window, and a cautious step towards the median
			*hw *= 0.25f;
			*meanp = mean - 0.5f * *hw * zq;
		} else {
			float u = psel_ndtri(0.5f + 0.5f * f);
			if (fabsf(zq) > 0.4f) u = psel_offcentre_u(zq, f, u);
			*hw = PSEL_DIV(*hw, u);
			*meanp = mean - *hw * zq;
		}
		return 1;
	}
	f = f < 0.05f ? 0.05f : (f > 0.95f ? 0.95f : f);
	float u = psel_ndtri(0.5f + 0.5f * f);
	if (fabsf(zq) > 0.4f) u = psel_offcentre_u(zq, f, u);
	const float srob = PSEL_DIV(*hw, u);			// scale of the bulk
	const float p1 = mean - srob * zq;			// where half of the samples are below
	const float sq = PSEL_SQRT(fn);
	const float wr = fmaxf(6.f, 3.0f * PSEL_SQRT(sq));	// half-width of the centre window in ranks
	const float w = wr * srob * 2.5066283f * rfn;
	float rlo = FLT_MAX, rhi = FLT_MAX;
	if (tails) {
		if (mode == 1) {	// percentile: thresholds median * sig (rejection_float.c:31-44), the median within w of p1
			rlo = fabsf(p1) * s0 - 2.f * w; rhi = fabsf(p1) * s1 - 2.f * w;
		} else {		// thresholds sig * sigma; the bulk's sigma is known to ~1.7 / sqrt(n) relative (1 sd)
			const float mf = PSEL_DIV(sq, sq + 6.f);
			rlo = mf * s0 * srob; rhi = mf * s1 * srob;
		}
		rlo = rlo > w ? rlo : w; rhi = rhi > w ? rhi : w;
	}
	if (!(w > 0.f) || !(w < FLT_MAX) || !(fabsf(p1) < FLT_MAX)) return 2;
	pl->p1 = p1; pl->w = w;
	if (rlo >= FLT_MAX || rhi >= FLT_MAX) { pl->cT = p1; pl->rT = FLT_MAX; pl->blo = -FLT_MAX; pl->bhi = FLT_MAX; return 0; }
	const float lo = p1 - rlo, hi = p1 + rhi;
	pl->cT = 0.5f * (lo + hi); pl->rT = 0.5f * (hi - lo);
	const float eps = (fabsf(pl->cT) + pl->rT) * 4e-7f + FLT_MIN;	// a sample the tail test keeps out of the lists
	pl->blo = pl->cT - pl->rT - eps; pl->bhi = pl->cT + pl->rT + eps;	// satisfies |fl(x - cT)| <= rT
	return 0;
}

// every sample is centre (small or degenerate stacks)
SGPU_HD inline void psel_plan_all(PselPlan *pl) {
	pl->p1 = 0.f; pl->w = FLT_MAX; pl->cT = 0.f; pl->rT = FLT_MAX; pl->blo = -FLT_MAX; pl->bhi = FLT_MAX;
}

// after the extraction: do the ranks a rejection can ask the median of lie inside the centre
// window?  (the median of the kept set moves by half the difference of low and high rejections;
// +-3 ranks are provided for, more ends in a hand-over)
SGPU_HD inline bool psel_window_ok(int n, int cl, int cc) {
	int lo = (n - 1) / 2 - 3, hi = n / 2 + 3;
	lo = lo < 0 ? 0 : lo; hi = hi > n - 1 ? n - 1 : hi;
	return cl <= lo && cl + cc - 1 >= hi;
}

// the window moved to where the counts say the median is, and widened
SGPU_HD inline void psel_replan(int n, int cl, int cc, PselPlan *pl) {
	const float fn = (float) n;
	// ranks per unit of value near the window: what the window itself held, else the plan's guess
	const float dens = cc > 4 ? (float) cc / (2.f * pl->w) : 3.f * sqrtf(sqrtf(fn)) / pl->w;
	const float shift = ((float) (n - 1) * 0.5f - ((float) cl + 0.5f * (float) cc)) / dens;
	const float p1 = pl->p1 + shift, w = pl->w * 1.5f + fabsf(shift) * 0.25f;
	float lo = pl->cT - pl->rT, hi = pl->cT + pl->rT;
	if (pl->rT < FLT_MAX) {		// keep the tails outside the centre window
		if (lo > p1 - w) lo = p1 - w;
		if (hi < p1 + w) hi = p1 + w;
		pl->cT = 0.5f * (lo + hi); pl->rT = 0.5f * (hi - lo);
		const float eps = (fabsf(pl->cT) + pl->rT) * 4e-7f + FLT_MIN;
		pl->blo = pl->cT - pl->rT - eps; pl->bhi = pl->cT + pl->rT + eps;
	}
	pl->p1 = p1; pl->w = w;
}

// ---- phase A, group level ------------------------------------------------------------------------
// The control flow of a lane group, shared by the kernel (stack_psel.cuh: samples in registers,
// shuffles inside the passes) and the host replay (tests/helpers/psel_host.cu).  `P` provides
//   any(b)                            true when b holds for any group that runs in lockstep with this one
//   sums(is_mean, &n, &S1, &S2, &mn, &mx)   nulls become absent in mean mode
//   shifted_sums(K, &S1, &S2)         sum (x - K), sum (x - K)^2 of the kept samples
//   count(mean, hw, &c_neg, &c_in)    kept samples below mean / within hw of it
//   extract(on, plan, &below, &over)  fills the lane lists (nothing when !on); kept samples below plan.p1; lists overflowed
//   classify(plan, &cc, &ccb, &nl, &m)  list entries inside the centre window / of those, below plan.p1 / entries below the window / all
template <class P>
SGPU_HD __forceinline__ void psel_group(P &ps, int is_mean, bool tails, int mode, float s0, float s1, PselHdr &h) {	// mode: 0 sigma, 1 percentile
	double S1, S2; float mn, mx; int n;
	ps.sums(is_mean, &n, &S1, &S2, &mn, &mx);
	h.n = n; h.S1 = S1; h.S2 = S2; h.K = 0.0; h.mn = mn; h.mx = mx; h.flags = 0; h.cl = 0; h.nl = 0; h.cc = 0; h.m = 0;
	PselPlan pl; psel_plan_all(&pl);
	bool live = true;
	if (!(S2 < (double) FLT_MAX * (double) FLT_MAX)) { h.flags = PSEL_WEIRD; live = false; }	// a NaN or an infinity among the samples
	else if (n <= 1 || mn == mx) { h.flags = PSEL_TRIVIAL; live = false; }
	// plain sums lose the variance to cancellation when the spread is tiny beside the level: re-sum about the mean
	const double rn = rcp_d(n > 0 ? n : 1);
	const bool ill = live && !((S2 - S1 * S1 * rn) * 1e5 > S2);	// variance * (n - 1) against mean square * n
	if (ps.any(ill)) {
		const double K = (double) (float) (S1 * rn);
		double T1, T2;
		ps.shifted_sums(K, &T1, &T2);
		if (ill) { S1 = T1; S2 = T2; h.K = K; h.S1 = S1; h.S2 = S2; }
	}
	// where to count
	float mean = 0.f, hw = 0.f;
	bool want = live && psel_plan_count(n, S1, S2, h.K, mn, mx, &mean, &hw);
	bool all_centre = live && !want;
#pragma unroll 1
	for (int attempt = 0; attempt < 4; attempt++) {
		if (!ps.any(want)) break;
		int c_neg, c_in;
		ps.count(mean, hw, &c_neg, &c_in);
		if (want) {
			const int r = psel_plan_window(n, &mean, &hw, c_neg, c_in, tails, mode, s0, s1, attempt, &pl);
			if (r == 0) want = false;
			else if (r == 2) { want = false; all_centre = true; }
		}
	}
	if (want) all_centre = true;
	if (all_centre) {
		psel_plan_all(&pl);
		if (n > PSEL_CAP) { h.flags = PSEL_HANDOVER; live = false; }	// too many to sort by insertion
	}
	// extraction, repeated with a moved window when the centre missed the median's ranks
	bool need = live;
#pragma unroll 1
	for (int attempt = 0; attempt < 3; attempt++) {
		if (!ps.any(need)) break;
		int below; bool over;
		ps.extract(need, pl, &below, &over);
		int cc, ccb, nl, m;
		ps.classify(pl, &cc, &ccb, &nl, &m);
		if (need) {
			const int cl = below - ccb;
			if (over || m > PSEL_ROWS - 2) { h.flags = PSEL_HANDOVER; need = false; }
			else if (pl.w >= FLT_MAX || psel_window_ok(n, cl, cc)) { h.cl = cl; h.nl = nl; h.cc = cc; h.m = m; need = false; }
			else psel_replan(n, cl, cc, &pl);
		}
	}
	if (need) h.flags = PSEL_HANDOVER;
	h.p1 = pl.p1; h.w = pl.w; h.blo = pl.blo; h.bhi = pl.bhi;
}

// ---- phase B -------------------------------------------------------------------------------------
// B1 (still the lane group): the m extracted samples of a pixel are sorted into the first m rows of
// its column by value buckets, one per lane — T equal slices of the centre window, the first and the
// last open-ended — each lane inserting the samples of its bucket behind those of the buckets below.
// Whatever the rounding, psel_bucket is monotone in x, so the rows end up globally ascending:
// [lows | centre | highs] = ranks [0, nl), [cl, cl + cc), [n - nh, n) of the stack.
SGPU_HD __forceinline__ int psel_bucket(float x, float p1, float w, int T) {
	if (T == 1) return 0;
	float s = (x - p1 + w) * ((float) T / (2.f * w));
	s = s > 0.f ? s : 0.f;
	s = s < (float) (T - 1) ? s : (float) (T - 1);
	return (int) s;
}

// ascending insertion of x into rows [base, base + have)
template <class Store>
SGPU_HD __forceinline__ void psel_insert(const Store &col, int base, int have, float x) {
	int i = base + have;
	while (i > base) { const float y = col.get(i - 1); if (!(y > x)) break; col.set(i, y); i--; }
	col.set(i, x);
}

// lane t's share of B1; `base`: entries in the buckets below t.  Lists: cnt(u) entries of lane u, entry(u, k).
template <class Store, class Lists>
SGPU_HD __forceinline__ void psel_sort_bucket(const Store &col, const Lists &ls, int T, int t, int base, float p1, float w) {
	int have = 0;
	for (int u = 0; u < T; u++) {
		const int m = ls.cnt(u);
		for (int k = 0; k < m; k++) {
			const float x = ls.entry(u, k);
			if (psel_bucket(x, p1, w, T) == t) psel_insert(col, base, have++, x);
		}
	}
}

// B2 (one thread per pixel): the column as a partially known sorted stack.  Rows [0, nl) are ranks
// [0, nl), rows [nl, nl + cc) ranks [cl, cl + cc), rows [nl + cc, m) ranks [n - nh, n); rows m and
// m + 1 hold stand-ins for ranks nl and n - nh - 1: the minimum / maximum of the stack when that tail
// list is empty (exact), else a bound on the first sample past the list.  Every predicate the
// rejections evaluate is monotone, so a KEPT bound means a kept sample, and psel_pixel refuses any
// result that rejected one.
template <class Store>
struct PartialAcc {
	const Store &col;
	int n, nl, cl, cc, nh;
	float mn;
	bool trivial;
	mutable bool miss;	// a rank that was not extracted was read
	SGPU_HD __forceinline__ float operator()(int i) const {
		if (trivial) return mn;
		int row;
		if ((unsigned) (i - cl) < (unsigned) cc) row = nl + i - cl;
		else if (i < nl) row = i;
		else if (i >= n - nh && i < n) row = nl + cc + i - (n - nh);
		else if (i == nl) row = nl + cc + nh;
		else if (i == n - nh - 1) row = nl + cc + nh + 1;
		else { miss = true; row = nl + cc + nh + (i < cl ? 0 : 1); }
		return col.get(row);
	}
	SGPU_HD __forceinline__ void set(int, float) const { miss = true; }	// no rejection here compacts in place
};

// 1 / i: from a table of correctly rounded reciprocals when there is one (device: a load instead of
// MUFU.RCP64H and its Newton steps), else rcp_d — the same value either way
SGPU_HD __forceinline__ double psel_rcp(const double *tab, int i) {
#ifdef __CUDA_ARCH__
	if (tab) return __ldg(tab + i);
#endif
	(void) tab;
	return rcp_d(i);
}

// sd_interval (sorted.cuh) for this engine's B2: one square root; the float-mean term of the upper
// bound (sum (x - m~)^2 = sum (x - m)^2 + n (m - m~)^2) is folded into the relative margin unless it is
// a noticeable part of the variance (levels thousands of sigmas above zero), which takes a second root
SGPU_HD __forceinline__ bool psel_sd_bounds(const Sums &S, int nn, double rn, double rn1, float *lo, float *hi) {
	const double var = (S.S2 - S.S1 * S.S1 * rn) * rn1;
	if (!(var >= 0.0)) return false;
	const double e = fabs(S.K + S.S1 * rn) * 5.9604645e-08;	// >= half an ulp of the float mean
	const double t = e * e * ((double) nn * rn1);
	const float s = sqrtf((float) var);
	float sh = s;
	if (t > var * 1e-7) sh = sqrtf((float) (var + t));
	*lo = s * (1.f - SGPU_DELTA_SD); *hi = sh * (1.f + (SGPU_DELTA_SD + 1e-7f));
	return *hi < FLT_MAX;
}

// one iteration of the reference's loop (rejection_float.c:174-209) from the state "s.a lows and s.bh highs
// rejected, sums S1 / S2 of the kept": what it rejects, and whether the loop goes on
enum { SIG_STOP = 0, SIG_LAST = 1, SIG_CONT = 2, SIG_FAIL = 3 };	// nothing rejected (state final) | state after it final | another iteration | hand over
struct SigState { int status, a, bh; double S1, S2; };
struct SigCtx {
	int n, nl, cc, nh, m, cl;
	float lstand, hstand, s0, s1;
	double K;
	const double *rcp_tab;
};

template <bool U16, class Store>
SGPU_HD __forceinline__ SigState psel_sigma_step(const Store &col, const SigCtx &x, SigState s) {
	s.status = SIG_STOP;
	const int a = s.a, bh = s.bh, nn = x.n - a - bh, r = a + bh;
	if (nn - r <= 4) return s;		// the guard refuses every rejection of this pass
	// the next two candidates of each side and the median's rows, all loaded before the arithmetic
	const float l0 = a < x.nl ? col.get(a) : x.lstand, l1 = a + 1 < x.nl ? col.get(a + 1) : x.lstand;
	const float h0 = bh < x.nh ? col.get(x.m - 1 - bh) : x.hstand, h1 = bh + 1 < x.nh ? col.get(x.m - 2 - bh) : x.hstand;
	bool miss = false;
	auto cen = [&](int rank) { int q = rank - x.cl; if ((unsigned) q >= (unsigned) x.cc) { miss = true; q = 0; } return col.get(x.nl + q); };
	const float medf = (float) median_sorted<U16>(cen, a, nn);
	Sums S; S.K = x.K; S.S1 = s.S1; S.S2 = s.S2;
	float slo, shi;
	s.status = SIG_FAIL;
	if (!psel_sd_bounds(S, nn, psel_rcp(x.rcp_tab, nn), psel_rcp(x.rcp_tab, nn - 1), &slo, &shi) || miss) return s;
	const float tl = shi * x.s0, th = shi * x.s1;	// sigma_clipping_float: `sigma * sigmalow`
	int a2 = a, b2 = bh;
	float vl = l0, vh = h0;		// V_low[a2], V_high[b2]
	while (medf - vl > tl) {
		if (++a2 > x.nl) return s;	// the stand-in went: samples that were never extracted are in play
		vl = a2 == a + 1 ? l1 : (a2 < x.nl ? col.get(a2) : x.lstand);
	}
	while (vh - medf > th) {
		if (++b2 > x.nh) return s;
		vh = b2 == bh + 1 ? h1 : (b2 < x.nh ? col.get(x.m - 1 - b2) : x.hstand);
	}
	if (medf - vl > slo * x.s0 || vh - medf > slo * x.s1) return s;	// the low end of the interval rejects more
	const int K = (a2 - a) + (b2 - bh);
	if (K == 0) { s.status = SIG_STOP; return s; }
	if (nn - r - K < 4) return s;	// the guard would cut this pass short: order dependent
	if (a2 > a) {
		S.sub(l0);
		if (a2 > a + 1) {
			S.sub(l1);
#pragma unroll 1
			for (int i = a + 2; i < a2; i++) S.sub(col.get(i));
		}
	}
	if (b2 > bh) {
		S.sub(h0);
		if (b2 > bh + 1) {
			S.sub(h1);
#pragma unroll 1
			for (int i = bh + 2; i < b2; i++) S.sub(col.get(x.m - 1 - i));
		}
	}
	s.a = a2; s.bh = b2; s.S1 = S.S1; s.S2 = S.S2;
	s.status = (x.n - a2 - b2 <= 3) ? SIG_LAST : SIG_CONT;
	return s;
}

// Sigma clipping (fast_sigma, sorted.cuh) restated in ROW space, with the dependent chain cut short: B2 is
// latency bound — its pixels in flight are fixed by the shared memory a tile takes.  The clip walks run over
// V_low = [lows..., stand-in] and V_high = [highs (descending)..., stand-in] by row index, medians are centre
// rows, reciprocals come from a table, a rejected stand-in is a hand-over at once.  Decisions and arithmetic
// are fast_sigma's (same sd interval up to its margin, same float predicates, same down-dated sums);
// tests/test_psel_host.py runs it against the oracle.  (Running the next pass speculatively on the group's other
// lanes — one hypothesis per lane about what this pass rejects — was tried: the lanes diverge and it was slower.)
template <bool U16, class Store>
SGPU_HD __forceinline__ void psel_sigma(const Store &col, const PselHdr &h, int nl, int cc, int nh, const RejCfg &c, const double *rcp_tab,
		int rej[2], FastRes &o) {
	o.status = FAST_FALLBACK;
	SigCtx x;
	x.n = h.n; x.nl = nl; x.cc = cc; x.nh = nh; x.m = nl + cc + nh; x.cl = h.cl;
	x.lstand = nl ? h.blo : h.mn; x.hstand = nh ? h.bhi : h.mx; x.s0 = c.s0; x.s1 = c.s1; x.K = h.K; x.rcp_tab = rcp_tab;
	{
		bool miss = false;
		auto cen = [&](int rank) { int q = rank - x.cl; if ((unsigned) q >= (unsigned) cc) { miss = true; q = 0; } return col.get(nl + q); };
		const double median = median_sorted<U16>(cen, 0, x.n);
		if (miss || (U16 ? ((float) median == 0.f) : (median == 0.0))) return;
	}
	SigState s; s.status = SIG_CONT; s.a = 0; s.bh = 0; s.S1 = h.S1; s.S2 = h.S2;
#pragma unroll 1
	for (int pass = 0; s.status == SIG_CONT; pass++) {
		if (pass == 64) return;	// far more passes than a sane stack takes
		s = psel_sigma_step<U16>(col, x, s);
	}
	if (s.status == SIG_FAIL) return;
	rej[0] += s.a; rej[1] += s.bh;
	Sums S; S.K = x.K; S.S1 = s.S1; S.S2 = s.S2;
	o.status = FAST_DONE; o.kept = x.n - s.a - s.bh; o.sum = S.sum(o.kept);
	o.pmin = s.a < nl ? col.get(s.a) : x.lstand; o.pmax = s.bh < nh ? col.get(x.m - 1 - s.bh) : x.hstand;
}

// Store: get(row) / set(row, v) on the pixel's column: m + 2 rows, the first m sorted by B1
template <bool U16, class Store, typename OrigFn>
SGPU_HD double psel_pixel(const Store &col, const PselHdr &h, int N, int is_mean, const RejCfg &c,
		const double *wts, OrigFn orig, int rej[2], int *status, const double *rcp_tab = nullptr) {
	*status = FAST_FALLBACK;
	if (h.flags & (PSEL_WEIRD | PSEL_HANDOVER)) return 0.0;
	const int n = h.n;
	int nl = 0, nh = 0, cc = 0;
	if (!(h.flags & PSEL_TRIVIAL)) {
		nl = h.nl; cc = h.cc; nh = h.m - nl - cc;
		if (nl > PSEL_CAPT || cc > PSEL_CAP || nh > PSEL_CAPT || nh < 0 || h.cl < nl || h.cl + cc > n - nh) return 0.0;
		col.set(h.m, nl ? h.blo : h.mn);
		col.set(h.m + 1, nh ? h.bhi : h.mx);
	}
	const PartialAcc<Store> acc{ col, n, nl, h.cl, cc, nh, h.mn, (h.flags & PSEL_TRIVIAL) != 0, false };
	double result;
	int st = FAST_DONE;
	if (is_mean) {
		Sums S0; S0.K = h.K; S0.S1 = h.S1; S0.S2 = h.S2;
		FastRes o;
		result = 0.0;
		if (n > 0) {
			// only what this engine serves is instantiated (the others stay with the lane-sort engine)
			o.status = FAST_DONE; o.kept = n; o.sum = 0.0; o.pmin = o.pmax = 0.f;
			if (n == 1 || c.type == SGPU_NO_REJEC) fast_all(acc, n, &S0, o);
			else if (c.type == SGPU_PERCENTILE) fast_percentile<U16>(acc, n, c, &S0, rej, o);
			else if (c.type == SGPU_SIGMA && !acc.trivial) psel_sigma<U16>(col, h, nl, cc, nh, c, rcp_tab, rej, o);
			else if (c.type == SGPU_SIGMA) fast_sigma<U16>(acc, n, c, &S0, rej, o);
			else o.status = FAST_FALLBACK;
			if (o.status != FAST_DONE || o.kept == 0) o.status = FAST_FALLBACK;
			st = o.status;
			// a stand-in is fine as long as it was KEPT (every rejected rank must be a real sample)
			if (acc.miss || rej[0] > nl || rej[1] > nh) st = FAST_FALLBACK;
			if (st == FAST_DONE) {
				const int a = rej[0], b = n - rej[1];		// the kept ranks [a, b)
				const bool lo_b = a > 0 && a == nl, hi_b = b < n && b == n - nh;	// a whole tail list went: its stand-in became an extreme
				if (wts && (lo_b || hi_b) && !acc.trivial) {
					// the weighted combine re-selects by value between the extremes of the kept set
					// (median_and_mean.c:1045-1070); where a bound stood in for one of them, find it
					const float L = lo_b ? col.get(a - 1) : 0.f, Hh = hi_b ? col.get(nl + cc) : 0.f;	// the innermost rejected samples
					float pmin = FLT_MAX, pmax = -FLT_MAX;
					for (int f = 0; f < N; f++) {
						const float v = orig(f);
						if (v == 0.f) continue;
						if (lo_b && v > L && v < pmin) pmin = v;
						if (hi_b && v < Hh && v > pmax) pmax = v;
					}
					if (lo_b) o.pmin = pmin;
					if (hi_b) o.pmax = pmax;
				}
				result = fast_combine(o, N, wts, orig);
			}
		}
	} else {
		result = median_sorted<U16>(acc, 0, n);	// all N samples, zeros included (median_and_mean.c:1711-1715)
		if (acc.miss) st = FAST_FALLBACK;
	}
	if (st != FAST_DONE) { rej[0] = rej[1] = 0; return 0.0; }
	*status = FAST_DONE;
	return result;
}

}  // namespace sgpu
