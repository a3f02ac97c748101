// reg.cuh — per-pixel logic of the REGISTER engine (kernel: stack_reg.cuh): mean with sigma /
// winsorized / percentile / no rejection and the median stack on float samples, WITHOUT sorting
// the pixel stack.
//
// A group of T lanes owns one pixel and keeps its N samples in registers (RE per lane, NaN in
// unused slots).  Everything the reference computes per clip pass is obtained from passes over
// those registers that cost a few instructions per sample:
//   sums     sum(x-K), sum((x-K)^2) in double about a sample K from the bulk: the real-number sd,
//            which the reference's float sd (siril_stats_float_sd, statistics.h:82-107) matches to
//            a few 1e-7 relative;
//   median   quickmedian_float's value (sorting.c:240-273) is the middle order statistic(s).  One
//            counting pass with a bracket mean +- 3.5 sd/sqrt(n) parks the samples inside the
//            bracket in lane-private lists (the pixel's own column of the shared-memory tile),
//            interpolation narrows the bracket on those short lists until <= 24 samples are left,
//            and a rank count sorts them: a WINDOW of consecutive order statistics around the
//            median with their absolute ranks.  Clip passes only remove extreme ranks (the
//            reference's predicates are monotone in the sample value), so later medians are
//            look-ups in the same window;
//   clip     one pass per iteration: a sample is a candidate if the low end of the sd interval
//            rejects it; candidates (rare) are decided with both ends of the interval exactly as
//            sorted.cuh does, removed from the running sums and overwritten with NaN, which no
//            later comparison selects.
// Decisions are identical to the reference's or the pixel is handed over (FAST_FALLBACK) to the
// exact engine: same protocol, same interval argument as sorted.cuh.  Pixels with null (zero) or non-finite samples, the
// guard regime `N - r <= 4` and medians that fall out of the window twice are handed over too.
//
// __host__ __device__ with T == 1 on the host: tests/helpers/reg_host.cu runs the same code on
// the CPU against the oracle.
#pragma once
#include "sorted.cuh"

namespace sgpu {

#ifdef SGPU_FAST_DEBUG	// host-side test builds only: the FIRST reason a stack was handed over
static int g_reg_reason = 0, g_reg_exact = 0, g_reg_inner = 0, g_reg_passes = 0, g_reg_rebuilt = 0;
SGPU_HD inline void reg_why(int x) {
#ifndef __CUDA_ARCH__
	if (!g_reg_reason) g_reg_reason = x;
#endif
}
#define REG_WHY(x) reg_why(x)
#else
#define REG_WHY(x) ((void) 0)
#endif

#define REG_WCAP 24		// capacity of the sorted window
// half-width of the interval put around the float sd the reference would compute, relative to the
// real-number sd: its float deviations squared in float (3 roundings, 1.8e-7 on the variance), the
// float conversion, sqrtf (and the 1.134f product), plus our own sqrtf: < 4e-7 in all.  The
// winsorized sigma also inherits 1.6x the error of the clamp bounds it was iterated from.
#define REG_DELTA_SD 6.0e-7f
#define REG_DELTA_WS 1.5e-6f

// ---- the T lanes of a pixel ----------------------------------------------------------------------
template <int T>
struct Group {
	unsigned mask;	// the group's lanes in the warp
	int t;		// this lane's index in the group
	SGPU_HD __forceinline__ int sum(int x) const {
#ifdef __CUDA_ARCH__
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x += __shfl_xor_sync(mask, x, m);
#endif
		return x;
	}
	SGPU_HD __forceinline__ double sum(double x) const {
#ifdef __CUDA_ARCH__
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x += __shfl_xor_sync(mask, x, m);
#endif
		return x;
	}
	SGPU_HD __forceinline__ int any(int x) const {
#ifdef __CUDA_ARCH__
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x |= __shfl_xor_sync(mask, x, m);
#endif
		return x;
	}
	SGPU_HD __forceinline__ float minf(float x) const {
#ifdef __CUDA_ARCH__
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x = fminf(x, __shfl_xor_sync(mask, x, m));
#endif
		return x;
	}
	SGPU_HD __forceinline__ float maxf(float x) const {
#ifdef __CUDA_ARCH__
#pragma unroll
		for (int m = 1; m < T; m <<= 1) x = fmaxf(x, __shfl_xor_sync(mask, x, m));
#endif
		return x;
	}
	// exclusive prefix sum over the lanes of the group
	SGPU_HD __forceinline__ int exscan(int x) const {
#ifdef __CUDA_ARCH__
		int incl = x;
#pragma unroll
		for (int m = 1; m < T; m <<= 1) { const int y = __shfl_up_sync(mask, incl, m, T); if (t >= m) incl += y; }
		return incl - x;
#else
		return 0;
#endif
	}
	SGPU_HD __forceinline__ float bcast(float x, int src) const {
#ifdef __CUDA_ARCH__
		return __shfl_sync(mask, x, src, T);
#else
		return x;
#endif
	}
	SGPU_HD __forceinline__ void sync() const {
#ifdef __CUDA_ARCH__
		__syncwarp(mask);
#endif
	}
};

SGPU_HD __forceinline__ float reg_nan() {
#ifdef __CUDA_ARCH__
	return __int_as_float(0x7fc00000);
#else
	union { unsigned u; float f; } c; c.u = 0x7fc00000u; return c.f;
#endif
}

// ---- sums -----------------------------------------------------------------------------------------
// sum(x-K), sum((x-K)^2) in double over the samples still in the registers.  ALL: every slot may
// hold NaN (after rejections); otherwise only the slots of rows >= RF can (padding), tested by index.
template <int T, int RE, int RF, bool ALL>
SGPU_HD __forceinline__ void reg_sums(const float (&v)[RE], const Group<T> &g, int N, float K, double *S1, double *S2) {
	double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
	const double Kd = (double) K;
#pragma unroll
	for (int j = 0; j < RE; j++) {
		double y = (double) v[j] - Kd;	// exact
		if (ALL) { if (!(v[j] == v[j])) y = 0.0; }
		else if (j >= RF) { if (!(j * T + g.t < N)) y = 0.0; }
		if (j & 1) { b1 += y; b2 = fma(y, y, b2); } else { a1 += y; a2 = fma(y, y, a2); }
	}
	*S1 = g.sum(a1 + b1); *S2 = g.sum(a2 + b2);
}

// ---- the window of order statistics around the median ---------------------------------------------
struct RegWin {
	int R0;		// rank (among the samples in the registers when it was built) of sw[0]
	int cnt;	// sw[0..cnt) ascending
	int a0;		// how many low-end rejections had happened when it was built
	int flat;	// all cnt entries equal fv (heavy ties): sw[] is not filled
	float fv;
};

// Storage `St`: put(e, x) / get(e) — entry e of this lane's private list; w(i) — unsorted window
// buffer, sw(i) — sorted window, both shared by the group, REG_WCAP entries each.

// count pass over the lane-private lists: entries below L, entries in [L, U) and their range
template <int T, class St>
SGPU_HD __forceinline__ void list_count(const St &st, int cnt, float L, float U, int *lo, int *in, float *mn, float *mx) {
	int l = 0, i = 0;
	float a = FLT_MAX, b = -FLT_MAX;
	for (int e = 0; e < cnt; e++) {
		const float x = st.get(e);
		const bool inside = (x >= L) && (x < U);
		l += x < L; i += inside;
		if (inside) { a = fminf(a, x); b = fmaxf(b, x); }
	}
	*lo = l; *in = i; *mn = a; *mx = b;
}
// keep the entries in [L, U) (in place, order preserved)
template <int T, class St>
SGPU_HD __forceinline__ int list_filter(St &st, int cnt, float L, float U) {
	int c = 0;
	for (int e = 0; e < cnt; e++) {
		const float x = st.get(e);
		if ((x >= L) && (x < U)) { st.put(c, x); c++; }
	}
	return c;
}

// Build the window for the n samples (non-NaN) in the registers: the order statistics around
// ranks klo..khi (the median position(s)); mean / sd locate the first bracket.  false: not found.
template <int T, int RE, class St>
SGPU_HD bool reg_window(const float (&v)[RE], const Group<T> &g, int n, float mean, float sd, St &st, RegWin &win) {
	const int khi = n >> 1, klo = (n & 1) ? khi : khi - 1;
	float h = 3.5f * sd / sqrtf((float) n);
	if (!(h < FLT_MAX)) { REG_WHY(101); return false; }
	if (!(h > 0.f)) h = fabsf(mean) * 1e-6f + FLT_MIN;	// all samples equal (or nearly): any bracket around them
	float L = mean - h, U = mean + h;
	int cnt = 0, Clo = 0, Cin = 0;
	// level 1: registers -> lane-private lists
#pragma unroll 1
	for (int attempt = 0; ; attempt++) {
		int lo = 0; cnt = 0;
#pragma unroll
		for (int j = 0; j < RE; j++) {
			const float x = v[j];
			const bool below = x < L;
			lo += below;
			if (!below && x < U) { st.put(cnt, x); cnt++; }
		}
		Clo = g.sum(lo); Cin = g.sum(cnt);
		if (Clo <= klo && Clo + Cin > khi) break;
		if (attempt == 2) { REG_WHY(102); return false; }
		const float step = (attempt ? 40.f : 6.f) * h;
		if (Clo > klo) { U = L; L = L - step; } else if (Clo + Cin <= khi) { L = U; U = U + step; } else { L -= step; U += step; }
	}
	// level 2: narrow the bracket on the lists until the window fits (or is a run of one value).
	// The next guess inverts a piecewise-linear model of the rank-vs-value curve through the
	// bracket ends and one interior knot (Mv, Mr) learnt from the previous count.
	int margin = 8, Mr = -1;
	float Mv = 0.f;
	win.flat = 0;
	for (int tries = 0; Cin > REG_WCAP; tries++) {
		if (tries == 12) { REG_WHY(103); return false; }
		float L2, U2;
		{
			const float want[2] = { (float) (klo - margin) - 0.5f, (float) (khi + margin) + 1.5f };
			float got[2];
#pragma unroll
			for (int q = 0; q < 2; q++) {
				float r0 = (float) Clo, v0 = L, r1 = (float) (Clo + Cin), v1 = U;
				if (Mr >= 0) { if (want[q] <= (float) Mr) { r1 = (float) Mr; v1 = Mv; } else { r0 = (float) Mr; v0 = Mv; } }
				float f = r1 > r0 ? (want[q] - r0) / (r1 - r0) : 1.f;
				f = f < 0.f ? 0.f : (f > 1.f ? 1.f : f);
				got[q] = f <= 0.f ? v0 : (f >= 1.f ? v1 : v0 + f * (v1 - v0));
			}
			L2 = got[0]; U2 = got[1];
		}
		if (!(L2 < U2) || !(L2 >= L) || !(U2 <= U)) { L2 = L; U2 = U; }
		const bool whole = (L2 == L && U2 == U);	// the model cannot split the bracket: look at the range of what is in it
		int lo2, in2; float mn, mx;
		list_count<T>(st, cnt, L2, U2, &lo2, &in2, &mn, &mx);
		lo2 = g.sum(lo2); in2 = g.sum(in2); mn = g.minf(mn); mx = g.maxf(mx);
		const int b0 = Clo + lo2, b1 = b0 + in2;	// ranks [b0, b1) fall inside the guess
#ifdef REG_TRACE
		printf("try %d: L %.8g U %.8g Clo %d Cin %d margin %d knot %.8g/%d | L2 %.8g U2 %.8g lo2 %d in2 %d mn %.8g mx %.8g\n", tries, L, U, Clo, Cin, margin, Mv, Mr, L2, U2, lo2, in2, mn, mx);
#endif
		if (khi < b0) { U = L2; Cin = lo2; if (!(Mr >= 0 && Mv < U)) Mr = -1; }			// the median lies below the guess
		else if (klo >= b1) { L = U2; Cin -= lo2 + in2; Clo = b1; if (!(Mr >= 0 && Mv > L)) Mr = -1; }	// above it
		else if (klo < b0) { U = U2; Cin = lo2 + in2; Mv = L2; Mr = b0; }				// across its lower edge
		else if (khi >= b1) { L = L2; Cin -= lo2; Clo = b0; Mv = U2; Mr = b1; }			// across its upper edge
		else if (mn == mx) { win.flat = 1; win.fv = mn; Clo = b0; Cin = in2; break; }		// a run of equal samples
		else if (whole) {	// split between the extremes of what is left
			const float mid = mn + (mx - mn) * 0.5f;
			if (margin > 1) { margin >>= 1; continue; }
			int lom, inm; float d0, d1;
			list_count<T>(st, cnt, L, mid > mn ? mid : mx, &lom, &inm, &d0, &d1);
			inm = g.sum(inm);
			if (khi < Clo + inm) { U = mid > mn ? mid : mx; Cin = inm; }
			else if (klo >= Clo + inm) { L = mid > mn ? mid : mx; Clo += inm; Cin -= inm; }
			else { REG_WHY(104); return false; }
			Mr = -1;
		}
		else { L = L2; U = U2; Clo = b0; Cin = in2; Mr = -1; if (margin > 1) margin >>= 1; }
		if (Mr >= 0 && !(Mv > L && Mv < U)) Mr = -1;
		cnt = list_filter<T>(st, cnt, L, U);
	}
	win.R0 = Clo; win.cnt = Cin;
	if (win.flat) return true;
	// gather the lists into the window buffer, then sort it by rank counting
	const int off = g.exscan(cnt);
	for (int e = 0; e < cnt; e++) st.w(off + e) = st.get(e);
	g.sync();
	{	// every lane ranks its first two window entries in one sweep
		const int i0 = g.t, i1 = g.t + T;
		const float x0 = i0 < Cin ? st.w(i0) : 0.f, x1 = i1 < Cin ? st.w(i1) : 0.f;
		int r0 = 0, r1 = 0;
		for (int q = 0; q < Cin; q++) {
			const float y = st.w(q);
			r0 += (y < x0) || (y == x0 && q < i0);
			r1 += (y < x1) || (y == x1 && q < i1);
		}
		if (i0 < Cin) st.sw(r0) = x0;
		if (i1 < Cin) st.sw(r1) = x1;
		for (int i = g.t + 2 * T; i < Cin; i += T) {	// host build (T == 1) and wider windows
			const float x = st.w(i);
			int r = 0;
			for (int q = 0; q < Cin; q++) { const float y = st.w(q); r += (y < x) || (y == x && q < i); }
			st.sw(r) = x;
		}
	}
	g.sync();
	return true;
}

// median of the kept rank range [a, a + nn) (ranks counted as when the window was built, plus the
// low-end rejections since); false: outside the window
template <class St>
SGPU_HD __forceinline__ bool win_median(const St &st, const RegWin &win, int a, int nn, double *med) {
	const int base = a - win.a0 - win.R0;
#ifdef REG_TRACE
	printf("win_median: a %d nn %d a0 %d R0 %d cnt %d\n", a, nn, win.a0, win.R0, win.cnt);
#endif
	const int k = nn >> 1;
	const int ihi = base + k, ilo = (nn & 1) ? ihi : ihi - 1;
	if (ilo < 0 || ihi >= win.cnt) return false;
	if (win.flat) { *med = (nn < 9 && !(nn & 1)) ? (win.fv + win.fv) / 2.0 : (double) win.fv; return true; }
	if (nn & 1) { *med = (double) st.sw(ihi); return true; }
	const float x = st.sw(ilo), y = st.sw(ihi);
	// sorting.c:366-415: below 9 samples the sorting network adds the middles in float
	*med = (nn < 9) ? (x + y) / 2.0 : ((double) x + (double) y) / 2.0;
	return true;
}

// ---- the engine -----------------------------------------------------------------------------------
struct RegState {
	double S1, S2;	// running sums about K over the kept samples
	float K;
	int nanl;	// NaN slots of THIS lane (padding + rejected)
};

SGPU_HD __forceinline__ double reg_var(const RegState &s, int n) { return (s.S2 - s.S1 * s.S1 * rcp_d(n)) * rcp_d(n - 1); }
SGPU_HD __forceinline__ double reg_mean(const RegState &s, int n) { return (double) s.K + s.S1 * rcp_d(n); }

// centre for the sums: the median of three samples of the group's first lane (a sample from the
// bulk keeps x - K exact or nearly so), then the sums; false: non-finite samples
template <int T, int RE, int RF>
SGPU_HD __forceinline__ bool reg_init(const float (&v)[RE], const Group<T> &g, int N, RegState &s) {
	const float x0 = v[0], x1 = v[1], x2 = v[2];
	s.K = g.bcast(fmaxf(fminf(x0, x1), fminf(fmaxf(x0, x1), x2)), 0);
	reg_sums<T, RE, RF, false>(v, g, N, s.K, &s.S1, &s.S2);
	const int mine = (N - g.t + T - 1) / T;	// samples this lane holds
	s.nanl = RE - (mine < RE ? mine : RE);
	return s.S2 < 1e300;
}

// one clip pass: a sample is a candidate if the low end of the sd interval rejects it; candidates
// are decided with both ends, taken out of the sums and overwritten with NaN.  false: a sample
// sits between the two ends (cannot be decided here).
template <int T, int RE>
SGPU_HD __forceinline__ bool reg_clip(float (&v)[RE], const Group<T> &g, float medf, float tl_lo, float tl_hi, float th_lo, float th_hi,
		float K, int *nlo, int *nhi, double *rs1, double *rs2) {
	int lo = 0, hi = 0, amb = 0;
	double r1 = 0.0, r2 = 0.0;
	const float nth = -th_lo;
	// screen without branches first: most passes of most pixels have no candidate at all
	bool any = false;
#pragma unroll
	for (int j = 0; j < RE; j++) {
		const float d = medf - v[j];
		any |= (d > tl_lo) | (d < nth);
	}
	if (g.any(any)) {
#pragma unroll
		for (int j = 0; j < RE; j++) {
			const float d = medf - v[j];	// `median - pixel`; `pixel - median` is its exact negation
			if (d > tl_lo || d < nth) {
				if (d > tl_hi) lo++;
				else if (-d > th_hi) hi++;
				else amb = 1;
				const double y = (double) v[j] - (double) K;
				r1 += y; r2 = fma(y, y, r2);
				v[j] = reg_nan();
			}
		}
	}
	*nlo = lo; *nhi = hi; *rs1 = r1; *rs2 = r2;
	return !g.any(amb);
}

// The winsorized sigma (rejection_float.c:223-237) of the n samples in the registers, as an interval
// [*slo, *shi] that holds the reference's float value.  Where the reference's convergence test is
// too close to call, the iteration goes on and the sigma it would have stopped with is kept inside
// the interval (the clip pass hands the pixel over only if a sample falls between the two ends).
// returns 0 ok, 1 hand the pixel over.
template <int T, int RE>
SGPU_HD int reg_winsor_sigma(const float (&v)[RE], const Group<T> &g, int n, const RegState &s, float medf, float *slo, float *shi) {
	double varw = reg_var(s, n), meanw = 0.0;
	if (!(varw >= 0.0)) { REG_WHY(106); return 1; }
	float sigma = sqrtf((float) varw), sigma0;	// `sigma = sd(stack)`, before the first clamp
	float cl = -FLT_MAX, cu = FLT_MAX, clo = FLT_MAX, chi = 0.f;
	const double rn = rcp_d(n), rn1 = rcp_d(n - 1);
	for (int it = 0; ; it++) {
		const float m0 = medf - 1.5f * sigma, m1 = medf + 1.5f * sigma;
		const float ncl = clampw(cl, m0, m1), ncu = clampw(cu, m0, m1);
		if (ncl < cl || ncu > cu) { REG_WHY(107); return 1; }	// clamp bounds only ever tighten with sane data
		cl = ncl; cu = ncu;
		// sums of the clamped samples about medf; a NaN slot clamps to cl and is taken out again below
		double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
#pragma unroll
		for (int j = 0; j < RE; j++) {
			const double y = (double) (fminf(fmaxf(v[j], cl), cu) - medf);
			if (j & 1) { b1 += y; b2 = fma(y, y, b2); } else { a1 += y; a2 = fma(y, y, a2); }
		}
		const double yc = (double) (cl - medf), nn = (double) s.nanl;
		const double W1 = g.sum((a1 + b1) - nn * yc), W2 = g.sum((a2 + b2) - nn * yc * yc);
		varw = (W2 - W1 * W1 * rn) * rn1;
		meanw = (double) medf + W1 * rn;
		if (!(varw >= 0.0)) { REG_WHY(108); return 1; }
		sigma0 = sigma;
		sigma = 1.134f * sqrtf((float) varw);
#if defined(SGPU_FAST_DEBUG) && !defined(__CUDA_ARCH__)
		g_reg_inner++;
#endif
		const float d = fabsf(sigma - sigma0), thr = sigma0 * 0.0005f;
		if (it > 64) { REG_WHY(109); return 1; }
		if (fabsf(d - thr) <= 2.f * REG_DELTA_WS * sigma0) {	// the reference may stop here or go on
#if defined(SGPU_FAST_DEBUG) && !defined(__CUDA_ARCH__)
			g_reg_exact++;
#endif
			float l, h;
			if (!sd_interval(varw, meanw, n, REG_DELTA_WS, 1.134f, &l, &h)) { REG_WHY(105); return 1; }
			clo = l < clo ? l : clo; chi = h > chi ? h : chi;
			continue;
		}
		if (!(d > thr)) break;
	}
	if (!sd_interval(varw, meanw, n, REG_DELTA_WS, 1.134f, slo, shi)) return 1;
	*slo = clo < *slo ? clo : *slo; *shi = chi > *shi ? chi : *shi;
	return 0;
}

// Mean with rejection of the N samples in the registers (no nulls among them: the kernel hands
// pixels with zeros over).  Types: NO_REJEC, PERCENTILE, SIGMA, WINSORIZED.
template <int T, int RE, int RF, bool WINSOR, class St>
SGPU_HD int reg_mean_and_reject(float (&v)[RE], const Group<T> &g, int N, const RejCfg &c, St &st, int rej[2], double *result) {
	RegState s;
	int n = N;
	if (!reg_init<T, RE, RF>(v, g, N, s)) { REG_WHY(110); return FAST_FALLBACK; }
	int a = 0, r = 0, built = 0;
	RegWin win; win.a0 = 0; win.cnt = 0; win.R0 = 0; win.flat = 0; win.fv = 0.f;
	while (c.type != SGPU_NO_REJEC) {
		if (c.type != SGPU_PERCENTILE && n - r <= 4) break;	// the guard refuses every rejection of this pass
		double median;
		if (!built || !win_median(st, win, a, n, &median)) {
			if (built == 7) { REG_WHY(116); return FAST_FALLBACK; }
#if defined(SGPU_FAST_DEBUG) && !defined(__CUDA_ARCH__)
			g_reg_rebuilt += built > 0;
#endif
			built++;
			const double var0 = reg_var(s, n);
			if (!(var0 >= 0.0)) { REG_WHY(117); return FAST_FALLBACK; }
			if (!reg_window<T, RE>(v, g, n, (float) reg_mean(s, n), sqrtf((float) var0), st, win)) { REG_WHY(118); return FAST_FALLBACK; }
			win.a0 = a;
			if (!win_median(st, win, a, n, &median)) { REG_WHY(119); return FAST_FALLBACK; }
		}
		const float medf = (float) median;
		float tl_lo, tl_hi, th_lo, th_hi;
		if (c.type == SGPU_PERCENTILE) {	// rejection_float.c:31-44, 160-173: one pass, exact float predicates
			if (median == 0.0) { REG_WHY(114); return FAST_FALLBACK; }
			tl_lo = tl_hi = medf * c.s0; th_lo = th_hi = medf * c.s1;
		} else {
			float slo, shi;
			if (!WINSOR || c.type == SGPU_SIGMA) {
				if (c.type != SGPU_SIGMA) { REG_WHY(128); return FAST_FALLBACK; }	// a kernel built without the winsorized iteration
				if (!sd_interval(reg_var(s, n), reg_mean(s, n), n, REG_DELTA_SD, 1.f, &slo, &shi)) { REG_WHY(120); return FAST_FALLBACK; }
			} else {
				const int rv = reg_winsor_sigma<T, RE>(v, g, n, s, medf, &slo, &shi);
				if (rv) { REG_WHY(121); return FAST_FALLBACK; }
			}
			tl_lo = slo * c.s0; tl_hi = shi * c.s0; th_lo = slo * c.s1; th_hi = shi * c.s1;
		}
#if defined(SGPU_FAST_DEBUG) && !defined(__CUDA_ARCH__)
		g_reg_passes++;
#endif
		int nlo, nhi; double rs1, rs2;
		if (!reg_clip<T, RE>(v, g, medf, tl_lo, tl_hi, th_lo, th_hi, s.K, &nlo, &nhi, &rs1, &rs2)) { REG_WHY(122); return FAST_FALLBACK; }
		s.nanl += nlo + nhi;
		nlo = g.sum(nlo); nhi = g.sum(nhi);
		const int Kr = nlo + nhi;
		if (c.type == SGPU_PERCENTILE) {
			if (n - Kr <= 0) { REG_WHY(115); return FAST_FALLBACK; }	// kept == 0: median of the compacted stack
			s.S1 -= g.sum(rs1);
			rej[0] += nlo; rej[1] += nhi; n -= Kr;
			break;
		}
		if (Kr == 0) break;
		if (n - r - Kr < 4) { REG_WHY(123); return FAST_FALLBACK; }	// the guard would cut this pass short: order dependent
		s.S1 -= g.sum(rs1); s.S2 -= g.sum(rs2);
		rej[0] += nlo; rej[1] += nhi; r += Kr; a += nlo; n -= Kr;
		if (n <= 3) break;
	}
	*result = (s.S1 + (double) n * (double) s.K) / (double) n;
	return FAST_DONE;
}

// median stack (median_and_mean.c:1711-1715): all N samples, zeros included
template <int T, int RE, int RF, class St>
SGPU_HD int reg_median(const float (&v)[RE], const Group<T> &g, int N, St &st, double *result) {
	RegState s;
	if (!reg_init<T, RE, RF>(v, g, N, s)) { REG_WHY(124); return FAST_FALLBACK; }
	const double var0 = reg_var(s, N);
	if (!(var0 >= 0.0)) { REG_WHY(125); return FAST_FALLBACK; }
	RegWin win; win.a0 = 0;
	if (!reg_window<T, RE>(v, g, N, (float) reg_mean(s, N), sqrtf((float) var0), st, win)) { REG_WHY(126); return FAST_FALLBACK; }
	if (!win_median(st, win, 0, N, result)) { REG_WHY(127); return FAST_FALLBACK; }
	return FAST_DONE;
}

}  // namespace sgpu
