// TEST INFRASTRUCTURE — replays the selection engine (siril_b200/csrc/psel.cuh) on the HOST: the
// group-level control flow (psel_group) and phase B (psel_pixel) are the product's own code; the
// streaming passes the kernel runs over registers (stack_psel.cuh) are restated here over plain
// arrays with the same float operations, lane by lane, so that the CPU-only suite can check the
// engine's decisions against the oracle.
#include <algorithm>
#include <cmath>
#include <vector>
#include "../../siril_b200/csrc/psel.cuh"

using namespace sgpu;

namespace {

const int LCHUNK = 8;	// stack_psel.cuh: samples between overflow checks
int lcap_of(int T) { return psel_list_cap(T); }

struct HostPasses {
	int T, R;
	std::vector<std::vector<float>> v;	// v[t][j]: sample j * T + t (NaN: absent)
	std::vector<std::vector<float>> lists;
	int passes = 0;
	bool lockstep = false;	// the worst the other groups of a warp can do: every pass runs again, with this group switched off
	bool any(bool b) { return b || lockstep; }
	static unsigned sgn(float x) { unsigned u; memcpy(&u, &x, 4); return u >> 31; }
	void sums(int is_mean, int *n, double *S1, double *S2, float *mn, float *mx) {
		passes++;
		double s1 = 0, s2 = 0; float lo = INFINITY, hi = -INFINITY; int cnt = 0;
		for (int t = 0; t < T; t++) {
			double a1 = 0, a2 = 0, b1 = 0, b2 = 0;
			for (int j = 0; j < R; j++) {
				const float x = v[t][j];
				if (x != x && !(j * T + t < cntN)) continue;	// static padding
				const double xd = (double) x;
				if (j & 1) { b1 += xd; b2 = fma(xd, xd, b2); } else { a1 += xd; a2 = fma(xd, xd, a2); }
				lo = fminf(lo, x); hi = fmaxf(hi, x); cnt++;
			}
			s1 += a1 + b1; s2 += a2 + b2;
		}
		if (is_mean && lo <= 0.f && hi >= 0.f) {	// nulls: drop them (rejection_float.c:128-135)
			lo = INFINITY; hi = -INFINITY;
			for (int t = 0; t < T; t++) for (int j = 0; j < R; j++) {
				float &x = v[t][j];
				if (x == 0.f) { x = PSEL_NAN; cnt--; }
				lo = fminf(lo, x); hi = fmaxf(hi, x);
			}
		}
		*n = cnt; *S1 = s1; *S2 = s2; *mn = lo; *mx = hi;
	}
	int cntN = 0;
	void shifted_sums(double K, double *S1, double *S2) {
		passes++;
		double s1 = 0, s2 = 0;
		for (int t = 0; t < T; t++) for (int j = 0; j < R; j++) {
			const float x = v[t][j];
			if (x != x) continue;
			const double y = (double) x - K; s1 += y; s2 = fma(y, y, s2);
		}
		*S1 = s1; *S2 = s2;
	}
	void count(float mean, float hw, int *c_neg, int *c_in) {
		passes++;
		unsigned a = 0, b = 0;
		for (int t = 0; t < T; t++) for (int j = 0; j < R; j++) {
			const float d = v[t][j] - mean;
			a += sgn(d); b += sgn(fabsf(d) - hw);
		}
		*c_neg = (int) a; *c_in = (int) b;
	}
	void extract(bool on, const PselPlan &pl, int *below, bool *over) {
		passes++;
		unsigned c = 0; bool ov = false;
		for (int t = 0; t < T; t++) {
			if (on) lists[t].clear();
			bool stop = !on;
			for (int j = 0; j < R; j++) {
				if (j % LCHUNK == 0 && (int) lists[t].size() > lcap_of(T) - LCHUNK) { stop = true; ov = true; }
				const float x = v[t][j];
				const float d1 = x - pl.p1, d2 = x - pl.cT;
				c += sgn(d1);
				if (!stop && (fabsf(d2) > pl.rT || fabsf(d1) < pl.w)) lists[t].push_back(x);
			}
		}
		*below = (int) c; *over = ov && on;
	}
	void classify(const PselPlan &pl, int *cc, int *ccb, int *nl, int *m) {
		int a = 0, b = 0, l = 0, all = 0;
		for (int t = 0; t < T; t++) for (float x : lists[t]) {
			const float d = x - pl.p1;
			if (fabsf(d) < pl.w) { a++; b += d < 0.f; } else l += d < 0.f;
			all++;
		}
		*cc = a; *ccb = b; *nl = l; *m = all;
	}
};

struct HostCol {
	float *p;
	float get(int i) const { return p[i]; }
	void set(int i, float x) const { p[i] = x; }
};
struct HostLists {
	const HostPasses &hp;
	int cnt(int t) const { return (int) hp.lists[t].size(); }
	float entry(int t, int k) const { return hp.lists[t][k]; }
};

}  // namespace

// returns status (FAST_DONE / FAST_FALLBACK); *result, rej[2] filled when FAST_DONE.
// diag[0..5]: header flags, streaming passes run, centre entries, low entries, high entries, kept n
extern "C" int psel_host_pixel(int is_float, const void *stack, int N, int T, int is_mean, int rejtype, float s0, float s1,
		const double *weights, int rej[2], double *result, int diag[6], int lockstep) {
	RejCfg c{ is_mean ? rejtype : 0, s0, s1, nullptr, N, 0.f, 0.f, nullptr };
	std::vector<float> o(N);
	for (int i = 0; i < N; i++) o[i] = is_float ? ((const float *) stack)[i] : (float) ((const u16 *) stack)[i];
	HostPasses hp;
	hp.T = T; hp.R = (N + T - 1) / T; hp.cntN = N; hp.lockstep = lockstep != 0;
	hp.v.assign(T, std::vector<float>(hp.R, PSEL_NAN));
	hp.lists.assign(T, {});
	for (int i = 0; i < N; i++) hp.v[i % T][i / T] = o[i];
	PselHdr h{};
	const bool tails = is_mean && (rejtype == SGPU_SIGMA || rejtype == SGPU_PERCENTILE);
	psel_group(hp, is_mean, tails, rejtype == SGPU_PERCENTILE ? 1 : 0, s0, s1, h);
	std::vector<float> colbuf(std::max(N, (int) PSEL_ROWS) + 3, -12345.f);
	HostCol col{ colbuf.data() };
	HostLists ls{ hp };
	rej[0] = rej[1] = 0;
	int status;
	auto orig = [&](int f) { return o[f]; };
	if (!(h.flags & (PSEL_WEIRD | PSEL_HANDOVER | PSEL_TRIVIAL))) {	// B1 as the kernel leaves it: every entry, sorted
		std::vector<float> lo, ce, hi;
		for (int t = 0; t < T; t++) for (float x : hp.lists[t]) {
			const float d = x - h.p1;
			(fabsf(d) < h.w ? ce : (d < 0.f ? lo : hi)).push_back(x);
		}
		std::sort(lo.begin(), lo.end()); std::sort(ce.begin(), ce.end()); std::sort(hi.begin(), hi.end());
		int r = 0;
		for (float x : lo) col.set(r++, x);
		for (float x : ce) col.set(r++, x);
		for (float x : hi) col.set(r++, x);
	}
	*result = is_float ? psel_pixel<false>(col, h, N, is_mean, c, weights, orig, rej, &status)
			   : psel_pixel<true>(col, h, N, is_mean, c, weights, orig, rej, &status);
	if (diag) {
		int cc = 0, nl = 0, nh = 0;
		for (int t = 0; t < T; t++) for (float x : hp.lists[t]) {
			const float d = x - h.p1;
			if (fabsf(d) < h.w) cc++; else if (d < 0.f) nl++; else nh++;
		}
		diag[0] = h.flags; diag[1] = hp.passes; diag[2] = cc; diag[3] = nl; diag[4] = nh; diag[5] = h.n;
	}
	return status;
}
