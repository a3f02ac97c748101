// TEST INFRASTRUCTURE — runs the product's sorted-stack rejection functions (sorted.cuh, phase B
// of the sorted engine) on the HOST against stacks sorted with std::sort, so that the CPU-only
// suite can check their arithmetic and their fallback decisions against the oracle.
#include <algorithm>
#include <vector>
#include "../../siril_b200/csrc/sorted.cuh"

using namespace sgpu;

struct HostAcc {
	float *p;
	float operator()(int i) const { return p[i]; }
	void set(int i, float v) { p[i] = v; }
};

// returns status (FAST_DONE / FAST_FALLBACK); *mean, rej[2] filled when FAST_DONE
extern "C" int sorted_host_mean_and_reject(int is_float, const void *stack, int N, int rejtype, float s0, float s1,
		const float *crit, const double *weights, int layer, int rej[2], double *mean, int use_s0) {
	RejCfg c{ rejtype, s0, s1, crit, N, 0.f, 0.f };
	if (rejtype == 6) {
		volatile float mx = (N - 1) * 0.5f, acc = 0.f;
		for (int j = 0; j < N; ++j) { volatile float dx = j - mx; volatile float sq = dx * dx; volatile float df = sq - acc; volatile float xf = 1.f / (j + 1); volatile float pr = df * xf; acc = acc + pr; }
		c.m_x = mx; c.m_dx2 = 1.f / acc;
	}
	const double *w = weights ? weights + (size_t) layer * N : nullptr;
	std::vector<float> o(N), s;
	for (int i = 0; i < N; i++) o[i] = is_float ? ((const float *) stack)[i] : (float) ((const u16 *) stack)[i];
	for (int i = 0; i < N; i++) if (o[i] != 0.f) s.push_back(o[i]);
	std::sort(s.begin(), s.end());
	const int n = (int) s.size();
	s.resize(N + 1);
	HostAcc acc{ s.data() };
	rej[0] = rej[1] = 0;
	int status;
	auto orig = [&](int f) { return o[f]; };
	// phase A's pre-accumulated sums: 8 lane partials around a sample from the bulk, then a tree
	Sums S0, part[8];
	if (use_s0 && n > 0) {
		const float K = s[((n / 2) / 16) * 16];
		for (int t = 0; t < 8; t++) part[t].init(K);
		for (int i = 0; i < n; i++) part[(i / 16) % 8].add(s[i]);
		for (int m = 1; m < 8; m <<= 1) for (int t = 0; t < 8; t += 2 * m) { part[t].S1 += part[t + m].S1; part[t].S2 += part[t + m].S2; }
		S0 = part[0];
	}
	const Sums *ps = (use_s0 && n > 0) ? &S0 : nullptr;
	*mean = is_float ? fast_mean_and_reject<false>(acc, n, N, c, w, orig, ps, rej, &status)
			 : fast_mean_and_reject<true>(acc, n, N, c, w, orig, ps, rej, &status);
	return status;
}

extern "C" double sorted_host_div_by_int(double a, int n) { return div_by_int(a, n); }

extern "C" double sorted_host_median(int is_float, const void *stack, int N) {
	std::vector<float> s(N);
	for (int i = 0; i < N; i++) s[i] = is_float ? ((const float *) stack)[i] : (float) ((const u16 *) stack)[i];
	std::sort(s.begin(), s.end());
	HostAcc acc{ s.data() };
	return is_float ? median_sorted<false>(acc, 0, N) : median_sorted<true>(acc, 0, N);
}
#ifdef SGPU_FAST_DEBUG
extern "C" int sorted_host_reason(void) { return sgpu::g_fast_reason; }
#endif
