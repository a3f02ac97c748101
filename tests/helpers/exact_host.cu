// TEST INFRASTRUCTURE — runs the product's thread-per-pixel device functions (exact.cuh) on the
// HOST (they are __host__ __device__), so the CPU-only test-suite can check the exact engine's
// arithmetic against the oracle without a GPU.  The GPU tests still go through the C ABI.
#include <vector>
#include <cstring>
#include "../../siril_b200/csrc/exact.cuh"

using namespace sgpu;

extern "C" double exact_host_mean_and_reject(int is_float, const void *stack, int N, int rejtype, float s0, float s1,
		const float *crit, const double *weights, int layer, int rej[2]) {
	RejCfg c{ rejtype, s0, s1, crit, N, 0.f, 0.f };
	if (rejtype == 6) {
		volatile float mx = (N - 1) * 0.5f, acc = 0.f;
		for (int j = 0; j < N; ++j) { volatile float dx = j - mx; volatile float sq = dx * dx; volatile float df = sq - acc; volatile float xf = 1.f / (j + 1); volatile float pr = df * xf; acc = acc + pr; }
		c.m_x = mx; c.m_dx2 = 1.f / acc;
	}
	const double *w = weights ? weights + (size_t) layer * N : nullptr;
	std::vector<u16> aux(N + 1);
	Col<u16> ca{ aux.data(), 1 };
	rej[0] = rej[1] = 0;
	if (is_float) {
		std::vector<float> s((const float *) stack, (const float *) stack + N), o(s);
		Col<float> cs{ s.data(), 1 };
		return mean_and_reject<float>(cs, N, c, w, ca, [&](int f) { return o[f]; }, rej);
	}
	std::vector<u16> s((const u16 *) stack, (const u16 *) stack + N), o(s);
	Col<u16> cs{ s.data(), 1 };
	return mean_and_reject<u16>(cs, N, c, w, ca, [&](int f) { return o[f]; }, rej);
}

// the same with the pixel's gathered feather-mask / drizzle-weight stacks (either may be null): PixWeights with unit strides
extern "C" double exact_host_mean_and_reject_planes(int is_float, const void *stack, int N, int rejtype, float s0, float s1,
		const float *crit, const double *weights, const float *mstack, const float *dstack, int rej[2]) {
	RejCfg c{ rejtype, s0, s1, crit, N, 0.f, 0.f };
	if (rejtype == 6) {
		volatile float mx = (N - 1) * 0.5f, acc = 0.f;
		for (int j = 0; j < N; ++j) { volatile float dx = j - mx; volatile float sq = dx * dx; volatile float df = sq - acc; volatile float xf = 1.f / (j + 1); volatile float pr = df * xf; acc = acc + pr; }
		c.m_x = mx; c.m_dx2 = 1.f / acc;
	}
	const PixWeights pw = { weights, mstack, dstack, 1, 1, nullptr };
	std::vector<u16> aux(N + 1);
	Col<u16> ca{ aux.data(), 1 };
	rej[0] = rej[1] = 0;
	if (is_float) {
		std::vector<float> s((const float *) stack, (const float *) stack + N), o(s);
		Col<float> cs{ s.data(), 1 };
		return mean_and_reject<float>(cs, N, c, pw, ca, [&](int f) { return o[f]; }, rej);
	}
	std::vector<u16> s((const u16 *) stack, (const u16 *) stack + N), o(s);
	Col<u16> cs{ s.data(), 1 };
	return mean_and_reject<u16>(cs, N, c, pw, ca, [&](int f) { return o[f]; }, rej);
}

extern "C" double exact_host_quickmedian(int is_float, void *stack, int N) {
	if (is_float) { Col<float> c{ (float *) stack, 1 }; return quickmedian_col(c, N); }
	Col<u16> c{ (u16 *) stack, 1 }; return quickmedian_col(c, N);
}
