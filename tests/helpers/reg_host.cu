// TEST INFRASTRUCTURE — runs the product's register-engine per-pixel logic (reg.cuh) on the HOST
// (one "lane" holding the whole stack) so that the CPU-only suite can check its arithmetic, its
// median window and its hand-over decisions against the oracle.
#include <vector>
#define SGPU_FAST_DEBUG 1
#include "../../siril_b200/csrc/reg.cuh"

using namespace sgpu;

constexpr int HRE = 512;

struct HostSt {
	float list[HRE], wb[REG_WCAP + 8], sb[REG_WCAP + 8];
	void put(int e, float x) { list[e] = x; }
	float get(int e) const { return list[e]; }
	float &w(int i) { return wb[i]; }
	float w(int i) const { return wb[i]; }
	float &sw(int i) { return sb[i]; }
	float sw(int i) const { return sb[i]; }
};

// status (FAST_DONE / FAST_FALLBACK); *mean, rej[2] filled when FAST_DONE.  is_mean == 0: median stack.
extern "C" int reg_host_stack(const float *stack, int N, int is_mean, int rejtype, float s0, float s1, int rej[2], double *result) {
	g_reg_reason = 0;
	if (N > HRE || N < 3) return FAST_FALLBACK;
	static float v[HRE];
	float (&vr)[HRE] = v;
	for (int i = 0; i < HRE; i++) v[i] = i < N ? stack[i] : reg_nan();
	if (is_mean) for (int i = 0; i < N; i++) if (v[i] == 0.f) return FAST_FALLBACK;	// the kernel hands these over
	Group<1> g{ 1u, 0 };
	static HostSt st;
	RejCfg c{ rejtype, s0, s1, nullptr, N, 0.f, 0.f };
	rej[0] = rej[1] = 0;
	if (!is_mean) return reg_median<1, HRE, 0>(vr, g, N, st, result);
	return reg_mean_and_reject<1, HRE, 0, true>(vr, g, N, c, st, rej, result);
}

extern "C" int reg_host_reason(void) { return sgpu::g_reg_reason; }
extern "C" void reg_host_counters(int out[4]) { out[0] = sgpu::g_reg_exact; out[1] = sgpu::g_reg_inner; out[2] = sgpu::g_reg_passes; out[3] = sgpu::g_reg_rebuilt; }
