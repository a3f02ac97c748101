// Compare-exchange forms for the register sort networks (B200, sm_100a): time a 32-value
// in-register bitonic merge with
//   A  min = FMNMX, max = FMNMX                      (2 ALU-pipe ops)
//   B  min = FMNMX, max = a + b - min on the bit patterns, both adds as IMAD with a run-time 1 / -1
//      (1 ALU + 2 FMA-pipe ops; exact: min is bitwise one of the inputs)
//   C  like B with plain integer adds (ptxas picks IADD3 or IMAD.IADD)
//   M  two of three exchanges as B, one as A
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false cex.cu -o cex
#include <cstdio>
#include <cuda_runtime.h>
#define R 32
template <int F> __device__ __forceinline__ void ce(float &a, float &b, int one, int mone, int k) {
	const float mn = fminf(a, b);
	if (F == 0 || (F == 3 && (k % 3) == 2)) { b = fmaxf(a, b); a = mn; return; }
	const int ia = __float_as_int(a), ib = __float_as_int(b), im = __float_as_int(mn);
	int mx;
	if (F == 2) mx = ia + ib - im;
	else { int t; asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(ia), "r"(one), "r"(ib)); asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(mx) : "r"(im), "r"(mone), "r"(t)); }
	a = mn; b = __int_as_float(mx);
}
template <int F> __global__ void __launch_bounds__(256) k(float *out, const float *in, int iters, int one, int mone, long long *cyc) {
	float v[R];
	for (int j = 0; j < R; j++) v[j] = in[(threadIdx.x * R + j + blockIdx.x) & 4095];
	long long t0 = clock64();
	for (int it = 0; it < iters; it++) {
		int kk = 0;
#pragma unroll
		for (int m = R >> 1; m > 0; m >>= 1) {
#pragma unroll
			for (int j = 0; j < R; j++) if ((j & m) == 0) { ce<F>(v[j], v[j | m], one, mone, kk); kk++; }
		}
		v[it & (R - 1)] += 1.0f;	// keep the network from being hoisted
	}
	long long t1 = clock64();
	float s = 0; for (int j = 0; j < R; j++) s += v[j] * (j + 1);
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int F> double run(const char *name, const float *in, float *out, long long *cyc, float *h_out) {
	const int iters = 2000, threads = 256, blocks = 148 * 4;	// 32 warps per SM
	k<F><<<blocks, threads>>>(out, in, 10, 1, -1, cyc);
	k<F><<<blocks, threads>>>(out, in, iters, 1, -1, cyc);
	long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	cudaMemcpy(h_out, out, 4096 * 4, cudaMemcpyDeviceToHost);
	const double ce_per_clk = 4.0 * threads * iters * (R / 2 * 5) / (double) h;	// lane compare-exchanges per clk per SM
	printf("%-26s %7.1f lane-CE/clk/SM   checksum %.6e  %s\n", name, ce_per_clk, (double) h_out[5] + h_out[999], cudaGetErrorString(cudaGetLastError()));
	return ce_per_clk;
}
int main() {
	float *in, *out, h[4096]; long long *cyc;
	for (int i = 0; i < 4096; i++) h[i] = (float) ((i * 2654435761u) % 100003) * 1e-3f - 20.f;
	cudaMalloc(&in, 4096 * 4); cudaMalloc(&out, 148 * 4 * 256 * 4); cudaMalloc(&cyc, 8);
	cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
	static float ho[4096];
	run<0>("A: FMNMX + FMNMX", in, out, cyc, ho);
	run<1>("B: FMNMX + 2 IMAD", in, out, cyc, ho);
	run<2>("C: FMNMX + int adds", in, out, cyc, ho);
	run<3>("M: 2/3 B, 1/3 A", in, out, cyc, ho);
	return 0;
}
