// Instruction-throughput probe for the ops the stacking kernel leans on (B200, sm_100a).
// Prints lane-results per clock per SM for each op class.  nvcc -arch=sm_100a -O3 rates.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
#define U 8
template <int OP> __global__ void k(float *out, float a, float b, int n, long long *cyc) {
	float x[U]; double d[U]; int c[U];
	for (int u = 0; u < U; u++) { x[u] = a + threadIdx.x * 1e-3f + u; d[u] = x[u]; c[u] = 0; }
	__shared__ float sm[1024 * 8];
	for (int i = threadIdx.x; i < 1024 * 8; i += blockDim.x) sm[i] = i;
	__syncthreads();
	long long t0 = clock64();
	for (int i = 0; i < n; i++) {
#pragma unroll
		for (int u = 0; u < U; u++) {
			if (OP == 0) x[u] = fmaf(x[u], a, b);
			if (OP == 1) x[u] = fminf(x[u], fmaxf(b, x[(u + 1) % U]));      // 2 FMNMX
			if (OP == 2) { if (x[u] < b + i) c[u]++; }                      // FSETP + pred IADD
			if (OP == 3) d[u] = fma(d[u], (double) a, (double) b);           // DFMA
			if (OP == 4) { d[u] = (double) x[u] + d[u]; }                    // F2D + DADD
			if (OP == 5) { x[u] = (float) (d[u]) + x[u]; d[u] += 1.0; }       // D2F + FADD + DADD
			if (OP == 6) x[u] = __shfl_xor_sync(0xffffffffu, x[u], 1 + (u & 3)) + a; // SHFL + FADD
			if (OP == 7) x[u] += sm[(threadIdx.x + (i * U + u) * 32) & 8191];   // LDS + FADD
			if (OP == 8) x[u] = x[u] + a;                                     // FADD
			if (OP == 9) { float y = x[u] - a; c[u] += (fabsf(y) > b); x[u] = y; }  // FADD + FSETP + IADD
			if (OP == 10) { float t = (float) ((double) x[u] * (double) a - (double) b); x[u] = t; }  // F2D DMUL DADD D2F
			if (OP == 11) { c[u] += __float_as_int(x[u]) - i; c[u] = max(c[u], i); }  // IADD + IMNMX
			if (OP == 12) { sm[(threadIdx.x + ((i * U + u) & 7) * 1024 + c[u]) & 8191] = x[u]; }  // STS
		}
	}
	long long t1 = clock64();
	float s = 0; for (int u = 0; u < U; u++) s += x[u] + (float) d[u] + c[u];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s + sm[threadIdx.x];
	if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int OP> void run(const char *name, int ops_per) {
	float *out; long long *cyc, h;
	cudaMalloc(&out, 148 * 1024 * 4 * 4); cudaMalloc(&cyc, 8);
	const int threads = 1024, blocks = 148;   // 32 warps per SM, 8 per SMSP
	k<OP><<<blocks, threads>>>(out, 1.0001f, 0.5f, 16, cyc);
	k<OP><<<blocks, threads>>>(out, 1.0001f, 0.5f, ITER, cyc);
	cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	cudaError_t e = cudaDeviceSynchronize();
	double per_clk = (double) threads * ITER * U / (double) h;
	printf("%-28s %8.1f iterations(lanes)/clk/SM  (%d ops each -> %.1f lane-ops/clk/SM) %s\n", name, per_clk, ops_per, per_clk * ops_per, e ? cudaGetErrorString(e) : "");
	cudaFree(out); cudaFree(cyc);
}
int main() {
	run<0>("FFMA", 1); run<8>("FADD", 1); run<1>("2xFMNMX", 2); run<2>("FSETP+@IADD", 2); run<3>("DFMA", 1);
	run<4>("F2D+DADD", 2); run<5>("D2F+FADD+DADD", 3); run<6>("SHFL+FADD", 2); run<7>("LDS+FADD", 2);
	run<9>("FADD+FSETP+IADD", 3); run<10>("F2D,DMUL,DADD,D2F", 4); run<11>("IADD+IMNMX", 2); run<12>("STS", 1);
	return 0;
}
