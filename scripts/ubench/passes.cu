// Pass-throughput probe for the selection engine (stack_psel): every thread keeps R samples in
// registers and runs one candidate "pass body" over them again and again with a changing pivot.
// Prints samples per clock per SM for each body, at 4 and 8 warps per SM sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false passes.cu -o passes
#include <cstdio>
#include <cuda_runtime.h>
#define R 64
#define ITER 512

template <int OP> __global__ void __launch_bounds__(512) k(float *out, float a, float b, int n, long long *cyc) {
	float v[R];
	__shared__ float sm[512 * 16];
#pragma unroll
	for (int j = 0; j < R; j++) v[j] = a + (threadIdx.x * 37 + j * 101) % 1000 * 1e-3f;
	for (int i = threadIdx.x; i < 512 * 16; i += blockDim.x) sm[i] = 0.f;
	__syncthreads();
	float p = b, acc = 0.f;
	unsigned c0 = 0, c1 = 0;
	double s1 = 0.0, s2 = 0.0;
	long long t0 = clock64();
	for (int i = 0; i < n; i++) {
		if (OP == 0) {	// FSETP + predicated IADD
#pragma unroll
			for (int j = 0; j < R; j++) c0 += (v[j] < p);
		}
		if (OP == 1) {	// FADD + sign bit by shift
#pragma unroll
			for (int j = 0; j < R; j++) c0 += __float_as_uint(v[j] - p) >> 31;
		}
		if (OP == 2) {	// FADD + IMAD.HI
#pragma unroll
			for (int j = 0; j < R; j++) { const unsigned db = __float_as_uint(v[j] - p); if (j & 1) asm("mad.hi.u32 %0, %1, 2, %0;" : "+r"(c1) : "r"(db)); else asm("mad.hi.u32 %0, %1, 2, %0;" : "+r"(c0) : "r"(db)); }
		}
		if (OP == 3) {	// C pass: below pivot + inside a window, both by sign bits
#pragma unroll
			for (int j = 0; j < R; j++) {
				const float d = v[j] - p;
				c0 += __float_as_uint(d) >> 31;
				c1 += __float_as_uint(fabsf(d) - a) >> 31;
			}
		}
		if (OP == 4) {	// E pass: interesting = far tail or centre window -> predicated append; below count
			unsigned ptr = (unsigned) __cvta_generic_to_shared(sm + threadIdx.x), ptr0 = ptr;
#pragma unroll
			for (int j = 0; j < R; j++) {
				const float d = v[j] - p;
				const float u = fabsf(d) - a;	// a = (r + w) / 2
				c0 += __float_as_uint(d) >> 31;
				asm volatile("{\n.reg .pred q;\nsetp.ge.f32 q, %2, %3;\n@q st.shared.f32 [%0], %1;\n@q add.u32 %0, %0, 2048;\n}" : "+r"(ptr) : "f"(v[j]), "f"(fabsf(u)), "f"(b) : "memory");
				if (j % 16 == 15) ptr = ptr0;
			}
			c1 += ptr;
		}
		if (OP == 5) {	// P1 sums: F2D + DADD + DFMA + min |x|
#pragma unroll
			for (int j = 0; j < R; j++) {
				const double x = (double) v[j];
				s1 += x; s2 = fma(x, x, s2);
				acc = fminf(acc, fabsf(v[j]));
			}
			p += 1e-7f;
		}
		if (OP == 6) {	// P1 sums with two accumulator pairs
			double t1 = 0.0, t2 = 0.0;
#pragma unroll
			for (int j = 0; j < R; j += 2) {
				const double x = (double) v[j], y = (double) v[j + 1];
				s1 += x; s2 = fma(x, x, s2);
				t1 += y; t2 = fma(y, y, t2);
			}
			s1 += t1; s2 += t2;
		}
		if (OP == 7) {	// F2D alone (+ one DADD per 4 to keep it live)
#pragma unroll
			for (int j = 0; j < R; j += 4) {
				const double x = (double) v[j] , y = (double) v[j + 1], z = (double) v[j + 2], w = (double) v[j + 3];
				s1 += __hiloint2double(__double2hiint(x) ^ __double2hiint(y), __double2loint(z) ^ __double2loint(w));
			}
		}
		if (OP == 8) {	// 3-input min / max
#pragma unroll
			for (int j = 0; j < R; j += 2) acc = fminf(acc, fminf(v[j], v[j + 1]));
			p += acc;
		}
		if (OP == 9) {	// clip scan: FADD + FSETP |d| > r accumulated as a max
#pragma unroll
			for (int j = 0; j < R; j++) acc = fmaxf(acc, fabsf(v[j] - p));
		}
		if (OP == 10) {	// float sums: FADD (x - K), FADD, FFMA
			float f1 = 0.f, f2 = 0.f;
#pragma unroll
			for (int j = 0; j < R; j++) { const float y = v[j] - p; f1 += y; f2 = fmaf(y, y, f2); }
			acc += f1 + f2;
		}
		if (OP == 11) {	// integer F2D by hand + DADD + DFMA
#pragma unroll
			for (int j = 0; j < R; j++) {
				const unsigned u = __float_as_uint(v[j]);
				const double x = __hiloint2double((int) (((u >> 3) & 0x0fffffffu) + 0x38000000u) | (int) (u & 0x80000000u), (int) (u << 29));
				s1 += x; s2 = fma(x, x, s2);
			}
		}
		if (OP == 12) {	// C pass with FSETP + IADD for the window
#pragma unroll
			for (int j = 0; j < R; j++) {
				const float d = v[j] - p;
				c0 += __float_as_uint(d) >> 31;
				c1 += (fabsf(d) < a);
			}
		}
		if (OP == 13) {	// FADD + sign add, four accumulators
			unsigned e0 = 0, e1 = 0, e2 = 0, e3 = 0;
#pragma unroll
			for (int j = 0; j < R; j += 4) {
				e0 += __float_as_uint(v[j] - p) >> 31; e1 += __float_as_uint(v[j + 1] - p) >> 31;
				e2 += __float_as_uint(v[j + 2] - p) >> 31; e3 += __float_as_uint(v[j + 3] - p) >> 31;
			}
			c0 += e0 + e1; c1 += e2 + e3;
		}
		if (OP == 14) {	// FSETP + @IADD, four accumulators
			unsigned e0 = 0, e1 = 0, e2 = 0, e3 = 0;
#pragma unroll
			for (int j = 0; j < R; j += 4) { e0 += (v[j] < p); e1 += (v[j + 1] < p); e2 += (v[j + 2] < p); e3 += (v[j + 3] < p); }
			c0 += e0 + e1; c1 += e2 + e3;
		}
		p += 1e-6f;
	}
	long long t1 = clock64();
	float s = acc + (float) s1 + (float) s2 + c0 + c1 + p;
#pragma unroll
	for (int j = 0; j < R; j++) s += v[j];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s + sm[threadIdx.x];
	if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP> void run(const char *name) {
	float *out; long long *cyc, h;
	cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 8);
	for (int threads = 256; threads <= 512; threads *= 2) {
		k<OP><<<148, threads>>>(out, 0.4f, 0.05f, 8, cyc);
		k<OP><<<148, threads>>>(out, 0.4f, 0.05f, ITER, cyc);
		cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
		cudaError_t e = cudaDeviceSynchronize();
		printf("%-44s %2d warps/SM: %7.2f samples/clk/SM %s\n", name, threads / 32, (double) threads * ITER * R / (double) h, e ? cudaGetErrorString(e) : "");
	}
	cudaFree(out); cudaFree(cyc);
}

int main() {
	run<0>("count: FSETP + @IADD");
	run<1>("count: FADD + (bits >> 31) add");
	run<2>("count: FADD + IMAD.HI");
	run<13>("count: FADD + sign add, 4 accumulators");
	run<14>("count: FSETP + @IADD, 4 accumulators");
	run<3>("C pass: 2 FADD + 2 sign adds");
	run<12>("C pass: FADD + sign add + FSETP + IADD");
	run<4>("E pass: 2 FADD + sign add + FSETP + @STS,@IADD");
	run<5>("P1: F2D + DADD + DFMA + FMNMX");
	run<6>("P1: F2D + DADD + DFMA, 2 chains");
	run<11>("P1: integer F2D + DADD + DFMA");
	run<7>("F2D alone");
	run<8>("FMNMX 3-input form");
	run<9>("clip scan: FADD + FMNMX");
	run<10>("float sums: 2 FADD + FFMA");
	return 0;
}
