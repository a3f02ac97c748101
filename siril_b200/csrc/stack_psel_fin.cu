// stack_psel_fin.cu — second kernel of the selection engine's split form: one THREAD per pixel takes the
// record the stack kernel left (psel.cuh: PSEL_REC_*), sorts its entries in registers (a 64-input
// network), parks them in a conflict-free shared-memory column and runs phase B2 (psel_pixel).
// The stack kernel's CTAs are sized by the tile they hold (four per SM, sixteen warps); the scalar tail of a
// pixel — a dependent chain of a few hundred instructions — ran there either on one lane in four (the other
// three repeating it) or on one warp in four.  Here it runs without that redundancy at full occupancy; the
// price is the record's trip through HBM (320 bytes per pixel each way).
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(64) {
#include "gen/sortnet_64.inc"
}

#define FIN_THREADS 128
#define FIN_ROWS (PSEL_REC_ENTRIES + 2)

struct FinCol {	// row i of this thread's column: bank = thread, whatever the row
	float *base;
	__device__ __forceinline__ float get(int i) const { return base[i * FIN_THREADS]; }
	__device__ __forceinline__ void set(int i, float x) const { base[i * FIN_THREADS] = x; }
};

__global__ void __launch_bounds__(FIN_THREADS) stack_psel_finish_kernel(const StackParams p, unsigned chunk_px) {
	__shared__ float cols[FIN_ROWS * FIN_THREADS];
	__shared__ unsigned cta_rej[2];
	if (threadIdx.x == 0) { cta_rej[0] = 0; cta_rej[1] = 0; }
	__syncthreads();
	const unsigned i = blockIdx.x * FIN_THREADS + threadIdx.x;
	const size_t npix = (size_t) p.H * p.W;
	const int layer = p.layer_base, lane = threadIdx.x & 31, N = p.N;
	int rej[2] = { 0, 0 };
	if (i < chunk_px) {
		const size_t pix = p.pix_base + i;
		const uint4 *rq = (const uint4 *) (p.records + (size_t) i * PSEL_REC_WORDS);
		const uint4 q0 = __ldg(rq), q1 = __ldg(rq + 1), q2 = __ldg(rq + 2), q3 = __ldg(rq + 3);
		PselHdr h;
		h.S1 = __hiloint2double((int) q0.y, (int) q0.x); h.S2 = __hiloint2double((int) q0.w, (int) q0.z);
		h.K = __hiloint2double((int) q1.y, (int) q1.x); h.mn = __uint_as_float(q1.z); h.mx = __uint_as_float(q1.w);
		h.blo = __uint_as_float(q2.x); h.bhi = __uint_as_float(q2.y); h.n = (int) q2.z; h.cl = (int) q2.w;
		h.nl = (int) q3.x; h.cc = (int) q3.y; h.m = (int) q3.z; h.flags = (int) q3.w;
		h.p1 = 0.f; h.w = 0.f; h.tid0 = 0;
		const bool live = !(h.flags & (PSEL_WEIRD | PSEL_HANDOVER | PSEL_TRIVIAL));
		const FinCol col{ cols + threadIdx.x };
		if (!(h.flags & PSEL_DONE)) {
		if (live) {
			float e[PSEL_REC_ENTRIES];
#pragma unroll
			for (int k = 0; k < PSEL_REC_ENTRIES / 4; k++) {
				uint4 v = make_uint4(0u, 0u, 0u, 0u);
				if (4 * k < h.m) v = __ldg(rq + PSEL_REC_HDR / 4 + k);	// nothing was written past entry m
				e[4 * k] = 4 * k < h.m ? __uint_as_float(v.x) : SGPU_INF; e[4 * k + 1] = 4 * k + 1 < h.m ? __uint_as_float(v.y) : SGPU_INF;
				e[4 * k + 2] = 4 * k + 2 < h.m ? __uint_as_float(v.z) : SGPU_INF; e[4 * k + 3] = 4 * k + 3 < h.m ? __uint_as_float(v.w) : SGPU_INF;
			}
			LocalNet<64>::template run<64>(e);	// (a 48-input network for the warps without a longer list was tried: the second copy of the code cost more than its 159 comparators saved)
#pragma unroll
			for (int k = 0; k < PSEL_REC_ENTRIES; k++) col.set(k, e[k]);
		}
		int status;
		double result;
		const int x = (int) (pix % p.W);
		const double *w = (p.is_mean && p.weights) ? p.weights + (size_t) layer * N : nullptr;
		if (p.is_u16) {
			auto orig = [&](int ff) { return (float) load_norm<u16>(p, ff, layer, pix, x); };
			result = psel_pixel<true>(col, h, N, p.is_mean, p.rc, w, orig, rej, &status, p.rcp_tab);
		} else {
			auto orig = [&](int ff) { return load_norm<float>(p, ff, layer, pix, x); };
			result = psel_pixel<false>(col, h, N, p.is_mean, p.rc, w, orig, rej, &status, p.rcp_tab);
		}
		if (status == FAST_DONE) {
			if (p.is_u16) store_result<u16>(p, (size_t) layer * p.out_plane + pix, result, rej);
			else store_result<float>(p, (size_t) layer * p.out_plane + pix, result, rej);
		} else {
			rej[0] = rej[1] = 0;
			p.list[atomicAdd(p.list_count, 1u)] = ((unsigned) layer << 30) | (unsigned) pix;
		}
		}	// !PSEL_DONE
	}
	if (p.is_mean && p.rc.type != SGPU_NO_REJEC) {
		__syncwarp();
		unsigned lo = rej[0], hi = rej[1];
		lo = __reduce_add_sync(0xffffffffu, lo); hi = __reduce_add_sync(0xffffffffu, hi);
		if (lane == 0 && (lo | hi)) { atomicAdd(&cta_rej[0], lo); atomicAdd(&cta_rej[1], hi); }
		__syncthreads();
		if (threadIdx.x == 0 && (cta_rej[0] | cta_rej[1])) {
			atomicAdd(&p.irej[layer * 2], (unsigned long long) cta_rej[0]);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) cta_rej[1]);
		}
	}
}

}  // namespace sgpu

int sgpu_launch_psel_finish(sgpu_ctx *ctx, const sgpu::StackParams &p, unsigned chunk_px) {
	if (!chunk_px) return SGPU_OK;
	sgpu::stack_psel_finish_kernel<<<(chunk_px + FIN_THREADS - 1) / FIN_THREADS, FIN_THREADS, 0, ctx->stream>>>(p, chunk_px);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}
