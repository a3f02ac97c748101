// stack_exact.cu — kernel + launcher of the thread-per-pixel engine (see exact.cuh).
//
// One CTA handles `blockDim.x` consecutive pixels of one layer.  For every frame the CTA reads
// the same pixel range of that frame's plane — consecutive threads read consecutive samples, so
// each warp issues full 128-byte coalesced HBM reads on the frame-major layout — applies the
// per-frame normalization in double exactly as median_and_mean.c:1644-1686 does, and parks the
// value in the thread's shared-memory column tile[f][tid] (conflict-free: consecutive threads
// hit consecutive banks).  The rejection / combine then runs entirely out of shared memory.
//
// LIST mode: the sorted engine (stack_sorted.cu) hands over the pixels it cannot decide; the same
// kernel then walks p.list (entries (layer << 30) | pixel, count on the device) grid-stride.
#include "kparams.cuh"

namespace sgpu {

// the pixel's rejection / combine on the stack already parked in the column `st`
template <typename T>
__device__ __forceinline__ void exact_pixel_body(const StackParams &p, int layer, size_t pix, Col<T> st, Col<u16> aux, int rej[2]) {
	const size_t npix = (size_t) p.H * p.W;
	const int x = (int) (pix % p.W);
	const int N = p.N;
	double result;
	if (p.is_mean) {
		const double *w = p.weights ? p.weights + (size_t) layer * N : nullptr;
		auto orig = [&](int ff) { return load_norm<T>(p, ff, layer, pix, x); };
		if (p.mask || p.drizz) {	// feather masks / drizzle weights ride along as per-sample weights (median_and_mean.c:1687-1692)
			const PixWeights pw = { w, p.mask ? p.mask + pix : nullptr, p.drizz ? p.drizz + (size_t) layer * npix + pix : nullptr,
				npix, (size_t) p.C * npix, p.shiftx };
			result = mean_and_reject<T>(st, N, p.rc, pw, aux, orig, rej);
		} else
			result = mean_and_reject<T>(st, N, p.rc, w, aux, orig, rej);
	} else {
		result = quickmedian_col(st, N);	// median_and_mean.c:1711-1715
	}
	store_result<T>(p, (size_t) layer * p.out_plane + pix, result, rej);
}

template <typename T>
__device__ __forceinline__ void exact_pixel(const StackParams &p, int layer, size_t pix, Col<T> st, Col<u16> aux, int rej[2]) {
	const int x = (int) (pix % p.W);
	const int N = p.N;
	int f = 0;
	for (; f + 4 <= N; f += 4) {	// 4 independent loads in flight per thread
		const T v0 = load_norm<T>(p, f, layer, pix, x), v1 = load_norm<T>(p, f + 1, layer, pix, x);
		const T v2 = load_norm<T>(p, f + 2, layer, pix, x), v3 = load_norm<T>(p, f + 3, layer, pix, x);
		st.set(f, v0); st.set(f + 1, v1); st.set(f + 2, v2); st.set(f + 3, v3);
	}
	for (; f < N; f++) st.set(f, load_norm<T>(p, f, layer, pix, x));
	exact_pixel_body<T>(p, layer, pix, st, aux, rej);
}

// LIST mode: one hand-over pixel per WARP.  Its samples sit a whole plane apart, so nothing coalesces; what the warp
// offers is 32 loads in flight instead of 4, and — the bigger part — lane 0 alone replays the pixel: 32 listed pixels on
// the 32 lanes of one warp would each take their own path through the rejection and serialize.
template <typename T>
__device__ __forceinline__ void exact_pixel_warp(const StackParams &p, int layer, size_t pix, Col<T> st0, Col<u16> aux0, int lane, int rej[2]) {
	const int x = (int) (pix % p.W);
	for (int f = lane; f < p.N; f += 32) st0.set(f, load_norm<T>(p, f, layer, pix, x));
	__syncwarp();
	if (lane == 0) exact_pixel_body<T>(p, layer, pix, st0, aux0, rej);
	__syncwarp();
}

template <typename T, bool SMEM, bool LIST>
__global__ void __launch_bounds__(256) stack_exact_kernel(const StackParams p) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const size_t npix = (size_t) p.H * p.W;
	Col<T> st; Col<u16> aux;
	if (SMEM) {
		st.p = (T *) smem_raw + threadIdx.x; st.s = blockDim.x;
		aux.p = (u16 *) (smem_raw + (size_t) p.N * blockDim.x * sizeof(T)) + threadIdx.x; aux.s = blockDim.x;
	} else {
		const size_t nthreads = (size_t) gridDim.x * gridDim.y * blockDim.x;
		const size_t gtid = ((size_t) blockIdx.y * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
		st.p = (T *) p.scratch + gtid; st.s = (int) nthreads;
		aux.p = (u16 *) ((T *) p.scratch + (size_t) p.N * nthreads) + gtid; aux.s = (int) nthreads;
	}
	if (!LIST) {
		const int layer = blockIdx.y;
		const size_t pix = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
		int rej[2] = { 0, 0 };
		if (pix < npix) exact_pixel<T>(p, layer, pix, st, aux, rej);
		if (p.is_mean && p.rc.type != SGPU_NO_REJEC) {	// irej[channel][low/high], :1748-1757
			unsigned lo = rej[0], hi = rej[1];
			lo = __reduce_add_sync(0xffffffffu, lo); hi = __reduce_add_sync(0xffffffffu, hi);
			if ((threadIdx.x & 31) == 0 && (lo | hi)) {
				atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
				atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
			}
		}
	} else {
		const unsigned count = *p.list_count;
		const int lane = threadIdx.x & 31;
		const unsigned nwarps = gridDim.x * (blockDim.x >> 5);
		Col<T> st0 = { st.p - lane, st.s }; Col<u16> aux0 = { aux.p - lane, aux.s };	// the column of the warp's lane 0
		if (SMEM) {	// one column per WARP: a CTA of 8 warps holds 8 stacks, so thousands of warps are resident and every listed pixel has its own
			const int nw = blockDim.x >> 5, wi = threadIdx.x >> 5;
			st0.p = (T *) smem_raw + wi; st0.s = nw;
			aux0.p = (u16 *) (smem_raw + (size_t) p.N * nw * sizeof(T)) + wi; aux0.s = nw;
		}
		for (unsigned e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); e < count; e += nwarps) {
			const unsigned ent = p.list[e];
			const int layer = (int) (ent >> 30);
			int rej[2] = { 0, 0 };
			exact_pixel_warp<T>(p, layer, (size_t) (ent & 0x3fffffffu), st0, aux0, lane, rej);
			if (lane == 0 && p.is_mean && (rej[0] | rej[1])) {
				if (rej[0]) atomicAdd(&p.irej[layer * 2], (unsigned long long) rej[0]);
				if (rej[1]) atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) rej[1]);
			}
		}
	}
}

}  // namespace sgpu

using namespace sgpu;

void sgpu_fill_stack_params(StackParams &p, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej) {
	const sgpu_block_desc &d = blk->d;
	p = StackParams{};
	p.frames = blk->dev; p.N = d.nb_frames; p.C = d.nb_layers; p.H = d.height; p.W = d.width;
	p.normalize = a->normalize; p.scale = dp.scale; p.offset = dp.offset; p.mul = dp.mul;
	p.weights = a->is_mean ? dp.weights : nullptr;
	p.shiftx = dp.shiftx;
	p.mask = a->is_mean ? blk->mask : nullptr; p.drizz = a->is_mean ? blk->drizz : nullptr;
	p.rc.type = a->is_mean ? a->type_of_rejection : SGPU_NO_REJEC;
	p.rc.s0 = a->sig[0]; p.rc.s1 = a->sig[1]; p.rc.crit = dp.crit; p.rc.nb_frames = d.nb_frames;
	p.rc.m_x = 0.f; p.rc.m_dx2 = 0.f; p.rc.xf = nullptr;
	p.rc.rcp = p.rc.type == SGPU_GESDT ? sgpu_rcp_table(blk->ctx) : nullptr;	// the sorted engines' Grubbs loop reads 1 / n from it
	if (p.rc.type == SGPU_LINEARFIT) {	// median_and_mean.c:1491-1499 (host, float)
		volatile float mx = (d.nb_frames - 1) * 0.5f, acc = 0.f;
		for (int j = 0; j < d.nb_frames; ++j) {
			volatile float dx = j - mx;
			volatile float sq = dx * dx;
			volatile float diff = sq - acc;
			volatile float xf = 1.f / (j + 1);
			volatile float prod = diff * xf;
			acc = acc + prod;
		}
		p.rc.m_x = mx; p.rc.m_dx2 = 1.f / acc;
	}
	p.is_mean = a->is_mean; p.output_norm = a->output_norm; p.out32 = a->use_32bit_output;
	p.norm16 = a->source_bitpix == 8 && a->output_norm && !a->use_32bit_output;
	p.rejmaps = (a->is_mean && a->create_rejmaps) ? (a->merge_lowhigh_rejmaps ? 1 : 2) : 0;
	p.out_plane = a->out_plane_stride ? a->out_plane_stride : (size_t) d.height * d.width;
	p.out = d_out; p.rej_low = d_rej_low; p.rej_high = d_rej_high; p.irej = (unsigned long long *) d_irej;
	p.use_aux = a->is_mean && (p.rc.type == SGPU_MAD || p.rc.type == SGPU_GESDT);
}

// CTA shape of the exact engine for this stack depth: largest CTA (<= 256 threads) whose tile
// leaves room for two CTAs per SM; else one CTA; else global columns (bd = 128, smem = false)
static void exact_shape(sgpu_ctx *ctx, size_t per_thread, int *bd_out, bool *smem_out) {
	int bd = 0;
	const size_t half = ctx->smem_optin / 2 - 1024;
	for (int cand = 256; cand >= 32; cand >>= 1) if (per_thread * cand <= half) { bd = cand; break; }
	if (!bd) for (int cand = 64; cand >= 32; cand >>= 1) if (per_thread * cand <= ctx->smem_optin) { bd = cand; break; }
	*smem_out = bd != 0;
	*bd_out = bd ? bd : 128;
}

template <typename K>
static int launch_kernel(sgpu_ctx *ctx, K kern, dim3 grid, int bd, size_t shm, const StackParams &p) {
	if (shm > 48 * 1024)
		SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm));
	kern<<<grid, bd, shm, ctx->stream>>>(p);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}

int sgpu_launch_exact(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej) {
	const sgpu_block_desc &d = blk->d;
	StackParams p;
	sgpu_fill_stack_params(p, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	const size_t per_thread = (size_t) d.nb_frames * blk->elem() + (p.use_aux ? (size_t) d.nb_frames * 2 : 0);
	const size_t npix = (size_t) d.height * d.width;
	if (npix == 0 || d.nb_frames == 0) return SGPU_OK;
	int bd; bool smem;
	exact_shape(ctx, per_thread, &bd, &smem);
	const size_t shm = smem ? per_thread * bd : 0;
	dim3 grid((unsigned) ((npix + bd - 1) / bd), (unsigned) d.nb_layers);
	if (!smem) {
		const size_t nthreads = (size_t) grid.x * grid.y * bd;
		if ((size_t) d.nb_frames * nthreads > 0x7fffffffull)
			return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "stack too deep for one launch (%d frames x %zu px)", d.nb_frames, npix);
		int rv = sgpu_ensure_scratch(ctx, per_thread * nthreads);
		if (rv) return rv;
		p.scratch = ctx->d_scratch;
	}
	if (d.dtype == SGPU_F32)
		return smem ? launch_kernel(ctx, stack_exact_kernel<float, true, false>, grid, bd, shm, p)
			    : launch_kernel(ctx, stack_exact_kernel<float, false, false>, grid, bd, shm, p);
	return smem ? launch_kernel(ctx, stack_exact_kernel<u16, true, false>, grid, bd, shm, p)
		    : launch_kernel(ctx, stack_exact_kernel<u16, false, false>, grid, bd, shm, p);
}

int sgpu_launch_exact_list(sgpu_ctx *ctx, sgpu_block *blk, StackParams p) {
	const sgpu_block_desc &d = blk->d;
	const size_t per_thread = (size_t) d.nb_frames * blk->elem() + (p.use_aux ? (size_t) d.nb_frames * 2 : 0);
	int bd; bool smem;
	exact_shape(ctx, per_thread, &bd, &smem);
	size_t shm = smem ? per_thread * bd : 0;
	// a fixed grid that walks the list, one listed pixel per warp: the count only exists on the device
	dim3 grid((unsigned) (ctx->sm_count * 2), 1);
	if (per_thread * 8 + 16 <= ctx->smem_optin) {	// shared-memory columns, one per warp
		smem = true; bd = 256; shm = per_thread * 8 + 16;
		const size_t per_sm = ctx->smem_optin / (shm + 1024);
		grid.x = (unsigned) (ctx->sm_count * (per_sm > 8 ? 8 : (per_sm < 1 ? 1 : per_sm)));
	}
	if (!smem) {
		int rv = sgpu_ensure_scratch(ctx, per_thread * (size_t) grid.x * bd);
		if (rv) return rv;
		p.scratch = ctx->d_scratch;
	}
	if (d.dtype == SGPU_F32)
		return smem ? launch_kernel(ctx, stack_exact_kernel<float, true, true>, grid, bd, shm, p)
			    : launch_kernel(ctx, stack_exact_kernel<float, false, true>, grid, bd, shm, p);
	return smem ? launch_kernel(ctx, stack_exact_kernel<u16, true, true>, grid, bd, shm, p)
		    : launch_kernel(ctx, stack_exact_kernel<u16, false, true>, grid, bd, shm, p);
}
