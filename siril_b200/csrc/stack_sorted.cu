// stack_sorted.cu — launcher of the sorted engine (kernel: stack_sorted.cuh): picks the variant
// (lanes per pixel, register slots per lane) for the stack depth, clears the fallback list,
// launches, then lets the exact engine redo the pixels on the list.
#include "stack_sorted.cuh"

using namespace sgpu;

bool sgpu_sorted_supports(const sgpu_block_desc &d, const sgpu_stack_args *a) {
	if (d.nb_frames > 2048 || (size_t) d.height * d.width >= (1u << 30)) return false;
	if (!a->is_mean) return true;
	switch (a->type_of_rejection) {
	case SGPU_NO_REJEC: case SGPU_PERCENTILE: case SGPU_SIGMA: case SGPU_LINEARFIT: case SGPU_GESDT: return true;
	case SGPU_WINSORIZED: return d.dtype == SGPU_F32;
	default: return false;	// MAD, SIGMEDIAN: exact engine
	}
}

#define TRY(LT, RE) if (per_lane <= RE) return sgpu_launch_sorted_variant<LT, RE>(ctx, p, grid, shm)
static int dispatch_sorted(sgpu_ctx *ctx, const StackParams &p, dim3 grid, size_t shm) {
	const int N = p.N;
	if (N <= 512) {
		const int per_lane = (N + 7) / 8;
		TRY(3, 4); TRY(3, 8); TRY(3, 13); TRY(3, 16); TRY(3, 20); TRY(3, 25); TRY(3, 32); TRY(3, 38); TRY(3, 44); TRY(3, 50); TRY(3, 57); TRY(3, 64);
	} else if (N <= 1024) {
		const int per_lane = (N + 15) / 16;
		TRY(4, 38); TRY(4, 44); TRY(4, 50); TRY(4, 57); TRY(4, 64);
	} else {
		const int per_lane = (N + 31) / 32;
		TRY(5, 38); TRY(5, 44); TRY(5, 50); TRY(5, 57); TRY(5, 64);
	}
	return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "no sorted-engine variant for %d frames", N);
}

int sgpu_launch_sorted(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej) {
	const sgpu_block_desc &d = blk->d;
	const size_t npix = (size_t) d.height * d.width;
	if (npix == 0) return SGPU_OK;
	StackParams p;
	sgpu_fill_stack_params(p, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	// fallback list: one counter + at most one entry per pixel
	const size_t list_bytes = ((npix * d.nb_layers + 1) * sizeof(unsigned) + 255) & ~(size_t) 255;
	int rv = sgpu_ensure_list(ctx, list_bytes); if (rv) return rv;
	p.list_count = (unsigned *) ctx->d_list; p.list = p.list_count + 1;
	SGPU_CUDA(ctx, cudaMemsetAsync(p.list_count, 0, sizeof(unsigned), ctx->stream));
	p.is_u16 = d.dtype == SGPU_U16;
	p.need_sums = a->is_mean && a->type_of_rejection != SGPU_LINEARFIT;
	const size_t shm = (size_t) TILE_PX * d.nb_frames * sizeof(float) + 2 * TILE_PX * sizeof(int) + TILE_PX * sizeof(Sums);
	if (shm > ctx->smem_optin) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "stack too deep for the sorted engine");
	dim3 grid((unsigned) ((npix + TILE_PX - 1) / TILE_PX), (unsigned) d.nb_layers);
	rv = dispatch_sorted(ctx, p, grid, shm);
	if (rv) return rv;
	return sgpu_launch_exact_list(ctx, blk, p);
}
