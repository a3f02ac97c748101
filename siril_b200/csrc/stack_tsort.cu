// stack_tsort.cu — launcher of the thread-sort engine (kernel: stack_tsort.cuh): picks the
// smallest compiled network that holds the stack, clears the hand-over list, launches, then lets
// the exact engine redo the pixels on the list.
#include <cstdlib>
#include "stack_tsort.cuh"

using namespace sgpu;

bool sgpu_sorted_supports(const sgpu_block_desc &d, const sgpu_stack_args *a);

bool sgpu_tsort_supports(const sgpu_block_desc &d, const sgpu_stack_args *a) {
	return d.nb_frames <= 128 && sgpu_sorted_supports(d, a);
}

struct TsortVariant { int ns, minb; int (*launch)(sgpu_ctx *, const StackParams &, dim3, size_t); };
static const TsortVariant VARIANTS[] = {
	{ 8, 8, sgpu_launch_tsort_variant<8, 8> }, { 16, 8, sgpu_launch_tsort_variant<16, 8> }, { 24, 8, sgpu_launch_tsort_variant<24, 8> },
	{ 32, 6, sgpu_launch_tsort_variant<32, 6> }, { 50, 4, sgpu_launch_tsort_variant<50, 4> }, { 64, 4, sgpu_launch_tsort_variant<64, 4> },
	{ 80, 3, sgpu_launch_tsort_variant<80, 3> }, { 100, 3, sgpu_launch_tsort_variant<100, 3> }, { 100, 2, sgpu_launch_tsort_variant<100, 2> }, { 128, 2, sgpu_launch_tsort_variant<128, 2> },
};

int sgpu_launch_tsort(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej) {
	const sgpu_block_desc &d = blk->d;
	const size_t npix = (size_t) d.height * d.width;
	if (npix == 0) return SGPU_OK;
	StackParams p;
	sgpu_fill_stack_params(p, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	const size_t list_bytes = ((npix * d.nb_layers + 1) * sizeof(unsigned) + 255) & ~(size_t) 255;
	int rv = sgpu_ensure_list(ctx, list_bytes); if (rv) return rv;
	p.list_count = (unsigned *) ctx->d_list; p.list = p.list_count + 1;
	SGPU_CUDA(ctx, cudaMemsetAsync(p.list_count, 0, sizeof(unsigned), ctx->stream));
	ctx->list_used = true;
	p.is_u16 = d.dtype == SGPU_U16;
	p.need_sums = a->is_mean && a->type_of_rejection != SGPU_LINEARFIT;
	const TsortVariant *var = nullptr;
	int want_minb = 0;	// tuning aid: SGPU_TSORT_MINB picks among variants of one size compiled for different occupancies
	if (const char *e = getenv("SGPU_TSORT_MINB")) want_minb = atoi(e);
	for (const TsortVariant &v : VARIANTS) if (v.ns >= d.nb_frames) {
		if (!var) var = &v;
		else if (v.ns == var->ns && v.minb == want_minb) var = &v;
		else if (v.ns != var->ns) break;
	}
	if (!var) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "no thread-sort variant for %d frames", d.nb_frames);
	const size_t shm = (size_t) d.nb_frames * TS_BD * sizeof(float);
	dim3 grid((unsigned) ((npix + TS_BD - 1) / TS_BD), (unsigned) d.nb_layers);
	rv = var->launch(ctx, p, grid, shm);
	if (rv) return rv;
	return sgpu_launch_exact_list(ctx, blk, p);
}
