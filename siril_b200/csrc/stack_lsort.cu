// stack_lsort.cu — launcher of the lane-sort engine (kernel: stack_lsort.cuh): picks the variant
// (lanes per pixel, register slots, rows that hold samples, sub-tiles and warps per CTA) for the
// stack depth, clears the hand-over list, launches, then lets the exact engine redo the list.
#include <cstdlib>
#include "stack_lsort.cuh"

using namespace sgpu;

int sgpu_make_tile_tmap(sgpu_ctx *ctx, CUtensorMap *tm, void *base, size_t npix, int layers, int frames, int box_px, int box_rows);
bool sgpu_sorted_supports(const sgpu_block_desc &d, const sgpu_stack_args *a);

bool sgpu_lsort_supports(const sgpu_block_desc &d, const sgpu_stack_args *a) {
	return d.nb_frames <= 1024 && sgpu_sorted_supports(d, a);
}

// 2^lt lanes per pixel, re register slots per lane of which the first rs can hold samples
// (the variant serves rf * 2^lt < N <= rs * 2^lt: rows below rf are always full), sub sub-tiles of 32 px and nw warps per CTA
struct LsortVariant { int lt, re, rs, sub, nw; int (*launch)(sgpu_ctx *, const StackParams &, const CUtensorMap *, dim3, size_t); };
#define LV(LT, RE, RS, RF, SUB, NW) { LT, RE, RS, SUB, NW, sgpu_launch_lsort_variant<LT, RE, RS, RF, SUB, NW> }
static const LsortVariant VARIANTS[] = {
	LV(1, 8, 8, 0, 4, 8),	// N <= 16
	LV(2, 8, 8, 4, 4, 8),	// 32
	LV(2, 16, 13, 8, 4, 8),	// 52
	LV(2, 16, 16, 13, 4, 8),	// 64
	LV(2, 20, 20, 16, 4, 8),	// 80
	LV(2, 24, 24, 20, 4, 8),	// 96
	LV(2, 32, 25, 20, 4, 8),	// 100
	LV(2, 32, 25, 20, 2, 4),	//   alternates of the same network (SGPU_LSORT_ALT = 1, 2, 3): tuning aid
	LV(2, 32, 25, 20, 4, 4),
	LV(2, 32, 25, 20, 2, 8),
	LV(2, 32, 32, 25, 4, 8),	// 128
	LV(3, 20, 20, 16, 2, 8),	// 160
	LV(3, 24, 24, 20, 2, 8),	// 192
	LV(3, 32, 25, 20, 2, 8),	// 200
	LV(3, 32, 32, 25, 2, 8),	// 256
	LV(4, 20, 20, 16, 1, 8),	// 320
	LV(4, 24, 24, 20, 1, 8),	// 384
	LV(4, 32, 25, 19, 1, 8),	// 400
	LV(4, 32, 32, 25, 1, 8),	// 512
	LV(5, 20, 20, 16, 1, 8),	// 640
	LV(5, 24, 24, 20, 1, 8),	// 768
	LV(5, 32, 32, 16, 1, 8),	// 1024
};

int sgpu_launch_lsort(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
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
	const LsortVariant *var = nullptr;
	for (const LsortVariant &v : VARIANTS) if ((v.rs << v.lt) >= d.nb_frames) { var = &v; break; }
	if (const char *e = getenv("SGPU_LSORT_ALT")) {	// alternates follow their primary in the table
		const int k = atoi(e);
		const ptrdiff_t nvar = (ptrdiff_t) (sizeof VARIANTS / sizeof VARIANTS[0]);
		if (var && k >= 0 && k <= 3 && (var - VARIANTS) + k < nvar) {
			const LsortVariant *alt = var + k;
			if (alt->lt == var->lt && alt->rs == var->rs && alt->re == var->re) var = alt;
		}
	}
	if (!var) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "no lane-sort variant for %d frames", d.nb_frames);
	// TMA tile load: f32 samples, no per-frame x shift, plane stride a multiple of 16 bytes
	CUtensorMap tmap;
	p.tile_rows = d.nb_frames; p.box_rows = d.nb_frames; p.use_tma = 0;
	const char *no_tma = getenv("SGPU_NO_TMA");
	if (d.dtype == SGPU_F32 && !a->shiftx && npix % 4 == 0 && ((uintptr_t) blk->dev & 15) == 0 && !(no_tma && no_tma[0] == '1')) {
		const int nbox = (d.nb_frames + 255) / 256;
		p.box_rows = (d.nb_frames + nbox - 1) / nbox;
		p.tile_rows = p.box_rows * nbox;
		if (sgpu_make_tile_tmap(ctx, &tmap, blk->dev, npix, d.nb_layers, d.nb_frames, TILE_PX, p.box_rows) == SGPU_OK) p.use_tma = 1;
		else { p.tile_rows = d.nb_frames; p.box_rows = d.nb_frames; }
	}
	const int pxc = var->sub * TILE_PX;
	const size_t shm = 1024 + (size_t) var->sub * TILE_PX * ((p.tile_rows + 7) & ~7) * sizeof(float) + (size_t) pxc * (sizeof(Sums) + 2 * sizeof(int)) + 16 + (size_t) d.nb_frames * 16 + (a->is_mean && a->type_of_rejection == SGPU_GESDT ? (size_t) pxc * SGPU_GESDT_WORDS * sizeof(unsigned) : 0);	// (the last term: coop_gesdt's decision bits)
	if (shm > ctx->smem_optin) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "stack too deep for the lane-sort engine");
	dim3 grid((unsigned) ((npix + pxc - 1) / pxc), (unsigned) d.nb_layers);
	rv = var->launch(ctx, p, p.use_tma ? &tmap : nullptr, grid, shm);
	if (rv) return rv;
	return sgpu_launch_exact_list(ctx, blk, p);
}
