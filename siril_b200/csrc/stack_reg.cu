// stack_reg.cu — launcher of the register engine (kernel: stack_reg.cuh): picks the variant
// (lanes per pixel, register slots per lane) for the stack depth, clears the hand-over list,
// launches, then lets the exact engine redo the pixels on the list.
#include <cstdlib>
#include "stack_reg.cuh"

using namespace sgpu;

int sgpu_make_tile_tmap(sgpu_ctx *ctx, CUtensorMap *tm, void *base, size_t npix, int layers, int frames, int box_px, int box_rows);

bool sgpu_reg_supports(const sgpu_block_desc &d, const sgpu_stack_args *a) {
	if (d.dtype != SGPU_F32 || d.nb_frames < 32 || d.nb_frames > 512 || (size_t) d.height * d.width >= (1u << 30)) return false;
	if (a->shiftx || ((size_t) d.height * d.width) % 4 != 0) return false;	// x shifts / planes that TMA cannot address: other engines
	if (!a->is_mean) return true;
	if (a->weights) return false;	// the weighted combine re-reads the original stack: other engines
	switch (a->type_of_rejection) {
	case SGPU_NO_REJEC: case SGPU_PERCENTILE: case SGPU_SIGMA: case SGPU_WINSORIZED: return true;
	default: return false;
	}
}

// every compiled variant: 2^lt lanes per pixel, re register slots per lane, the first rf rows
// always full (the variant serves rf * 2^lt < N <= re * 2^lt)
struct RegVariant { int lt, re, rf; int (*launch)(sgpu_ctx *, const StackParams &, const CUtensorMap *, dim3, size_t); int (*launch_w)(sgpu_ctx *, const StackParams &, const CUtensorMap *, dim3, size_t); };
static const RegVariant VARIANTS[] = {
	{ 2, 25, 8, sgpu_launch_reg_variant<2, 25, 8, 6, false>, sgpu_launch_reg_variant<2, 25, 8, 6, true> },	// 32 < N <= 100
	{ 2, 38, 25, sgpu_launch_reg_variant<2, 38, 25, 6, false>, sgpu_launch_reg_variant<2, 38, 25, 6, true> },	// .. 152
	{ 2, 64, 38, sgpu_launch_reg_variant<2, 64, 38, 4, false>, sgpu_launch_reg_variant<2, 64, 38, 4, true> },	// .. 256
	{ 3, 38, 32, sgpu_launch_reg_variant<3, 38, 32, 3, false>, sgpu_launch_reg_variant<3, 38, 32, 3, true> },	// 256 < N <= 304
	{ 3, 50, 38, sgpu_launch_reg_variant<3, 50, 38, 3, false>, sgpu_launch_reg_variant<3, 50, 38, 3, true> },	// .. 400
	{ 3, 64, 50, sgpu_launch_reg_variant<3, 64, 50, 2, false>, sgpu_launch_reg_variant<3, 64, 50, 2, true> },	// .. 512
};

int sgpu_launch_reg(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
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
	p.is_u16 = 0;
	const RegVariant *var = nullptr;
	for (const RegVariant &v : VARIANTS) if ((v.re << v.lt) >= d.nb_frames && (v.rf << v.lt) <= d.nb_frames) { var = &v; break; }
	if (!var) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "no register-engine variant for %d frames", d.nb_frames);
	// TMA tile load: no per-frame x shift, plane stride a multiple of 16 bytes
	CUtensorMap tmap;
	p.tile_rows = d.nb_frames; p.box_rows = d.nb_frames; p.use_tma = 0;
	const char *no_tma = getenv("SGPU_NO_TMA");
	if (!a->shiftx && npix % 4 == 0 && ((uintptr_t) blk->dev & 15) == 0 && !(no_tma && no_tma[0] == '1')) {
		const int nbox = (d.nb_frames + 255) / 256;
		p.box_rows = (d.nb_frames + nbox - 1) / nbox;
		p.tile_rows = p.box_rows * nbox;
		if (sgpu_make_tile_tmap(ctx, &tmap, blk->dev, npix, d.nb_layers, d.nb_frames, TILE_PX, p.box_rows) == SGPU_OK) p.use_tma = 1;
		else { p.tile_rows = d.nb_frames; p.box_rows = d.nb_frames; }
	}
	const size_t coef_bytes = (p.use_tma && a->normalize != SGPU_NO_NORM) ? (size_t) d.nb_frames * 16 : 0;
	const size_t shm = 1024 + (size_t) TILE_PX * p.tile_rows * sizeof(float) + (size_t) TILE_PX * 2 * REG_WCAP * sizeof(float) + coef_bytes + 16;
	if (shm > ctx->smem_optin) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "stack too deep for the register engine");
	dim3 grid((unsigned) ((npix + TILE_PX - 1) / TILE_PX), (unsigned) d.nb_layers);
	if (!p.use_tma) {	// no tensor map (unaligned block, TMA switched off): the lane-sort engine has a plain loader
		if (sgpu_lsort_supports(d, a)) return sgpu_launch_lsort(ctx, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "the register engine needs the TMA tile load");
	}
	rv = (a->is_mean && a->type_of_rejection == SGPU_WINSORIZED ? var->launch_w : var->launch)(ctx, p, &tmap, grid, shm);
	if (rv) return rv;
	return sgpu_launch_exact_list(ctx, blk, p);
}
