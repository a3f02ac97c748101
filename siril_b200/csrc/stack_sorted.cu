// stack_sorted.cu — launcher of the sorted engine (kernel: stack_sorted.cuh): picks the variant
// (lanes per pixel, register slots per lane) for the stack depth, clears the fallback list,
// launches, then lets the exact engine redo the pixels on the list.
#include <cstdlib>
#include "stack_sorted.cuh"

using namespace sgpu;

// cuTensorMapEncodeTiled through the runtime's driver entry point (libcuda is not linked)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
		const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 3-D map of the frame block [frame][layer][pixel] (f32) with a (box_px, 1, box_rows) box and
// the 128-byte swizzle sw_idx() mirrors
int sgpu_make_tile_tmap(sgpu_ctx *ctx, CUtensorMap *tm, void *base, size_t npix, int layers, int frames, int box_px, int box_rows) {
	static EncodeTiledFn encode = nullptr;
	if (!encode) {
		void *fn = nullptr; cudaDriverEntryPointQueryResult q;
		if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) return SGPU_GENERIC_ERROR;
		encode = (EncodeTiledFn) fn;
	}
	const cuuint64_t dims[3] = { (cuuint64_t) npix, (cuuint64_t) layers, (cuuint64_t) frames };
	const cuuint64_t strides[2] = { (cuuint64_t) npix * 4, (cuuint64_t) npix * 4 * layers };
	const cuuint32_t box[3] = { (cuuint32_t) box_px, 1, (cuuint32_t) box_rows }, estr[3] = { 1, 1, 1 };
	const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
			CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	(void) ctx;
	return r == CUDA_SUCCESS ? SGPU_OK : SGPU_GENERIC_ERROR;
}

bool sgpu_sorted_supports(const sgpu_block_desc &d, const sgpu_stack_args *a) {
	if (d.nb_frames > 2048 || (size_t) d.height * d.width >= (1u << 30)) return false;
	if (!a->is_mean) return true;
	switch (a->type_of_rejection) {
	case SGPU_NO_REJEC: case SGPU_PERCENTILE: case SGPU_SIGMA: case SGPU_MAD: case SGPU_SIGMEDIAN: case SGPU_LINEARFIT: case SGPU_GESDT: return true;
	case SGPU_WINSORIZED: return d.dtype == SGPU_F32;
	default: return false;	// 16-bit WINSORIZED: exact engine
	}
}

// every compiled variant: lanes per pixel 2^lt, register slots per lane re, warps per CTA nw
struct SortedVariant { int lt, re, nw; int (*launch)(sgpu_ctx *, const StackParams &, const CUtensorMap *, dim3, size_t); };
static const SortedVariant VARIANTS[] = {
	{ 3, 4, 4, sgpu_launch_sorted_variant<3, 4, 4> },
	{ 3, 8, 4, sgpu_launch_sorted_variant<3, 8, 4> },
	{ 3, 13, 4, sgpu_launch_sorted_variant<3, 13, 4> },
	{ 3, 16, 4, sgpu_launch_sorted_variant<3, 16, 4> },
	{ 3, 20, 4, sgpu_launch_sorted_variant<3, 20, 4> },
	{ 3, 25, 4, sgpu_launch_sorted_variant<3, 25, 4> },
	{ 3, 32, 4, sgpu_launch_sorted_variant<3, 32, 4> },
	{ 3, 38, 4, sgpu_launch_sorted_variant<3, 38, 4> },
	{ 3, 44, 4, sgpu_launch_sorted_variant<3, 44, 4> },
	{ 3, 50, 4, sgpu_launch_sorted_variant<3, 50, 4> },
	{ 3, 57, 4, sgpu_launch_sorted_variant<3, 57, 4> },
	{ 3, 64, 4, sgpu_launch_sorted_variant<3, 64, 4> },
	{ 4, 7, 8, sgpu_launch_sorted_variant<4, 7, 8> },
	{ 4, 13, 8, sgpu_launch_sorted_variant<4, 13, 8> },
	{ 4, 19, 4, sgpu_launch_sorted_variant<4, 19, 4> },
	{ 4, 19, 8, sgpu_launch_sorted_variant<4, 19, 8> },
	{ 4, 25, 8, sgpu_launch_sorted_variant<4, 25, 8> },
	{ 4, 38, 4, sgpu_launch_sorted_variant<4, 38, 4> },
	{ 4, 44, 4, sgpu_launch_sorted_variant<4, 44, 4> },
	{ 4, 50, 4, sgpu_launch_sorted_variant<4, 50, 4> },
	{ 4, 57, 4, sgpu_launch_sorted_variant<4, 57, 4> },
	{ 4, 64, 4, sgpu_launch_sorted_variant<4, 64, 4> },
	{ 5, 38, 4, sgpu_launch_sorted_variant<5, 38, 4> },
	{ 5, 44, 4, sgpu_launch_sorted_variant<5, 44, 4> },
	{ 5, 50, 4, sgpu_launch_sorted_variant<5, 50, 4> },
	{ 5, 57, 4, sgpu_launch_sorted_variant<5, 57, 4> },
	{ 5, 64, 4, sgpu_launch_sorted_variant<5, 64, 4> }
};

// default choice: 8 lanes per pixel up to 512 frames, 16 up to 1024, 32 beyond; smallest slot count that
// holds the stack.  SGPU_SORT_VARIANT="lt,re,nw" (tuning aid) forces a variant that fits.
static int dispatch_sorted(sgpu_ctx *ctx, const StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm) {
	const int N = p.N;
	int flt = -1, fre = -1, fnw = -1;
	if (const char *e = getenv("SGPU_SORT_VARIANT")) sscanf(e, "%d,%d,%d", &flt, &fre, &fnw);
	const SortedVariant *best = nullptr;
	for (const SortedVariant &v : VARIANTS) {
		if ((v.re << v.lt) < N) continue;
		if (flt >= 0) { if (v.lt == flt && v.re == fre && v.nw == fnw) { best = &v; break; } continue; }
		const int lt = N <= 512 ? 3 : (N <= 1024 ? 4 : 5);
		if (v.lt != lt || v.nw != 4) continue;
		if (!best || v.re < best->re) best = &v;
	}
	if (!best) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "no sorted-engine variant for %d frames", N);
	return best->launch(ctx, p, tmap, grid, shm);
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
	ctx->list_used = true;
	p.is_u16 = d.dtype == SGPU_U16;
	p.need_sums = a->is_mean && a->type_of_rejection != SGPU_LINEARFIT;
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
	const size_t shm = 1024 + (size_t) TILE_PX * p.tile_rows * sizeof(float) + 2 * TILE_PX * sizeof(int) + TILE_PX * sizeof(Sums) + 16;
	if (shm > ctx->smem_optin) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "stack too deep for the sorted engine");
	dim3 grid((unsigned) ((npix + TILE_PX - 1) / TILE_PX), (unsigned) d.nb_layers);
	rv = dispatch_sorted(ctx, p, p.use_tma ? &tmap : nullptr, grid, shm);
	if (rv) return rv;
	return sgpu_launch_exact_list(ctx, blk, p);
}
