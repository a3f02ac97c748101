// stack_psel.cu — launcher of the selection engine (kernel: stack_psel.cuh): picks the variant
// (lanes per pixel) for the stack depth, sizes shared memory, clears the hand-over list, launches,
// then lets the exact engine redo the list.
#include <cstdlib>
#include <mutex>
#include "stack_psel.cuh"

using namespace sgpu;

int sgpu_launch_psel_finish(sgpu_ctx *ctx, const sgpu::StackParams &p, unsigned chunk_px);
int sgpu_make_tile_tmap(sgpu_ctx *ctx, CUtensorMap *tm, void *base, size_t npix, int layers, int frames, int box_px, int box_rows);

// 2^lt lanes per pixel; serves nmin <= N <= nmax (about 40 to 80 samples per lane); sub sub-tiles of 32 px per CTA
struct PselVariant { int lt, nmin, nmax, sub; int (*launch)(sgpu_ctx *, const StackParams &, const CUtensorMap *, dim3, size_t); };
#define PV(LT, NMIN, NMAX, SUB) { LT, NMIN, NMAX, SUB, sgpu_launch_psel_variant<LT, SUB> }
static const PselVariant VARIANTS[] = { PV(0, 41, 80, 4), PV(1, 81, 160, 2), PV(2, 161, 320, 1), PV(3, 321, 640, 1), PV(4, 641, 1280, 1) };

static const PselVariant *pick(int N) {
	if (const char *e = getenv("SGPU_PSEL_LT")) {	// tuning aid: force the group size (any depth the kernel's loops can walk)
		const int lt = atoi(e);
		for (const PselVariant &v : VARIANTS) if (v.lt == lt && N >= 41 && N <= 1280) return &v;
	}
	for (const PselVariant &v : VARIANTS) if (N >= v.nmin && N <= v.nmax) return &v;
	return nullptr;
}

// correctly rounded reciprocals 1 / i for the B2 arithmetic (psel_rcp), filled once per device
#define PSEL_RCP_N SGPU_RCP_N
static __device__ double g_psel_rcp[PSEL_RCP_N];
const double *sgpu_rcp_table(sgpu_ctx *ctx) {
	static std::mutex mu;
	static const double *addr[64];
	if (ctx->device < 0 || ctx->device >= 64) return nullptr;
	std::lock_guard<std::mutex> lk(mu);
	if (!addr[ctx->device]) {
		static double host[PSEL_RCP_N];
		host[0] = 0.0;
		for (int i = 1; i < PSEL_RCP_N; i++) host[i] = 1.0 / (double) i;
		void *p = nullptr;
		if (cudaGetSymbolAddress(&p, g_psel_rcp) != cudaSuccess) return nullptr;
		if (cudaMemcpy(p, host, sizeof host, cudaMemcpyHostToDevice) != cudaSuccess) return nullptr;
		addr[ctx->device] = (const double *) p;
	}
	return addr[ctx->device];
}

bool sgpu_psel_supports(const sgpu_block_desc &d, const sgpu_stack_args *a) {
	if ((size_t) d.height * d.width >= (1u << 30) || !pick(d.nb_frames)) return false;
	if (!a->is_mean) return true;
	switch (a->type_of_rejection) {
	case SGPU_NO_REJEC: case SGPU_PERCENTILE: case SGPU_SIGMA: return true;
	default: return false;
	}
}

int sgpu_launch_psel(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej) {
	const sgpu_block_desc &d = blk->d;
	const size_t npix = (size_t) d.height * d.width;
	if (npix == 0) return SGPU_OK;
	const PselVariant *var = pick(d.nb_frames);
	if (!var) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "no selection-engine variant for %d frames", d.nb_frames);
	StackParams p;
	sgpu_fill_stack_params(p, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	const size_t list_bytes = ((npix * d.nb_layers + 1) * sizeof(unsigned) + 255) & ~(size_t) 255;
	int rv = sgpu_ensure_list(ctx, list_bytes); if (rv) return rv;
	p.list_count = (unsigned *) ctx->d_list; p.list = p.list_count + 1;
	SGPU_CUDA(ctx, cudaMemsetAsync(p.list_count, 0, sizeof(unsigned), ctx->stream));
	ctx->list_used = true;
	p.is_u16 = d.dtype == SGPU_U16;
	{ const char *e = getenv("SGPU_PSEL_STOP"); p.dbg_stop = e ? atoi(e) : 0; }
	p.rcp_tab = sgpu_rcp_table(ctx);	// null: the kernel computes the reciprocals
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
	const int T = 1 << var->lt, nt = var->sub * T * 32, pxc = var->sub * TILE_PX;
	const int rows8 = (p.tile_rows + 9) & ~7, lcap = psel_list_cap(T);
	const size_t shm = (size_t) var->sub * TILE_PX * rows8 * sizeof(float) + 16
		+ ((p.use_tma && a->normalize != SGPU_NO_NORM) ? (size_t) d.nb_frames * 16 : 0) + (size_t) lcap * nt * sizeof(float);
	size_t shm_pad = 0;
	if (const char *e = getenv("SGPU_PSEL_PAD")) shm_pad = (size_t) atol(e);	// tuning aid: extra shared memory per CTA (fewer CTAs per SM)
	if (shm + shm_pad > ctx->smem_optin) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "stack too deep for the selection engine");
	// Split form (default; SGPU_PSEL_SPLIT=0 keeps everything in the stack kernel): the stack kernel stops after the
	// extraction and psel_finish_kernel sorts and rejects, one thread per pixel.  The records between the two take
	// PSEL_REC_WORDS * 4 bytes per pixel of scratch; a block that needs more than half of the free memory is
	// processed in chunks of pixels.
	const char *se = getenv("SGPU_PSEL_SPLIT");
	p.split = !(se && se[0] == '0') && var->lt <= 2;	// deeper stacks extract more entries than a record holds
	p.layer_base = 0; p.pix_base = 0;
	if (p.split) {
		const size_t rec_bytes = (size_t) PSEL_REC_WORDS * sizeof(float);
		size_t chunk_px = npix;
		if (npix * rec_bytes > ctx->d_scratch_cap) {	// (only then: cudaMemGetInfo under load can hold the launches up for milliseconds)
			size_t freeb = 0, totalb = 0;
			SGPU_CUDA(ctx, cudaMemGetInfo(&freeb, &totalb));
			const size_t budget = ctx->d_scratch_cap + freeb / 2;
			if (chunk_px * rec_bytes > budget) chunk_px = budget / rec_bytes;
		}
		if (const char *ce = getenv("SGPU_PSEL_CHUNK")) { const size_t c = (size_t) atol(ce); if (c > 0 && c < chunk_px) chunk_px = c; }	// tests: force chunks
		if (chunk_px < npix) chunk_px = chunk_px / pxc * pxc;	// (a block that goes through whole keeps its ragged last tile: no second pair of launches for 16 pixels)
		if (chunk_px < (size_t) pxc) p.split = 0;	// no room for records: the one-kernel form
		else {
			rv = sgpu_ensure_scratch(ctx, chunk_px * rec_bytes);
			if (rv) return rv;
			p.records = (float *) ctx->d_scratch;
			const char *te = getenv("SGPU_PSEL_TIMING");
			const bool timing = te && te[0] == '1';
			for (int layer = 0; layer < d.nb_layers; layer++) {
				for (size_t base = 0; base < npix; base += chunk_px) {
					const size_t n = npix - base < chunk_px ? npix - base : chunk_px;
					p.layer_base = layer; p.pix_base = base;
					dim3 grid((unsigned) ((n + pxc - 1) / pxc), 1);
					cudaEvent_t ev[3] = { nullptr, nullptr, nullptr };
					if (timing) { for (auto &e : ev) cudaEventCreate(&e); cudaEventRecord(ev[0], ctx->stream); }
					rv = var->launch(ctx, p, p.use_tma ? &tmap : nullptr, grid, shm + shm_pad);
					if (rv) return rv;
					if (timing) cudaEventRecord(ev[1], ctx->stream);
					rv = sgpu_launch_psel_finish(ctx, p, (unsigned) n);
					if (rv) return rv;
					if (timing) {	// SGPU_PSEL_TIMING=1: the two kernels' durations, on stderr (serializes the stream)
						cudaEventRecord(ev[2], ctx->stream); cudaEventSynchronize(ev[2]);
						float a = 0.f, b = 0.f; cudaEventElapsedTime(&a, ev[0], ev[1]); cudaEventElapsedTime(&b, ev[1], ev[2]);
						fprintf(stderr, "psel layer %d pixels [%zu, +%zu): stack kernel %.3f ms, finish kernel %.3f ms\n", layer, base, n, a, b);
						for (auto &e : ev) cudaEventDestroy(e);
					}
				}
			}
			p.layer_base = 0; p.pix_base = 0;
			return sgpu_launch_exact_list(ctx, blk, p);
		}
	}
	dim3 grid((unsigned) ((npix + pxc - 1) / pxc), (unsigned) d.nb_layers);
	rv = var->launch(ctx, p, p.use_tma ? &tmap : nullptr, grid, shm + shm_pad);
	if (rv) return rv;
	return sgpu_launch_exact_list(ctx, blk, p);
}
