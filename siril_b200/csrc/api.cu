// api.cu — C ABI of libsgpu.so (include/sgpu.h): context, frame blocks, the stacking entry
// points and the host-side scalar helpers (planner, factors, weights).
#include <cstdarg>
#include <cmath>
#include <new>
#include <cstdlib>
#include <functional>
#include <cfloat>
#include <climits>
#include "common.cuh"

int sgpu_stack_host_impl(sgpu_ctx *ctx, const sgpu_block_desc *desc, const std::function<int(sgpu_block *, int, int)> &upload_one,
		const sgpu_stack_args *a, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]);

int sgpu_fail(sgpu_ctx *ctx, int code, const char *fmt, ...) {
	char buf[512];
	va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
	if (ctx) ctx->err = buf;
	return code;
}

int sgpu_cuda_fail(sgpu_ctx *ctx, cudaError_t e, const char *what) {
	const int code = (e == cudaErrorMemoryAllocation) ? SGPU_ALLOC_ERROR : SGPU_GENERIC_ERROR;
	return sgpu_fail(ctx, code, "CUDA error %d (%s) in %s", (int) e, cudaGetErrorString(e), what);
}

int sgpu_ensure_scratch(sgpu_ctx *ctx, size_t bytes) {
	if (bytes <= ctx->d_scratch_cap) return SGPU_OK;
	if (ctx->d_scratch) { cudaStreamSynchronize(ctx->stream); cudaFree(ctx->d_scratch); ctx->d_scratch = nullptr; ctx->d_scratch_cap = 0; }
	SGPU_CUDA(ctx, cudaMalloc(&ctx->d_scratch, bytes));
	ctx->d_scratch_cap = bytes;
	return SGPU_OK;
}

int sgpu_ensure_list(sgpu_ctx *ctx, size_t bytes) {
	if (bytes <= ctx->d_list_cap) return SGPU_OK;
	if (ctx->d_list) { cudaStreamSynchronize(ctx->stream); cudaFree(ctx->d_list); ctx->d_list = nullptr; ctx->d_list_cap = 0; }
	SGPU_CUDA(ctx, cudaMalloc(&ctx->d_list, bytes));
	ctx->d_list_cap = bytes;
	return SGPU_OK;
}

extern "C" {

int sgpu_abi_version(void) { return SGPU_ABI_VERSION; }

int sgpu_ctx_create(int device, sgpu_ctx **out) {
	if (!out) return SGPU_GENERIC_ERROR;
	*out = nullptr;
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count <= 0 || device < 0 || device >= count) {
		fprintf(stderr, "sgpu: no usable CUDA device %d (%s); this module has no CPU fallback\n",
				device, e == cudaSuccess ? "device ordinal out of range" : cudaGetErrorString(e));
		return SGPU_GENERIC_ERROR;
	}
	sgpu_ctx *ctx = new (std::nothrow) sgpu_ctx();
	if (!ctx) return SGPU_ALLOC_ERROR;
	ctx->device = device;
	if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return SGPU_GENERIC_ERROR; }
	cudaDeviceProp prop;
	if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return SGPU_GENERIC_ERROR; }
	ctx->sm_count = prop.multiProcessorCount;
	ctx->smem_optin = prop.sharedMemPerBlockOptin;
	{ const char *e = getenv("SGPU_STATS_FULL_HIST"); ctx->stats_full_hist = e && *e && *e != '0'; }
	if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
	    cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return SGPU_GENERIC_ERROR; }
	ctx->stream = ctx->own_stream;
	ctx->d_params_cap = 1 << 20;
	if (cudaMalloc(&ctx->d_params, ctx->d_params_cap) != cudaSuccess ||
	    cudaMallocHost(&ctx->h_params, ctx->d_params_cap) != cudaSuccess) { sgpu_ctx_destroy(ctx); return SGPU_ALLOC_ERROR; }
	*out = ctx;
	return SGPU_OK;
}

void sgpu_ctx_destroy(sgpu_ctx *ctx) {
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	cudaDeviceSynchronize();
	if (ctx->d_params) cudaFree(ctx->d_params);
	if (ctx->h_params) cudaFreeHost(ctx->h_params);
	if (ctx->d_scratch) cudaFree(ctx->d_scratch);
	if (ctx->d_list) cudaFree(ctx->d_list);
	if (ctx->d_arena) cudaFree(ctx->d_arena);
	sgpu_free_stage(ctx);
	if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
	if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
	delete ctx;
}

int sgpu_ctx_set_stream(sgpu_ctx *ctx, void *s) {
	if (!ctx) return SGPU_GENERIC_ERROR;
	ctx->stream = (cudaStream_t) s;	/* NULL is CUDA's legacy default stream, as everywhere in the runtime API */
	return SGPU_OK;
}

int sgpu_ctx_synchronize(sgpu_ctx *ctx) {
	if (!ctx) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return SGPU_OK;
}

const char *sgpu_last_error(const sgpu_ctx *ctx) { return ctx ? ctx->err.c_str() : "no context"; }
uint64_t sgpu_launch_count(const sgpu_ctx *ctx) { return ctx ? ctx->launches : 0; }
int sgpu_last_engine(const sgpu_ctx *ctx) { return ctx ? ctx->last_engine : 0; }
uint64_t sgpu_stats_window_misses(const sgpu_ctx *ctx) { return ctx ? ctx->stats_window_misses : 0; }
uint64_t sgpu_handover_count(sgpu_ctx *ctx) {
	if (!ctx || !ctx->d_list || !ctx->list_used) return 0;
	unsigned n = 0;
	if (cudaSetDevice(ctx->device) != cudaSuccess) return 0;
	if (cudaMemcpyAsync(&n, ctx->d_list, sizeof n, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) return 0;
	if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) return 0;
	return n;
}

int sgpu_ctx_set_cancel_callback(sgpu_ctx *ctx, int (*cb)(void *), void *user) {
	if (!ctx) return SGPU_GENERIC_ERROR;
	ctx->should_cancel = cb; ctx->cancel_user = user;
	return SGPU_OK;
}

int sgpu_mem_info(sgpu_ctx *ctx, size_t *free_bytes, size_t *total_bytes) {
	if (!ctx || !free_bytes || !total_bytes) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	SGPU_CUDA(ctx, cudaMemGetInfo(free_bytes, total_bytes));
	return SGPU_OK;
}

int sgpu_device_alloc(sgpu_ctx *ctx, size_t bytes, void **ptr) {
	if (!ctx || !ptr) return SGPU_GENERIC_ERROR;
	*ptr = nullptr;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	cudaError_t e = cudaMalloc(ptr, bytes ? bytes : 16);
	if (e != cudaSuccess) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
	e = cudaMemsetAsync(*ptr, 0, bytes ? bytes : 16, ctx->stream);
	if (e != cudaSuccess) { cudaFree(*ptr); *ptr = nullptr; return sgpu_cuda_fail(ctx, e, "cudaMemsetAsync"); }
	return SGPU_OK;
}
void sgpu_device_free(sgpu_ctx *ctx, void *ptr) { if (ptr) { if (ctx) cudaStreamSynchronize(ctx->stream); cudaFree(ptr); } }

int sgpu_download(sgpu_ctx *ctx, void *host_dst, const void *dev_src, size_t bytes) {
	if (!ctx || !host_dst || !dev_src) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	return SGPU_OK;
}

int sgpu_ipc_export(sgpu_ctx *ctx, void *dev_ptr, unsigned char handle[SGPU_IPC_HANDLE_BYTES]) {
	static_assert(sizeof(cudaIpcMemHandle_t) == SGPU_IPC_HANDLE_BYTES, "handle size");
	if (!ctx || !dev_ptr || !handle) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	cudaIpcMemHandle_t h;
	SGPU_CUDA(ctx, cudaIpcGetMemHandle(&h, dev_ptr));
	memcpy(handle, &h, sizeof h);
	return SGPU_OK;
}
int sgpu_ipc_open(sgpu_ctx *ctx, const unsigned char handle[SGPU_IPC_HANDLE_BYTES], void **peer_ptr) {
	if (!ctx || !handle || !peer_ptr) return SGPU_GENERIC_ERROR;
	*peer_ptr = nullptr;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	cudaIpcMemHandle_t h;
	memcpy(&h, handle, sizeof h);
	cudaError_t e = cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess);
	if (e != cudaSuccess) { *peer_ptr = nullptr; return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "cannot map the peer's image (cudaIpcOpenMemHandle: %s)", cudaGetErrorString(e)); }
	return SGPU_OK;
}
int sgpu_ipc_close(sgpu_ctx *ctx, void *peer_ptr) {
	if (!ctx) return SGPU_GENERIC_ERROR;
	if (!peer_ptr) return SGPU_OK;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	SGPU_CUDA(ctx, cudaIpcCloseMemHandle(peer_ptr));
	return SGPU_OK;
}

int sgpu_host_alloc(sgpu_ctx *ctx, size_t bytes, void **ptr) {
	if (!ctx || !ptr) return SGPU_GENERIC_ERROR;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	cudaError_t e = cudaMallocHost(ptr, bytes);
	if (e != cudaSuccess) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "pinned allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
	return SGPU_OK;
}
void sgpu_host_free(sgpu_ctx *ctx, void *ptr) { (void) ctx; if (ptr) cudaFreeHost(ptr); }

/* ---- blocks ---- */
static int check_desc(sgpu_ctx *ctx, const sgpu_block_desc *d) {
	if (!ctx || !d) return SGPU_GENERIC_ERROR;
	if (d->nb_frames < 1 || d->nb_layers < 1 || d->nb_layers > 3 || d->width < 1 || d->height < 0 ||
	    (d->dtype != SGPU_U16 && d->dtype != SGPU_F32))
		return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "bad block descriptor (%d frames, %d layers, %dx%d, dtype %d)",
				d->nb_frames, d->nb_layers, d->width, d->height, d->dtype);
	return SGPU_OK;
}

int sgpu_block_create(sgpu_ctx *ctx, const sgpu_block_desc *desc, sgpu_block **out) {
	int rv = check_desc(ctx, desc); if (rv) return rv;
	sgpu_block *b = new (std::nothrow) sgpu_block();
	if (!b) return SGPU_ALLOC_ERROR;
	b->ctx = ctx; b->d = *desc; b->owned = true;
	if (cudaSetDevice(ctx->device) != cudaSuccess) { delete b; return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "cudaSetDevice(%d) failed", ctx->device); }
	cudaError_t e = cudaMalloc(&b->dev, b->bytes() ? b->bytes() : 16);
	if (e != cudaSuccess) { delete b; return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "cannot allocate %zu bytes of HBM for the frame block: %s", b->bytes(), cudaGetErrorString(e)); }
	*out = b;
	return SGPU_OK;
}

int sgpu_block_wrap(sgpu_ctx *ctx, const sgpu_block_desc *desc, void *dev, sgpu_block **out) {
	int rv = check_desc(ctx, desc); if (rv) return rv;
	if (!dev) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "null device pointer");
	sgpu_block *b = new (std::nothrow) sgpu_block();
	if (!b) return SGPU_ALLOC_ERROR;
	b->ctx = ctx; b->d = *desc; b->owned = false; b->dev = dev;
	*out = b;
	return SGPU_OK;
}

void sgpu_block_destroy(sgpu_block *b) {
	if (!b) return;
	const bool own = (b->owned && b->dev) || (b->mask_owned && b->mask) || (b->drizz_owned && b->drizz);
	if (own) cudaStreamSynchronize(b->ctx->stream);
	if (b->owned && b->dev) cudaFree(b->dev);
	if (b->mask_owned && b->mask) cudaFree(b->mask);
	if (b->drizz_owned && b->drizz) cudaFree(b->drizz);
	delete b;
}

void *sgpu_block_device_ptr(sgpu_block *b) { return b ? b->dev : nullptr; }

int sgpu_block_upload_frame(sgpu_block *b, int frame, const void *host, int frame_height, int row0, int shifty) {
	if (!b || !host) return SGPU_GENERIC_ERROR;
	sgpu_ctx *ctx = b->ctx;
	const sgpu_block_desc &d = b->d;
	if (frame < 0 || frame >= d.nb_frames) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame %d out of range", frame);
	const size_t es = b->elem(), rowb = (size_t) d.width * es;
	// block row r <- frame row row0 + r - shifty, valid when inside [0, frame_height)
	long lo = (long) shifty - row0; if (lo < 0) lo = 0;			// first valid block row
	long hi = (long) frame_height + shifty - row0; if (hi > d.height) hi = d.height;	// one past last valid
	for (int l = 0; l < d.nb_layers; l++) {
		char *dst = (char *) b->dev + (((size_t) frame * d.nb_layers + l) * b->plane()) * es;
		const char *src = (const char *) host + (size_t) l * frame_height * rowb;
		if (hi <= lo) { SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) d.height * rowb, ctx->stream)); continue; }
		if (lo > 0) SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) lo * rowb, ctx->stream));
		if (hi < d.height) SGPU_CUDA(ctx, cudaMemsetAsync(dst + (size_t) hi * rowb, 0, (size_t) (d.height - hi) * rowb, ctx->stream));
		SGPU_CUDA(ctx, cudaMemcpyAsync(dst + (size_t) lo * rowb, src + (size_t) (row0 + lo - shifty) * rowb,
				(size_t) (hi - lo) * rowb, cudaMemcpyDefault, ctx->stream));
	}
	return SGPU_OK;
}

/* maximized framing (median_and_mean.c:160-190): the canvas that holds every registered frame, and the origin
 * args->offset; h02 / h12 are the frames' homography translations, rx / ry their sizes */
int sgpu_maximized_framing(int nb_frames, const int *rx, const int *ry, const double *h02, const double *h12, int upscale_at_stacking,
		long *naxes0, long *naxes1, int offset[2]) {
	if (!rx || !ry || !h02 || !h12 || !naxes0 || !naxes1 || !offset || nb_frames < 1) return SGPU_GENERIC_ERROR;
	const double scale = upscale_at_stacking ? 2. : 1.;
	double xmin = DBL_MAX, xmax = -DBL_MAX, ymin = DBL_MAX, ymax = -DBL_MAX;
	for (int i = 0; i < nb_frames; i++) {
		xmin = (xmin > h02[i] * scale) ? h02[i] * scale : xmin;
		ymin = (ymin > h12[i] * scale) ? h12[i] * scale : ymin;
		xmax = (xmax < h02[i] * scale + rx[i]) ? h02[i] * scale + rx[i] : xmax;
		ymax = (ymax < h12[i] * scale + ry[i]) ? h12[i] * scale + ry[i] : ymax;
	}
	*naxes0 = (int) xmax - (int) xmin + 1;
	*naxes1 = (int) ymax - (int) ymin + 1;
	offset[0] = (int) xmin;
	offset[1] = -(int) ymin;
	return SGPU_OK;
}

/* a frame of its own size into a block of the canvas' width (maximize_framing, variable-size sequences): what
 * seq_opened_read_region + rearrange_block_data (median_and_mean.c:363-380, :394-480) leave in data->pix[frame] — the
 * frame's rows at the LEFT of the block's rows, zero padded, the frame's TOP row on the canvas' top row before the y shift */
int sgpu_block_upload_frame_sized(sgpu_block *b, int frame, const void *host, int frame_width, int frame_height,
		int image_height, int row0, int shifty) {
	if (!b || !host) return SGPU_GENERIC_ERROR;
	sgpu_ctx *ctx = b->ctx;
	const sgpu_block_desc &d = b->d;
	if (frame < 0 || frame >= d.nb_frames) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame %d out of range", frame);
	if (frame_width < 1 || frame_width > d.width || frame_height < 1) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame of %d x %d does not fit a block %d wide", frame_width, frame_height, d.width);
	const size_t es = b->elem(), rowb = (size_t) d.width * es, frowb = (size_t) frame_width * es;
	// memory rows are bottom-up: canvas row r holds top-down row image_height - 1 - r, which the reference reads from the
	// frame's top-down row (image_height - 1 - r) + shifty, i.e. frame memory row r - shifty - (image_height - frame_height)
	const long sh = (long) shifty + ((long) image_height - frame_height);
	long lo = sh - row0; if (lo < 0) lo = 0;
	long hi = (long) frame_height + sh - row0; if (hi > d.height) hi = d.height;
	for (int l = 0; l < d.nb_layers; l++) {
		char *dst = (char *) b->dev + (((size_t) frame * d.nb_layers + l) * b->plane()) * es;
		const char *src = (const char *) host + (size_t) l * frame_height * frowb;
		SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) d.height * rowb, ctx->stream));
		if (hi <= lo) continue;
		SGPU_CUDA(ctx, cudaMemcpy2DAsync(dst + (size_t) lo * rowb, rowb, src + (size_t) (row0 + lo - sh) * frowb, frowb,
				frowb, (size_t) (hi - lo), cudaMemcpyDefault, ctx->stream));
	}
	return SGPU_OK;
}

/* ---- feather masks / drizzle weights: two more float plane sets of the block (median_and_mean.c:483-542) ---- */
static size_t weight_plane_floats(const sgpu_block *b, int which) {
	return (size_t) b->d.nb_frames * (which == SGPU_PLANE_DRIZZLE ? b->d.nb_layers : 1) * b->plane();
}

int sgpu_block_set_weight_planes(sgpu_block *b, float *d_mask, float *d_drizzle) {
	if (!b) return SGPU_GENERIC_ERROR;
	if ((b->mask_owned && b->mask) || (b->drizz_owned && b->drizz)) cudaStreamSynchronize(b->ctx->stream);
	if (b->mask_owned && b->mask) cudaFree(b->mask);
	if (b->drizz_owned && b->drizz) cudaFree(b->drizz);
	b->mask = d_mask; b->drizz = d_drizzle; b->mask_owned = b->drizz_owned = false;
	return SGPU_OK;
}

int sgpu_block_upload_weight_plane(sgpu_block *b, int which, int frame, const float *host, int frame_height, int row0, int shifty) {
	if (!b || !host) return SGPU_GENERIC_ERROR;
	sgpu_ctx *ctx = b->ctx;
	const sgpu_block_desc &d = b->d;
	if (which != SGPU_PLANE_MASK && which != SGPU_PLANE_DRIZZLE) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "unknown weight plane kind %d", which);
	if (frame < 0 || frame >= d.nb_frames) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame %d out of range", frame);
	float *&set = which == SGPU_PLANE_MASK ? b->mask : b->drizz;
	bool &owned = which == SGPU_PLANE_MASK ? b->mask_owned : b->drizz_owned;
	if (!set) {	// first upload: the plane set comes into being zeroed (the reference callocs its pool, :1402-1462)
		SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
		const size_t bytes = weight_plane_floats(b, which) * 4;
		cudaError_t e = cudaMalloc(&set, bytes ? bytes : 16);
		if (e != cudaSuccess) { set = nullptr; return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "cannot allocate %zu bytes of HBM for the %s planes: %s", bytes, which == SGPU_PLANE_MASK ? "mask" : "drizzle weight", cudaGetErrorString(e)); }
		owned = true;
		SGPU_CUDA(ctx, cudaMemsetAsync(set, 0, bytes ? bytes : 16, ctx->stream));
	}
	const int layers = which == SGPU_PLANE_DRIZZLE ? d.nb_layers : 1;
	const size_t rowb = (size_t) d.width * 4;
	long lo = (long) shifty - row0; if (lo < 0) lo = 0;
	long hi = (long) frame_height + shifty - row0; if (hi > d.height) hi = d.height;
	for (int l = 0; l < layers; l++) {
		char *dst = (char *) (set + ((size_t) frame * layers + l) * b->plane());
		const char *src = (const char *) host + (size_t) l * frame_height * rowb;
		if (hi <= lo) { SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) d.height * rowb, ctx->stream)); continue; }
		if (lo > 0) SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) lo * rowb, ctx->stream));
		if (hi < d.height) SGPU_CUDA(ctx, cudaMemsetAsync(dst + (size_t) hi * rowb, 0, (size_t) (d.height - hi) * rowb, ctx->stream));
		SGPU_CUDA(ctx, cudaMemcpyAsync(dst + (size_t) lo * rowb, src + (size_t) (row0 + lo - shifty) * rowb,
				(size_t) (hi - lo) * rowb, cudaMemcpyDefault, ctx->stream));
	}
	return SGPU_OK;
}

int sgpu_block_mask_ramp(sgpu_block *b, int feather_dist) {
	if (!b) return SGPU_GENERIC_ERROR;
	if (!b->mask) return sgpu_fail(b->ctx, SGPU_GENERIC_ERROR, "the block has no mask planes");
	if (feather_dist <= 0) return sgpu_fail(b->ctx, SGPU_GENERIC_ERROR, "feather distance must be positive");
	return sgpu_launch_mask_ramp(b->ctx, b->mask, weight_plane_floats(b, SGPU_PLANE_MASK), feather_dist);
}

/* ---- stacking ---- */
static int check_args(sgpu_ctx *ctx, const sgpu_block_desc &d, const sgpu_stack_args *a) {
	if (!a) return SGPU_GENERIC_ERROR;
	// median_and_mean.c:1279-1285
	if (d.nb_frames < 2) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "Select at least two frames for stacking. Aborting.");
	if (d.nb_frames < 3 && a->is_mean && a->type_of_rejection == SGPU_GESDT)
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "The Generalized Extreme Studentized Deviate Test needs at least three frames for stacking. Aborting.");
	if (a->type_of_rejection < SGPU_NO_REJEC || a->type_of_rejection > SGPU_GESDT) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "unknown rejection type %d", a->type_of_rejection);
	if (a->normalize < SGPU_NO_NORM || a->normalize > SGPU_MULTIPLICATIVE_SCALING) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "unknown normalization %d", a->normalize);
	if (a->normalize != SGPU_NO_NORM && (!a->scale || !a->offset || !a->mul)) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "normalization requested without coefficients");
	if (a->is_mean && a->type_of_rejection == SGPU_GESDT && !a->critical_value) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "GESDT needs critical values");
	if (a->is_mean && a->type_of_rejection == SGPU_GESDT && a->nb_critical > 0 && a->nb_critical < (int) floorf((float) d.nb_frames * a->sig[0]))
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "GESDT: %d critical values given, %d needed", a->nb_critical, (int) floorf((float) d.nb_frames * a->sig[0]));
	if (!a->use_32bit_output && d.dtype == SGPU_F32) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "16-bit output needs 16-bit input");
	if (a->out_plane_stride && a->out_plane_stride < (size_t) d.height * d.width)
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "output plane stride %zu is smaller than the block's plane (%d x %d)", a->out_plane_stride, d.width, d.height);
	return SGPU_OK;
}

// stage the small per-call arrays into the device parameter arena
static int stage_params(sgpu_ctx *ctx, const sgpu_block_desc &d, const sgpu_stack_args *a, StackDevParams &dp) {
	const size_t cn = (size_t) d.nb_layers * d.nb_frames;
	size_t off = 0;
	auto put = [&](const void *src, size_t bytes) -> const void * {
		if (!src) return nullptr;
		off = (off + 15) & ~(size_t) 15;
		if (off + bytes > ctx->d_params_cap) return nullptr;
		memcpy((char *) ctx->h_params + off, src, bytes);
		const void *dev = (char *) ctx->d_params + off;
		off += bytes;
		return dev;
	};
	// the pinned mirror is reused by every call: make sure the previous copy has left it
	SGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	const bool norm = a->normalize != SGPU_NO_NORM;
	dp.scale = (const double *) put(norm ? a->scale : nullptr, cn * 8);
	dp.offset = (const double *) put(norm ? a->offset : nullptr, cn * 8);
	dp.mul = (const double *) put(norm ? a->mul : nullptr, cn * 8);
	dp.weights = (const double *) put(a->is_mean ? a->weights : nullptr, cn * 8);
	dp.shiftx = (const int *) put(a->shiftx, (size_t) d.nb_frames * 4);
	int ncrit = 0;
	if (a->is_mean && a->type_of_rejection == SGPU_GESDT) ncrit = (int) floorf((float) d.nb_frames * a->sig[0]);	// int * float: a float product (median_and_mean.c:1479)
	dp.crit = (const float *) put(ncrit ? a->critical_value : nullptr, (size_t) ncrit * 4);
	if ((norm && (!dp.scale || !dp.offset || !dp.mul)) || (a->is_mean && a->weights && !dp.weights) ||
	    (a->shiftx && !dp.shiftx) || (ncrit && !dp.crit))
		return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "parameter arena too small for %d frames", d.nb_frames);
	if (off) SGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_params, ctx->h_params, off, cudaMemcpyHostToDevice, ctx->stream));
	return SGPU_OK;
}

int sgpu_stack_block(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej) {
	if (!ctx || !blk) return SGPU_GENERIC_ERROR;
	int rv = check_args(ctx, blk->d, a); if (rv) return rv;
	if (!d_out) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "null output");
	if (a->is_mean && a->create_rejmaps && (!d_rej_low || (!a->merge_lowhigh_rejmaps && !d_rej_high)))
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "rejection maps requested without buffers");
	if (a->is_mean && a->type_of_rejection != SGPU_NO_REJEC && !d_irej) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "null rejection counters");
	if (ctx->should_cancel && ctx->should_cancel(ctx->cancel_user)) return sgpu_fail(ctx, SGPU_CANCEL, "cancelled");
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	StackDevParams dp;
	rv = stage_params(ctx, blk->d, a, dp); if (rv) return rv;
	// engine: 0 automatic, 1 exact thread-per-pixel, 2 sorted or fail, 5 lane-sort or fail, 6 selection or fail
	// (3 and 4 were round 1's register and thread-sort engines: gone, the numbers stay reserved)
	if (a->engine == 3 || a->engine == 4) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "engine %d does not exist any more", a->engine);
	if (a->engine == 2 && !sgpu_sorted_supports(blk->d, a)) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "the sorted engine does not cover this configuration");
	ctx->list_used = false;
	sgpu_stack_args forced;
	const bool masked = a->is_mean && blk->mask;	// feather masks: per-sample weights in the combine — every engine but the selection engine carries them
	if (a->is_mean && blk->drizz) {
		// drizzle weights also decide what counts as a null sample before the rejection: served by the exact engine
		// (the others have no plane loads at that point); asking for another one is an error
		if (a->engine != 0 && a->engine != 1) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "drizzle weight planes are served by the exact engine only");
		forced = *a; forced.engine = 1; a = &forced;
	} else if (masked && a->engine == 6) {
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "the selection engine does not carry mask planes");
	} else
	if (a->engine == 0) if (const char *e = getenv("SGPU_ENGINE")) {	// tuning aid: force an engine from the environment
		forced = *a; forced.engine = atoi(e);
		if ((forced.engine == 2 && sgpu_sorted_supports(blk->d, a)) || (forced.engine == 5 && sgpu_lsort_supports(blk->d, a)) ||
		    (forced.engine == 6 && sgpu_psel_supports(blk->d, a)) || forced.engine == 1) a = &forced;
	}
	// automatic (by measurement on B200, profiles/): the selection engine for what it serves (no /
	// percentile / sigma rejection and the median, 64 frames and more), else
	static const int psel_min = getenv("SGPU_PSEL_MIN") ? atoi(getenv("SGPU_PSEL_MIN")) : 64;
	// (winsorized was tried in the selection engine: with a quarter of the stack beyond the clamp there is little to select —
	// C2 took 22.5 ms against the lane-sort engine's 16.4 — and it is not served there)
	if (((a->engine == 0 && blk->d.nb_frames >= psel_min && !masked) || a->engine == 6) && sgpu_psel_supports(blk->d, a))
		return ctx->last_engine = 6, sgpu_launch_psel(ctx, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	if (a->engine == 6) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "the selection engine does not cover this configuration");
	// the lane-sort engine wherever it applies (up to 1024 frames), else the sorted engine, else the
	// exact engine; the others only when asked for
	if ((a->engine == 0 || a->engine == 5) && sgpu_lsort_supports(blk->d, a))
		return ctx->last_engine = 5, sgpu_launch_lsort(ctx, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	if (a->engine == 5) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "the lane-sort engine does not cover this configuration");
	if (a->engine != 1 && sgpu_sorted_supports(blk->d, a))
		return ctx->last_engine = 2, sgpu_launch_sorted(ctx, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
	ctx->last_engine = 1;
	return sgpu_launch_exact(ctx, blk, a, dp, d_out, d_rej_low, d_rej_high, d_irej);
}

int sgpu_stack_host(sgpu_ctx *ctx, const sgpu_block_desc *desc, const void *const *frames,
		const int *shifty, const sgpu_stack_args *a,
		void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]) {
	if (!ctx || !desc || !frames || !out) return SGPU_GENERIC_ERROR;
	const int H = desc->height;
	return sgpu_stack_host_impl(ctx, desc, [&](sgpu_block *b, int f, int row0) {
		return sgpu_block_upload_frame(b, f, frames[f], H, row0, shifty ? shifty[f] : 0);
	}, a, out, rej_low, rej_high, irej);
}

}	// extern "C"

// the block loop behind sgpu_stack_host and sgpu_stack_fits: upload_one(block, frame, first image row)
// fills one frame of one row block on ctx->stream
int sgpu_stack_host_impl(sgpu_ctx *ctx, const sgpu_block_desc *desc, const std::function<int(sgpu_block *, int, int)> &upload_one,
		const sgpu_stack_args *a, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]) {
	if (!ctx || !desc || !out) return SGPU_GENERIC_ERROR;
	int rv = check_args(ctx, *desc, a); if (rv) return rv;
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	const int H = desc->height, W = desc->width, C = desc->nb_layers;
	const size_t npx = (size_t) C * H * W;
	const size_t in_es = desc->dtype == SGPU_F32 ? 4 : 2, out_es = a->use_32bit_output ? 4 : 2;
	const int nmaps = (a->is_mean && a->create_rejmaps) ? (a->merge_lowhigh_rejmaps ? 1 : 2) : 0;
	const bool renorm = a->use_32bit_output && a->output_norm;	// rescaled with the whole image's range once every block is in
	const size_t out_bytes = (npx * out_es + 255) & ~(size_t) 255, map_bytes = (npx * 2 + 255) & ~(size_t) 255;
	const size_t fixed = out_bytes + nmaps * map_bytes + 256;
	// Row blocks, as the reference's own block loop (median_and_mean.c:1552; rows per block from the
	// memory budget, stack_get_max_number_of_rows :1234): NBUF device buffers take turns, block k + 1 is
	// on its way over PCIe (copy stream) while block k is stacked and its rows go back (context stream),
	// so the device holds NBUF blocks and the outputs whatever the size of the sequence.
	const size_t row_bytes = (size_t) C * W * in_es * (size_t) desc->nb_frames;	// one image row of every frame and layer
	const size_t in_bytes = row_bytes * (size_t) H;
	int nblk = 1;
	if (in_bytes >= ((size_t) 256 << 20) && H >= 512) nblk = 8;
	{
		size_t fr = 0, tot = 0;
		SGPU_CUDA(ctx, cudaMemGetInfo(&fr, &tot));
		fr += ctx->d_arena_cap;		// what the cached arena holds is ours to reuse
		const size_t budget = fr > fixed + ((size_t) 512 << 20) ? (size_t) ((fr - fixed - ((size_t) 512 << 20)) * 0.9) : 0;
		const int nbuf = nblk > 1 ? 3 : 1;
		while ((size_t) nbuf * (((size_t) H + nblk - 1) / nblk + 4) * row_bytes > budget && nblk < H) nblk = nblk < 4 ? 4 : nblk * 2;
	}
	if (const char *e = getenv("SGPU_HOST_BLOCKS")) { const int n = atoi(e); if (n >= 1) nblk = n; }
	if (nblk > H) nblk = H > 0 ? H : 1;
	int rows = (H + nblk - 1) / nblk;
	if (nblk > 1) rows = (rows + 3) & ~3;		// keeps h * W a multiple of 4 samples (TMA tile load)
	if (rows < 1) rows = 1;
	nblk = H > 0 ? (H + rows - 1) / rows : 1;
	const int NBUF = nblk < 3 ? nblk : 3;
	const size_t blk_bytes = ((size_t) rows * row_bytes + 255) & ~(size_t) 255;
	// one cached device arena: [NBUF blocks | result | rejection maps | counters]
	const size_t need = (size_t) NBUF * blk_bytes + fixed;
	if (need > ctx->d_arena_cap) {
		if (ctx->d_arena) { cudaStreamSynchronize(ctx->stream); cudaFree(ctx->d_arena); ctx->d_arena = nullptr; ctx->d_arena_cap = 0; }
		cudaError_t e = cudaMalloc(&ctx->d_arena, need);
		if (e != cudaSuccess) return sgpu_fail(ctx, SGPU_ALLOC_ERROR, "cannot allocate %zu bytes of HBM for %d row blocks and the outputs: %s", need, NBUF, cudaGetErrorString(e));
		ctx->d_arena_cap = need;
	}
	char *d_buf = (char *) ctx->d_arena + (size_t) NBUF * blk_bytes;
	std::vector<cudaEvent_t> up(nblk, nullptr), done(nblk, nullptr);
	std::vector<sgpu_block *> blks(nblk, nullptr);
	cudaStream_t const compute = ctx->stream;
	auto cleanup = [&](int code) {
		ctx->stream = compute;
		cudaStreamSynchronize(ctx->copy_stream); cudaStreamSynchronize(compute);
		for (sgpu_block *b : blks) if (b) sgpu_block_destroy(b);
		for (cudaEvent_t ev : up) if (ev) cudaEventDestroy(ev);
		for (cudaEvent_t ev : done) if (ev) cudaEventDestroy(ev);
		return code;
	};
	for (int k = 0; k < nblk; k++) {
		sgpu_block_desc dk = *desc;
		dk.height = (k + 1) * rows <= H ? rows : H - k * rows;
		rv = sgpu_block_wrap(ctx, &dk, (char *) ctx->d_arena + (size_t) (k % NBUF) * blk_bytes, &blks[k]); if (rv) return cleanup(rv);
		if (cudaEventCreateWithFlags(&up[k], cudaEventDisableTiming) != cudaSuccess ||
		    cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
	}
	uint64_t *d_irej = (uint64_t *) (d_buf + out_bytes + nmaps * map_bytes);
	if (cudaMemsetAsync(d_irej, 0, 48, compute) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
	// uploads run on the copy stream when there is something to overlap with; a buffer is filled
	// again only once the stack of the block that used it has run
	auto upload = [&](int k) -> int {
		if (nblk > 1) ctx->stream = ctx->copy_stream;
		int r = SGPU_OK;
		if (k >= NBUF && cudaStreamWaitEvent(ctx->stream, done[k - NBUF], 0) != cudaSuccess) r = SGPU_GENERIC_ERROR;
		for (int f = 0; f < desc->nb_frames && !r; f++) r = upload_one(blks[k], f, k * rows);
		if (!r && cudaEventRecord(up[k], ctx->stream) != cudaSuccess) r = SGPU_GENERIC_ERROR;
		ctx->stream = compute;
		return r;
	};
	rv = upload(0); if (rv) return cleanup(rv);
	for (int k = 0; k < nblk; k++) {
		if (k + 1 < nblk) { rv = upload(k + 1); if (rv) return cleanup(rv); }
		if (cudaStreamWaitEvent(compute, up[k], 0) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
		// block k's outputs: [layer][h][W] at the offset of everything the blocks before it wrote
		const int h = blks[k]->d.height, row0 = k * rows;
		const size_t off = (size_t) C * row0 * W;
		char *d_out = d_buf + off * out_es;
		uint16_t *d_lo = nmaps >= 1 ? (uint16_t *) (d_buf + out_bytes) + off : nullptr;
		uint16_t *d_hi = nmaps == 2 ? (uint16_t *) (d_buf + out_bytes + map_bytes) + off : nullptr;
		sgpu_stack_args dense = *a; dense.out_plane_stride = 0;	// the row blocks' outputs are dense staging planes
		rv = sgpu_stack_block(ctx, blks[k], &dense, d_out, d_lo, d_hi, d_irej);
		if (rv) return cleanup(rv);
		if (cudaEventRecord(done[k], compute) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
		for (int l = 0; l < C; l++) {
			const size_t src = (size_t) l * h * W, dst = ((size_t) l * H + row0) * W, n = (size_t) h * W;
			if (!renorm && cudaMemcpyAsync((char *) out + dst * out_es, d_out + src * out_es, n * out_es, cudaMemcpyDeviceToHost, compute) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
			if (d_lo && rej_low && cudaMemcpyAsync(rej_low + dst, d_lo + src, n * 2, cudaMemcpyDeviceToHost, compute) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
			if (d_hi && rej_high && cudaMemcpyAsync(rej_high + dst, d_hi + src, n * 2, cudaMemcpyDeviceToHost, compute) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
		}
	}
	if (renorm) {	// norm_to_0_1_range (median_and_mean.c:557-582) over the whole result, which stayed on the device.
		// The blocks lie one after the other, [block][layer][h][W]: the scan order differs from the image's but
		// min / max do not care; the reference skips element 0 of the IMAGE, which is element 0 of block 0
		float mm[2];
		rv = sgpu_minmax_nonzero(ctx, (const float *) d_buf, npx, 1, mm); if (rv) return cleanup(rv);
		rv = sgpu_rescale_0_1(ctx, (float *) d_buf, npx, mm[0], mm[1]); if (rv) return cleanup(rv);
		for (int k = 0; k < nblk; k++) {
			const int h = blks[k]->d.height, row0 = k * rows;
			const size_t off = (size_t) C * row0 * W;
			for (int l = 0; l < C; l++) {
				const size_t src = off + (size_t) l * h * W, dst = ((size_t) l * H + row0) * W, n = (size_t) h * W;
				if (cudaMemcpyAsync((char *) out + dst * 4, d_buf + src * 4, n * 4, cudaMemcpyDeviceToHost, compute) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
			}
		}
	}
	uint64_t h_irej[6] = { 0 };
	if (cudaMemcpyAsync(h_irej, d_irej, 48, cudaMemcpyDeviceToHost, compute) != cudaSuccess) return cleanup(SGPU_GENERIC_ERROR);
	cudaError_t e = cudaStreamSynchronize(compute);
	if (e != cudaSuccess) return cleanup(sgpu_cuda_fail(ctx, e, "stack"));
	if (irej) memcpy(irej, h_irej, 48);
	return cleanup(SGPU_OK);
}

extern "C" {

/* ---- host-side scalar helpers ---- */

/* _compute_estimators_for_image (normalization.c:98-141): which statistic feeds which array */
int sgpu_estimators_from_stats(int normalize, int lite, const sgpu_imstats *st, double *offset, double *mul, double *scale) {
	if (!st || !offset || !mul || !scale) return SGPU_GENERIC_ERROR;
	*offset = 0.0; *mul = 1.0; *scale = 1.0;	/* init_coeffs, normalization.c:81-96 */
	switch (normalize) {
	default:
	case SGPU_ADDITIVE_SCALING: *scale = lite ? 1.5 * st->mad : st->scale; /* fallthrough */
	case SGPU_ADDITIVE: *offset = lite ? st->median : st->location; break;
	case SGPU_MULTIPLICATIVE_SCALING: *scale = lite ? 1.5 * st->mad : st->scale; /* fallthrough */
	case SGPU_MULTIPLICATIVE: *mul = lite ? st->median : st->location; break;
	}
	return SGPU_OK;
}

/* compute_factors_from_estimators (normalization.c:144-181) */
int sgpu_compute_factors(int normalize, int nb_layers, int N, int ref_index, int reglayer, int equalizeRGB,
		double *offset, double *mul, double *scale) {
	if (!offset || !mul || !scale || nb_layers < 1 || nb_layers > 3 || ref_index < 0 || ref_index >= N) return SGPU_GENERIC_ERROR;
	double o0[3], m0[3], s0[3];
	for (int l = 0; l < nb_layers; l++) {
		o0[l] = offset[l * N + ref_index]; m0[l] = mul[l * N + ref_index]; s0[l] = scale[l * N + ref_index];
	}
	const int rl = reglayer > -1 ? reglayer : 1;
	for (int l = 0; l < nb_layers; l++) {
		const int ref = equalizeRGB ? rl : l;
		for (int i = 0; i < N; i++) {
			double &ps = scale[l * N + i], &po = offset[l * N + i], &pm = mul[l * N + i];
			if (normalize == SGPU_ADDITIVE || normalize == SGPU_MULTIPLICATIVE) {
				if (normalize == SGPU_ADDITIVE) po = ps * po - o0[ref];
				else pm = (pm == 0) ? 1 : m0[ref] / pm;
			} else if (normalize == SGPU_MULTIPLICATIVE_SCALING) {
				ps = (ps == 0) ? 1 : s0[ref] / ps;
				pm = (pm == 0) ? 1 : m0[ref] / pm;
			} else {	/* ADDITIVE_SCALING and the switch's default */
				ps = (ps == 0) ? 1 : s0[ref] / ps;
				po = ps * po - o0[ref];
			}
		}
	}
	return SGPU_OK;
}

/* compute_noise_weights (median_and_mean.c:1111-1135) */
int sgpu_noise_weights(int nb_layers, int N, const double *scale, const double *bgnoise, double *weights) {
	if (!scale || !bgnoise || !weights) return SGPU_GENERIC_ERROR;
	for (int l = 0; l < nb_layers; l++) {
		double norm = 0.0;
		for (int i = 0; i < N; i++) {
			const double s = scale[l * N + i], b = bgnoise[l * N + i];
			weights[l * N + i] = 1.f / (s * s * b * b);
			norm += weights[l * N + i];
		}
		norm /= (double) N;
		for (int i = 0; i < N; i++) weights[l * N + i] /= norm;
	}
	return SGPU_OK;
}

/* the statistics cache's record in a .seq file: written by writeseqfile (io/seqfile.c:838-853), read back by readseqfile
 * (:544-611).  The reference prints with %lg — six significant digits — so a second run works from the ROUNDED numbers;
 * the same format keeps that behaviour. */
int sgpu_imstats_seq_record(int layer, int image_index, const sgpu_imstats *st, char *line, size_t cap) {
	if (!st || !line || layer < 0 || layer > 9) return SGPU_GENERIC_ERROR;
	const int n = snprintf(line, cap, "M%c-%d %ld %ld %lg %lg %lg %lg %lg %lg %lg %lg %lg %lg %lg %lg\n", '0' + layer, image_index,
			(long) st->total, (long) st->ngoodpix, st->mean, st->median, st->sigma, st->avgDev, st->mad, st->sqrtbwmv,
			st->location, st->scale, st->min, st->max, st->normValue, st->bgnoise);
	return (n < 0 || (size_t) n >= cap) ? SGPU_GENERIC_ERROR : SGPU_OK;
}

int sgpu_imstats_from_seq_record(const char *line, int *layer, int *image_index, sgpu_imstats *st) {
	if (!line || !layer || !image_index || !st) return SGPU_GENERIC_ERROR;
	if (line[0] != 'M' || line[1] == '\0') return SGPU_SEQUENCE_ERROR;
	const int cur = line[1] == '*' ? 0 : line[1] - '0';	// '*': statistics of the CFA channel (:549-560)
	if (cur < 0 || cur > 9 || line[2] != '-') return SGPU_SEQUENCE_ERROR;	// "sequence file format error"
	long total, ngood; int image;
	sgpu_imstats t;
	const int tokens = sscanf(line + 3, "%d %ld %ld %lg %lg %lg %lg %lg %lg %lg %lg %lg %lg %lg %lg", &image, &total, &ngood,
			&t.mean, &t.median, &t.sigma, &t.avgDev, &t.mad, &t.sqrtbwmv, &t.location, &t.scale, &t.min, &t.max, &t.normValue, &t.bgnoise);
	if (tokens != 15) return SGPU_SEQUENCE_ERROR;
	t.total = (double) total; t.ngoodpix = (double) ngood;
	*st = t; *layer = cur; *image_index = image;
	return SGPU_OK;
}

/* solve_overlap_coeffs (normalization.c:296-355): the (nb_frames - 1)^2 system of one layer, then LU with partial pivoting */
static int solve_overlap_coeffs(int nb_frames, const int *index, int index_ref, const size_t *Nij, const double *Mij, bool additive, double *coeffs) {
	const int N = nb_frames - 1;
	std::vector<double> A((size_t) N * N, 0.0), B(N, 0.0);
	auto n_ = [&](int a, int b) { return (double) Nij[(size_t) a * nb_frames + b]; };
	auto m_ = [&](int a, int b) { return Mij[(size_t) a * nb_frames + b]; };
	int c = 0;
	for (int i = 0; i < N; i++) {
		const int ii = index[i];
		B[i] = additive ? n_(ii, index_ref) * (m_(index_ref, ii) - m_(ii, index_ref)) : n_(ii, index_ref) * m_(index_ref, ii) * m_(ii, index_ref);
		for (int j = 0; j < N; j++) {
			const int ij = index[j];
			if (ii == ij) {
				for (int k = 0; k < nb_frames; k++)
					if (k != ii) A[c] += additive ? n_(ii, k) : n_(ii, k) * m_(ii, k) * m_(ii, k);
			} else {
				A[c] = additive ? -n_(ii, ij) : -n_(ii, ij) * m_(ii, ij) * m_(ij, ii);
				if (additive) B[i] += Nij[(size_t) ii * nb_frames + ij] * (m_(ij, ii) - m_(ii, ij));
			}
			c++;
		}
	}
	// LU decomposition with partial (row) pivoting, then forward / back substitution
	std::vector<int> perm(N);
	for (int i = 0; i < N; i++) perm[i] = i;
	for (int j = 0; j < N; j++) {
		int piv = j; double best = fabs(A[(size_t) j * N + j]);
		for (int i = j + 1; i < N; i++) if (fabs(A[(size_t) i * N + j]) > best) { best = fabs(A[(size_t) i * N + j]); piv = i; }
		if (piv != j) { for (int k = 0; k < N; k++) std::swap(A[(size_t) j * N + k], A[(size_t) piv * N + k]); std::swap(perm[j], perm[piv]); }
		const double ajj = A[(size_t) j * N + j];
		if (ajj == 0.0) return 1;	// singular: frames that overlap nothing
		for (int i = j + 1; i < N; i++) {
			const double l = A[(size_t) i * N + j] / ajj;
			A[(size_t) i * N + j] = l;
			for (int k = j + 1; k < N; k++) A[(size_t) i * N + k] -= l * A[(size_t) j * N + k];
		}
	}
	for (int i = 0; i < N; i++) { double v = B[perm[i]]; for (int k = 0; k < i; k++) v -= A[(size_t) i * N + k] * coeffs[k]; coeffs[i] = v; }
	for (int i = N - 1; i >= 0; i--) { double v = coeffs[i]; for (int k = i + 1; k < N; k++) v -= A[(size_t) i * N + k] * coeffs[k]; coeffs[i] = v / A[(size_t) i * N + i]; }
	return 0;
}

int sgpu_overlap_solve(int normalize, int nb_frames, int ref_index, const size_t *Nij, double *Mij, const double *Sij,
		double *offset, double *mul, double *scale) {
	if (!Nij || !Mij || !Sij || !offset || !mul || !scale || nb_frames < 2 || ref_index < 0 || ref_index >= nb_frames) return SGPU_GENERIC_ERROR;
	const int N = nb_frames - 1;
	std::vector<int> index; std::vector<double> coeffs(N, 0.0);
	for (int i = 0; i < nb_frames; i++) if (i != ref_index) index.push_back(i);
	if (normalize == SGPU_MULTIPLICATIVE_SCALING || normalize == SGPU_ADDITIVE_SCALING) {	// :875-888
		if (solve_overlap_coeffs(nb_frames, index.data(), ref_index, Nij, Sij, false, coeffs.data())) return SGPU_GENERIC_ERROR;
		for (int i = 0; i < N; i++) scale[index[i]] = coeffs[i];
		for (int ii = 0; ii < nb_frames; ii++) for (int jj = 0; jj < nb_frames; jj++) Mij[(size_t) ii * nb_frames + jj] *= scale[ii];
	}
	if (normalize == SGPU_ADDITIVE || normalize == SGPU_ADDITIVE_SCALING) {	// :890-897
		if (solve_overlap_coeffs(nb_frames, index.data(), ref_index, Nij, Mij, true, coeffs.data())) return SGPU_GENERIC_ERROR;
		for (int i = 0; i < N; i++) offset[index[i]] = -coeffs[i];
	}
	if (normalize == SGPU_MULTIPLICATIVE) {	// :899-906
		if (solve_overlap_coeffs(nb_frames, index.data(), ref_index, Nij, Mij, false, coeffs.data())) return SGPU_GENERIC_ERROR;
		for (int i = 0; i < N; i++) mul[index[i]] = coeffs[i];
	}
	return SGPU_OK;
}

/* compute_wfwhm_weights (median_and_mean.c:1137-1183): weighted_fwhm of the selected frames (regdata's float) */
int sgpu_wfwhm_weights(int nb_layers, int N, const float *weighted_fwhm, double *weights) {
	if (!weighted_fwhm || !weights || N < 1) return SGPU_GENERIC_ERROR;
	double fwhmmin = DBL_MAX, fwhmmax = -DBL_MAX;
	for (int i = 0; i < N; i++) {
		if (weighted_fwhm[i] < fwhmmin && weighted_fwhm[i] > 0) fwhmmin = weighted_fwhm[i];
		if (weighted_fwhm[i] > fwhmmax) fwhmmax = weighted_fwhm[i];
	}
	const double invdenom = 1. / (1. / (fwhmmin * fwhmmin) - 1. / (fwhmmax * fwhmmax));
	const double invfwhmax2 = 1. / (fwhmmax * fwhmmax);
	for (int l = 0; l < nb_layers; l++) {
		double norm = 0.0;
		double *pw = weights + (size_t) l * N;
		for (int i = 0; i < N; i++) {
			if (weighted_fwhm[i] > 0) {
				pw[i] = (1. / (weighted_fwhm[i] * weighted_fwhm[i]) - invfwhmax2) * invdenom;	// the square is a float product, as in the reference
				norm += pw[i];
			} else pw[i] = 0.;
		}
		norm /= (double) N;
		if (!norm) return SGPU_GENERIC_ERROR;	// NaN is "true" here too: the reference goes on with NaN weights
		for (int i = 0; i < N; i++) pw[i] /= norm;
	}
	return SGPU_OK;
}

/* compute_nbstars_weights (median_and_mean.c:1185-1230): number_of_stars of the selected frames */
int sgpu_nbstars_weights(int nb_layers, int N, const int *number_of_stars, double *weights) {
	if (!number_of_stars || !weights || N < 1) return SGPU_GENERIC_ERROR;
	int starmin = INT_MAX, starmax = 0;
	for (int i = 0; i < N; i++) {
		if (number_of_stars[i] < starmin) starmin = number_of_stars[i];
		if (number_of_stars[i] > starmax) starmax = number_of_stars[i];
	}
	const double invdenom = starmax == starmin ? 1.0 : 1. / (double) (starmax - starmin);
	for (int l = 0; l < nb_layers; l++) {
		double norm = 0.0;
		double *pw = weights + (size_t) l * N;
		for (int i = 0; i < N; i++) {
			if (starmax == starmin) pw[i] = 1.;
			else pw[i] = (double) (number_of_stars[i] - starmin) * (double) (number_of_stars[i] - starmin) * invdenom * invdenom;
			norm += pw[i];
		}
		norm /= (double) N;
		for (int i = 0; i < N; i++) pw[i] /= norm;
	}
	return SGPU_OK;
}

/* NBSTACK_WEIGHT (median_and_mean.c:150-159): the frame's STACKCNT keyword as its weight on every layer, not normalized */
int sgpu_nbstack_weights(int nb_layers, int N, const int *stack_count, double *weights) {
	if (!stack_count || !weights || N < 1) return SGPU_GENERIC_ERROR;
	for (int l = 0; l < nb_layers; l++)
		for (int i = 0; i < N; i++) weights[(size_t) l * N + i] = (double) stack_count[i];
	return SGPU_OK;
}

/* refine_blocks_candidate + stack_compute_parallel_blocks (median_and_mean.c:261-356) */
int sgpu_compute_parallel_blocks(sgpu_image_block *blocks, int max_blocks, long max_rows, const long naxes[3],
		int nb_threads, long *largest, int *nb_blocks) {
	if (!naxes || !largest || !nb_blocks || nb_threads < 1 || max_rows < 1) return SGPU_GENERIC_ERROR;
	const long H = naxes[1], C = naxes[2];
	auto ceil_mult = [](int x, int f) { const int r = x % f; return r ? x + f - r : x; };
	int cand = nb_threads;
	while ((max_rows * cand) / nb_threads < H * C) cand++;
	int factor = (C == 3) ? 3 : 1;
	if (nb_threads < 4) {
		if (factor != 1 && nb_threads % factor == 0) factor = nb_threads; else factor *= nb_threads;
		cand = ceil_mult(cand, factor);
	} else {
		const int slack = nb_threads < 8 ? 1 : 3;
		cand = ceil_mult(cand, factor);
		for (;;) { const int rem = cand % nb_threads; if (rem == 0 || rem >= nb_threads - slack) break; cand += factor; }
	}
	*nb_blocks = cand;
	const long hb = H * C / cand;
	long remainder = H % (cand / C);
	*largest = 0;
	long ch = 0, row = 0; int j = 0;
	do {
		if (j >= cand) return SGPU_GENERIC_ERROR;
		long end = row + hb - 1;
		if (remainder > 0) { end++; remainder--; }
		const long start = row, channel = ch;
		if (end >= H - 1 || (H - end < hb / 10)) {
			end = H - 1; row = 0; ch++;
			remainder = H - (cand / C * hb);
		} else row = end + 1;
		const long h = end - start + 1;
		if (blocks && j < max_blocks) { blocks[j].channel = channel; blocks[j].start_row = start; blocks[j].end_row = end; blocks[j].height = h; }
		if (*largest < h) *largest = h;
		j++;
	} while (ch < C);
	return SGPU_OK;
}

}  // extern "C"
