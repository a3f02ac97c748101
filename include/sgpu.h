/*
 * sgpu.h — C ABI of the B200 (sm_100a) rejection-stacking module.
 *
 * This is the drop-in boundary for ONE hot path of Siril (lock042/siril @ a5869d93): the
 * per-pixel mean-with-rejection / median combine with normalization and weighting, i.e. what
 * `args->method(args)` (stack_mean_with_rejection / stack_median, src/stacking/stacking.c:104,
 * src/stacking/median_and_mean.c:1103-1109) and do_normalization (src/stacking/normalization.c:46)
 * compute.  Plain C: pointers and sizes only, no glib, no torch, no C++ types.  Siril's glue
 * (integration/siril/src/gpu/gpu_stack.c) and the Python harness both bind
 * exactly these entry points.
 *
 * Conventions
 *  - every function returns an ST_* code of the reference (src/stacking/stacking.h:18-24):
 *      SGPU_OK 0, SGPU_GENERIC_ERROR -1, SGPU_SEQUENCE_ERROR -2, SGPU_CANCEL -9, SGPU_ALLOC_ERROR -10
 *    and never throws; sgpu_last_error() gives the text.
 *  - enum values are the reference's (src/core/settings.h:34-53, src/stacking/stacking.h:46-52).
 *  - images are in MEMORY order as Siril holds them (fits.fdata / fits.data: FITS, bottom-up
 *    rows), planes [layer][row][col].  A "block" is what data->pix[frame] is in the reference
 *    (src/stacking/median_and_mean.c:382-545): for every frame, the rows [row0, row0+height) of
 *    every layer with the frame's integer y-shift already applied at load time and zero fill
 *    outside the frame; layout [frame][layer][row][col] ("frame-major"), resident in HBM.
 *  - coefficient / weight arrays are [layer * nb_frames + frame] doubles, the layout of
 *    coeff.pscale[layer][frame] / weights (normalization.c:81-96, median_and_mean.c:1115-1120).
 *  - there is no CPU fallback: without a CUDA device every compute entry point fails with
 *    SGPU_GENERIC_ERROR.
 */
#ifndef SGPU_H
#define SGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SGPU_ABI_VERSION 5

enum {	/* stacking.h:18-24 */
	SGPU_OK = 0,
	SGPU_GENERIC_ERROR = -1,
	SGPU_SEQUENCE_ERROR = -2,
	SGPU_CANCEL = -9,
	SGPU_ALLOC_ERROR = -10
};

enum { SGPU_U16 = 0, SGPU_F32 = 1 };	/* data_type: DATA_USHORT, DATA_FLOAT */

enum {	/* rejection, settings.h:43-52 */
	SGPU_NO_REJEC = 0, SGPU_PERCENTILE, SGPU_SIGMA, SGPU_MAD, SGPU_SIGMEDIAN,
	SGPU_WINSORIZED, SGPU_LINEARFIT, SGPU_GESDT
};

enum {	/* normalization, settings.h:34-40 */
	SGPU_NO_NORM = 0, SGPU_ADDITIVE, SGPU_MULTIPLICATIVE, SGPU_ADDITIVE_SCALING,
	SGPU_MULTIPLICATIVE_SCALING
};

enum {	/* weightingType, stacking.h:46-52 */
	SGPU_NO_WEIGHT = 0, SGPU_NBSTARS_WEIGHT, SGPU_WFWHM_WEIGHT, SGPU_NOISE_WEIGHT, SGPU_NBSTACK_WEIGHT
};

typedef struct sgpu_ctx sgpu_ctx;
typedef struct sgpu_block sgpu_block;

/* ---- context -------------------------------------------------------------------------------- */
int sgpu_abi_version(void);
/* one context per GPU / per process rank; `device` is the CUDA ordinal */
int sgpu_ctx_create(int device, sgpu_ctx **ctx);
void sgpu_ctx_destroy(sgpu_ctx *ctx);
/* Run all work of this context on a caller-owned cudaStream_t.  NULL means CUDA's legacy default
 * stream (the runtime API's convention), not "none": a context that is never given a stream uses a
 * private non-blocking stream of its own. */
int sgpu_ctx_set_stream(sgpu_ctx *ctx, void *cuda_stream);
int sgpu_ctx_synchronize(sgpu_ctx *ctx);
const char *sgpu_last_error(const sgpu_ctx *ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
uint64_t sgpu_launch_count(const sgpu_ctx *ctx);
/* number of pixels the last sgpu_stack_block / sgpu_stack_host call of this context handed from the
 * selection / lane-sort / sorted engine to the exact engine (diagnostic; synchronizes the context stream) */
uint64_t sgpu_handover_count(sgpu_ctx *ctx);
/* engine the last stack call of this context ran on: 1 exact, 2 sorted, 5 lane-sort, 6 selection */
int sgpu_last_engine(const sgpu_ctx *ctx);
/* sgpu_image_stats: medians of this context that fell outside the shared-memory histogram window and
 * were recomputed on the full 65536-bin histogram (diagnostic; results are identical either way) */
uint64_t sgpu_stats_window_misses(const sgpu_ctx *ctx);
/* cooperative cancel: polled between row-block launches like processing_should_continue()
 * (median_and_mean.c:1563); return non-zero from the callback to get SGPU_CANCEL */
int sgpu_ctx_set_cancel_callback(sgpu_ctx *ctx, int (*should_cancel)(void *), void *user);

/* device memory as the runtime sees it (the glue sizes its row blocks from it, as the reference sizes
 * its own from get_max_memory_in_MB(), median_and_mean.c:1234) */
int sgpu_mem_info(sgpu_ctx *ctx, size_t *free_bytes, size_t *total_bytes);
/* zeroed device memory for the outputs of sgpu_stack_block, and an asynchronous device -> host copy on
 * the context stream (sgpu_ctx_synchronize completes it) */
int sgpu_device_alloc(sgpu_ctx *ctx, size_t bytes, void **dev_ptr);
void sgpu_device_free(sgpu_ctx *ctx, void *dev_ptr);
int sgpu_download(sgpu_ctx *ctx, void *host_dst, const void *dev_src, size_t bytes);

/* Outputs on a peer GPU ("output rows plus rejection maps gathered over NVLink").  One process per GPU:
 * the root exports the image it allocated with sgpu_device_alloc, every other rank opens it and hands sgpu_stack_block
 * pointers INTO it (its own rows; out_plane_stride = the image's plane): the stacking kernels' stores then travel over
 * NVLink / NVSwitch while the stack is still being computed and no gather follows — the assembly of the result image that
 * the reference does by writing every block into one fits (median_and_mean.c:1717-1734).  The writer's
 * sgpu_ctx_synchronize (then a barrier between the processes) makes the rows visible to the owner.
 * CUDA IPC underneath: same node, peer access between the two GPUs. */
#define SGPU_IPC_HANDLE_BYTES 64
int sgpu_ipc_export(sgpu_ctx *ctx, void *dev_ptr, unsigned char handle[SGPU_IPC_HANDLE_BYTES]);
int sgpu_ipc_open(sgpu_ctx *ctx, const unsigned char handle[SGPU_IPC_HANDLE_BYTES], void **peer_ptr);
int sgpu_ipc_close(sgpu_ctx *ctx, void *peer_ptr);

/* pinned host memory for the upload / download legs */
int sgpu_host_alloc(sgpu_ctx *ctx, size_t bytes, void **ptr);
void sgpu_host_free(sgpu_ctx *ctx, void *ptr);

/* ---- frame blocks (replaces stack_read_block_data's data->pix[], median_and_mean.c:382) ------ */
typedef struct {
	int nb_frames;	/* args->nb_images_to_stack */
	int nb_layers;	/* naxes[2] */
	int width;	/* naxes[0] */
	int height;	/* rows held by this block (a row shard; == naxes[1] on one GPU) */
	int dtype;	/* SGPU_U16 | SGPU_F32 */
} sgpu_block_desc;

int sgpu_block_create(sgpu_ctx *ctx, const sgpu_block_desc *desc, sgpu_block **blk);
/* wrap caller-owned device memory of exactly [nb_frames][nb_layers][height][width] elements */
int sgpu_block_wrap(sgpu_ctx *ctx, const sgpu_block_desc *desc, void *device_ptr, sgpu_block **blk);
void sgpu_block_destroy(sgpu_block *blk);
void *sgpu_block_device_ptr(sgpu_block *blk);
/* Copy rows of one full host frame ([nb_layers][frame_height][width], memory order) into the
 * block: block row r (r in [0,height)) is image row row0 + r and receives frame row
 * row0 + r - shifty, or zeros where that is outside [0, frame_height) — the read-time y shift of
 * median_and_mean.c:407-448 expressed in memory row order.  Asynchronous on the context stream;
 * `host_frame` should be pinned for the copy to overlap. */
int sgpu_block_upload_frame(sgpu_block *blk, int frame, const void *host_frame, int frame_height,
		int row0, int shifty);

/* Maximized framing / sequences of variable-size frames.  With args->maximize_framing the result is
 * the canvas that holds every registered frame (stack_open_all_files, median_and_mean.c:160-190): naxes0 x naxes1 and the
 * origin args->offset from the frames' sizes and homography translations (regparam[].H.h02 / .h12). */
int sgpu_maximized_framing(int nb_frames, const int *rx, const int *ry, const double *h02, const double *h12,
		int upscale_at_stacking, long *naxes0, long *naxes1, int offset[2]);
/* sgpu_block_upload_frame for a frame of its own size [nb_layers][frame_height][frame_width] into a block of the canvas'
 * width: what seq_opened_read_region + rearrange_block_data leave in data->pix[frame] (:363-380, :394-480) — the frame's
 * rows at the left of the block's rows, zero padded to the block's width, its TOP row on the canvas' top row before the y
 * shift.  image_height is the canvas height (naxes1), row0 the canvas (memory) row of block row 0. */
int sgpu_block_upload_frame_sized(sgpu_block *blk, int frame, const void *host_frame, int frame_width, int frame_height,
		int image_height, int row0, int shifty);

/* ---- feather masks and drizzle weights ---------------------------------------
 * Two optional float plane sets of a block, what data->mask[frame] and data->drizz[frame] are in the reference
 * (stack_read_block_data, median_and_mean.c:483-542): the feather mask of every frame, [nb_frames][height][width]
 * (one plane per frame, shared by the layers: the reference reads the same mask file for each channel block), and
 * the drizzle weights, [nb_frames][nb_layers][height][width].  With either present a mean stack takes the weighted
 * combine with n = drizzle * mask * weight per sample (mean_and_reject, :1043-1070, :1060-1066; gathered at the
 * sample's shifted column, :1687-1692), and samples with a zero drizzle weight are dropped with the null samples
 * before the rejection (rejection_float.c:116-126, median_and_mean.c:716-726).  Median stacks ignore them. */
enum { SGPU_PLANE_MASK = 0, SGPU_PLANE_DRIZZLE = 1 };
/* rows of one host plane set of one frame ([frame_height][width] mask, [nb_layers][frame_height][width] drizzle
 * weights) into the block, geometry as sgpu_block_upload_frame; the plane set is allocated (zeroed) on first use */
int sgpu_block_upload_weight_plane(sgpu_block *blk, int which, int frame, const float *host_planes,
		int frame_height, int row0, int shifty);
/* or: use caller-owned device memory of exactly those shapes (NULL: none) */
int sgpu_block_set_weight_planes(sgpu_block *blk, float *d_mask, float *d_drizzle);
/* the mask normalization of :513-521 on the block's mask planes, in place: a distance above feather_dist becomes
 * 1, one below it the smoothstep ramp of blending.c:34-50 (get_ramped_value, 1001-point table by truncation),
 * zero stays zero.  Upload the up-scaled distance masks (cvUpscaleBlendMask's output), then call this once. */
int sgpu_block_mask_ramp(sgpu_block *blk, int feather_dist);

/* ---- raw FITS data units ----------------------------------------------------
 * Replaces, for uncompressed primary-HDU images, the read leg of stack_read_block_data
 * (median_and_mean.c:382-545 -> seq_opened_read_region -> internal_read_partial_fits,
 * image_format_fits.c:927-1015): the host moves the file's bytes (pread -> pinned staging -> H2D, two
 * chunks in flight) and the device does what cfitsio + convert_data_ushort (:390-416) /
 * convert_floats (:648-672) do on the CPU: big-endian decode, +32768 for 16-bit data, BYTE
 * widening, the DATAMAX > 10 rescale by INV_USHRT_MAX_SINGLE, y shift and zero fill. */
typedef struct {
	int bitpix;		/* 8, 16, 32, -32, -64 as in the header */
	long naxes[3];		/* NAXIS1..3 (1 when absent) */
	double bzero, bscale;	/* 0 / 1 when absent */
	double datamax; int has_datamax;	/* DATAMAX (image_format_fits.c:995-1008) */
	long data_offset;	/* byte offset of the primary data unit */
	int device_readable;	/* 1: BYTE, SHORT/USHORT or FLOAT without physical scaling */
} sgpu_fits_info;
/* parse the primary header (80-byte cards, 2880-byte blocks, up to END); SGPU_SEQUENCE_ERROR when
 * the file is missing, not FITS or shorter than its header says */
int sgpu_fits_probe(const char *path, sgpu_fits_info *info);
/* data_unit: host memory holding the whole big-endian data unit [layer][frame_height][width] of one
 * frame (an mmap of the file, or pinned memory from sgpu_host_alloc: then the H2D copy reads it in
 * place).  bitpix 8 / 16 -> SGPU_U16 blocks, -32 -> SGPU_F32 blocks.  rescale_ushrt: 1 for float data whose
 * DATAMAX is above 10 (image_format_fits.c:1009-1011), 0 otherwise, -1 when the header has no DATAMAX:
 * the reference then decides per read from three or four samples of the region (:997-1003), and so
 * does this call, per layer, on the rows it uploads.  Geometry as sgpu_block_upload_frame.
 * Asynchronous on the context stream, except for waiting on its own staging buffers. */
int sgpu_block_upload_fits(sgpu_block *blk, int frame, const void *data_unit, int bitpix,
		int frame_height, int row0, int shifty, int rescale_ushrt);
/* the same, reading the rows this block needs from an open file descriptor with pread */
int sgpu_block_upload_fits_fd(sgpu_block *blk, int frame, int fd, long data_offset, int bitpix,
		int frame_height, int row0, int shifty, int rescale_ushrt);

/* ---- stacking (replaces the block loop of stack_mean_or_median, median_and_mean.c:1552-1760) -- */
typedef struct {
	int is_mean;			/* 1: stack_mean_with_rejection, 0: stack_median */
	int type_of_rejection;		/* args->type_of_rejection */
	float sig[2];			/* args->sig */
	const float *critical_value;	/* args->critical_value (GESDT), floor(nb_frames*sig[0]) entries (host) */
	int normalize;			/* args->normalize */
	const double *scale, *offset, *mul;	/* args->coeff.scale/offset/mul (host), NULL when NO_NORM */
	const double *weights;		/* args->weights (host) or NULL == NO_WEIGHT */
	const int *shiftx;		/* per-frame integer x shift (round_to_int(dx), :1621-1623) or NULL */
	int output_norm;		/* args->output_norm: skip the [0,1] clamp (rescale is sgpu_norm_to_0_1) */
	int use_32bit_output;		/* args->use_32bit_output: float result, else WORD result */
	int create_rejmaps;		/* args->create_rejmaps */
	int merge_lowhigh_rejmaps;	/* args->merge_lowhigh_rejmaps */
	int nb_critical;		/* entries critical_value holds; 0: not stated (trusted to be (int) floor((float) nb_frames * sig[0])) */
	int source_bitpix;		/* seq->bitpix of the frames as they were on disk; only 8 (BYTE_IMG) matters: with a 16-bit output and
					 * output_norm the result is scaled by 65535/255 (normalize_to16bit, median_and_mean.c:547-555,1729-1733); 0: unknown */
	size_t out_plane_stride;	/* sgpu_stack_block: elements between the layers of d_out / d_rej_low / d_rej_high; 0: height * width (dense).
					 * A row shard that writes its rows straight into a taller image (its own or a peer GPU's, see
					 * sgpu_ipc_open) passes the image's plane size and pointers to its first row. */
	int engine;			/* 0: automatic; 1: force the exact thread-per-pixel engine; 2: force the sorted engine; 5: force the lane-sort engine; 6: force the selection engine */
} sgpu_stack_args;

/* All outputs are DEVICE pointers: d_out [nb_layers][height][width] float or uint16_t,
 * d_rej_low/high same shape uint16_t (or NULL), d_irej 6 x uint64_t laid out irej[layer][low,high]
 * and ACCUMULATED into (zero it first).  Asynchronous on the context stream. */
int sgpu_stack_block(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *args,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej);

/* End-to-end host call: frames[f] points to a full host frame [nb_layers][height][width];
 * shifty may be NULL.  Uploads (H2D), stacks, applies output_norm, downloads (D2H) into the host
 * outputs and returns when they are complete.  irej may be NULL.  Large inputs go through in row
 * blocks like the reference's block loop (median_and_mean.c:1552): block k + 1 is uploaded on a
 * copy stream while block k is stacked and its rows are read back (pinned host frames overlap;
 * pageable ones still work).  output_norm needs the whole image's range and runs as one block.
 * sgpu_handover_count / sgpu_last_engine then describe the last row block. */
int sgpu_stack_host(sgpu_ctx *ctx, const sgpu_block_desc *desc, const void *const *frames,
		const int *shifty, const sgpu_stack_args *args,
		void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]);

/* sgpu_stack_host for a sequence of single-image FITS files on disk (what the `stack` command is given,
 * command.c:11729 -> stack_mean_or_median): the files are probed (same NAXISn / BITPIX, device
 * readable, else SGPU_SEQUENCE_ERROR as median_and_mean.c:1300-1310), and every row block is filled by
 * the raw reader (pread -> pinned staging -> H2D -> device decode) while the block before it is stacked.
 * out: host [NAXIS3][NAXIS2][NAXIS1] float or uint16_t, sized by the caller from sgpu_fits_probe. */
int sgpu_stack_fits(sgpu_ctx *ctx, int nb_frames, const char *const *paths, const int *shifty,
		const sgpu_stack_args *args, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]);

/* FITS sequence files (SEQ_FITSEQ: one frame per HDU, io/fits_sequence.c): the frames _find_hdus keeps (:39-150 — image HDUs
 * with NAXIS > 1 that are not ICC profiles, thumbnails, HDRLET / WCS / D2IM extensions; an HDU whose layer count or sample type
 * differs from the first image's is skipped, a different width or height fails the file), each with the offset of its data
 * unit: sgpu_block_upload_fits_fd reads them as it reads single-image files.  infos may be NULL to count. */
int sgpu_fitseq_probe(const char *path, sgpu_fits_info *infos, int max_frames, int *nb_frames);
/* sgpu_stack_fits over frames of one sequence file; frame_indices (NULL: 0 .. nb_frames - 1) index the probed frames */
int sgpu_stack_fitseq(sgpu_ctx *ctx, const char *path, int nb_frames, const int *frame_indices, const int *shifty,
		const sgpu_stack_args *args, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]);

/* SER files (SEQ_SER, io/ser.c): a 178-byte header, then frame_count frames of height x width samples, rows TOP-DOWN, 8-bit
 * or 16-bit in the byte order the header names — its "LittleEndian" field read the way Siril and the first SER programs read
 * it: 0 = little-endian data, 1 = big-endian (io/ser.h:71-80).  8-bit samples become WORDs unscaled and 16-bit ones are
 * byte-swapped as needed (ser_manage_endianess_and_depth, ser.c:521-539); the reads are read_area_from_image's (:1054-1090).
 * device_readable: one plane per frame (mono, or a colour filter array taken as mono — no demosaicing, no RGB / BGR). */
typedef struct {
	int color_id, little_endian, width, height, bit_depth, frame_count;	/* header fields as stored */
	int byte_depth, planes;		/* 1 | 2 bytes per sample; 1 | 3 planes (ser.c:345-352) */
	long data_offset;		/* SER_HEADER_LEN */
	int device_readable;
} sgpu_ser_info;
int sgpu_ser_probe(const char *path, sgpu_ser_info *info);
/* frame `ser_frame` of an open SER file into frame `frame` of a one-layer SGPU_U16 block: geometry as sgpu_block_upload_frame
 * (memory order is bottom-up: the file's rows are turned over on the device) */
int sgpu_block_upload_ser_fd(sgpu_block *blk, int frame, int fd, const sgpu_ser_info *info, int ser_frame, int row0, int shifty);
/* sgpu_stack_fits for frames of one SER file; frame_indices NULL: 0 .. nb_frames - 1; out is WORD or float [1][height][width] */
int sgpu_stack_ser(sgpu_ctx *ctx, const char *path, int nb_frames, const int *frame_indices, const int *shifty,
		const sgpu_stack_args *args, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]);

/* norm_to_0_1_range (median_and_mean.c:557-582) split for row shards: min/max over non-zero
 * samples (skip_first: the reference's scan starts at index 1 — set it on the shard that holds
 * element 0), then the rescale with the globally reduced min/max. */
int sgpu_minmax_nonzero(sgpu_ctx *ctx, const float *d_img, size_t n, int skip_first, float minmax[2]);
int sgpu_rescale_0_1(sgpu_ctx *ctx, float *d_img, size_t n, float mini, float maxi);

/* sgpu_minmax_nonzero for callers that cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Same inputs, outputs and results as sgpu_minmax_nonzero; the call allocates
 * nothing (no cudaMalloc / cudaFree, no host allocation): its per-chunk partials live in a file-scope
 * static device array of SGPU_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 * Limit: n <= SGPU_STATIC_MAX_PIXELS (one 8-byte partial per 65536 pixels, one slot for the result);
 * beyond it the call fails with SGPU_ALLOC_ERROR and a message in sgpu_last_error.  All calls (of all
 * contexts on a device) share the array: calls that overlap on the GPU race — serialize them. */
#define SGPU_STATIC_SCRATCH_BYTES 65536
#define SGPU_STATIC_MAX_PIXELS ((SGPU_STATIC_SCRATCH_BYTES / 8 - 1) * (size_t) 65536)	/* 536 805 376 */
#define SGPU_STATIC_MAX_DEVICES 64
int sgpu_minmax_nonzero_static(sgpu_ctx *ctx, const float *d_img, size_t n, int skip_first, float minmax[2]);

/* ---- sum / max / min stackers (replace sum_stacking_image_hook + _finalize_hook, sum.c:107-372,
 * and minmax_stacking_image_hook, stackminmax.c:103-184) ------------------------------------- */
#define SGPU_SIMPLE_SUM 0	/* stack_summing_generic (sum.c:418) */
#define SGPU_SIMPLE_MAX 1	/* stack_addmax (stackminmax.c:301) */
#define SGPU_SIMPLE_MIN 2	/* stack_addmin (stackminmax.c:305) */
/* One pass over the frames of a resident block.  shiftx / shifty: the integer registration shifts
 * (host arrays of nb_frames, or NULL); the block's rows must have been uploaded with the same
 * shifty (sgpu_block_upload_frame), row0 is the image row of block row 0 and frame_height the full
 * frame height: a pixel outside a shifted frame is skipped, as in the reference, not read as zero.
 * d_acc (device, [nb_layers][height][width]): SUM -> double (float samples) or uint64_t (WORD
 * samples) sums, and *max_out (host, may be NULL) their maximum over the block, never below 0;
 * MAX / MIN -> the result in the sample type.  SUM with max_out synchronizes the stream. */
int sgpu_simple_block(sgpu_ctx *ctx, sgpu_block *blk, int op, const int *shiftx, const int *shifty,
		int row0, int frame_height, void *d_acc, double *max_out);
/* sum.c:307-343 with the maximum over the WHOLE image (reduce it over row shards first): float
 * output = (float)(sum / max); WORD output = sum, scaled by 65535 / max when max > 65535.
 * Fails like the reference when max is 0 (sum.c:266, :275). */
int sgpu_sum_convert(sgpu_ctx *ctx, const void *d_acc, int acc_is_float, size_t n, double max,
		int output_32bits, void *d_out);

/* ---- normalization inputs (replaces compute_normalization, normalization.c:217) --------------- */
/* imstats fields in the reference's order (core/siril.h:407-413) */
typedef struct {
	double total, ngoodpix, mean, median, sigma, avgDev, mad, sqrtbwmv,
	       location, scale, min, max, normValue, bgnoise;
} sgpu_imstats;

/* statistics_internal_float / _ushort with STATS_NORM (lite: STATS_LITENORM) on one plane of
 * nx*ny samples; `plane` is a device pointer when on_device != 0, else host memory (uploaded).
 * lite == SGPU_STATS_NOISE_ONLY stops after STATS_BASIC | STATS_BGNOISE (min, max, ngoodpix, mean,
 * sigma, bgnoise): the noise estimate the reference logs for a stacked result (bgnoise_image_hook,
 * command.c:6901); median / mad / location / scale are left 0. */
#define SGPU_STATS_NOISE_ONLY 2
int sgpu_image_stats(sgpu_ctx *ctx, const void *plane, int on_device, int dtype, long nx, long ny,
		int lite, sgpu_imstats *out);

/* The same statistics of a frame held as ROW SHARDS, one process per GPU: `shard` is
 * rows [row0, row0 + ny) of a plane of full_ny rows.  The estimators are whole-frame quantities
 * (normalization.c:217-294 -> statistics_float.c:281-481), so every reduction over the plane — counts,
 * sums, minima / maxima, the integer histograms the medians are read from, the per-row noise values —
 * goes through `reduce` before it is used: an in-place all-reduce over all ranks of `count` elements
 * of host memory (what ncclAllReduce / MPI_Allreduce do; return 0 on success).  Every rank must make
 * the call for the same plane; every rank gets the whole frame's statistics — the counts and medians
 * bit-identical to the unsharded ones, the double sums to ~1e-15 (the summation order differs). */
enum { SGPU_RED_F32 = 0, SGPU_RED_F64 = 1, SGPU_RED_U32 = 2, SGPU_RED_U64 = 3 };
enum { SGPU_RED_SUM = 0, SGPU_RED_MIN = 1, SGPU_RED_MAX = 2 };
typedef int (*sgpu_reduce_fn)(void *user, void *host_buf, size_t count, int dtype, int op);
int sgpu_image_stats_sharded(sgpu_ctx *ctx, const void *shard, int on_device, int dtype, long nx, long ny,
		long row0, long full_ny, int lite, sgpu_reduce_fn reduce, void *user, sgpu_imstats *out);

/* the statistics cache across runs: one `M<layer>-<image> ...` line of a .seq file per cached imstats,
 * exactly as writeseqfile prints it (io/seqfile.c:838-853: %ld counts, %lg — six significant digits — for the rest) and
 * readseqfile parses it (:544-611; '*' in place of the layer marks CFA statistics and maps to layer 0).  A caller that
 * caches sgpu_image_stats results this way gets the reference's second-run behaviour, rounding included.
 * SGPU_SEQUENCE_ERROR for a line that is not a statistics record. */
int sgpu_imstats_seq_record(int layer, int image_index, const sgpu_imstats *st, char *line, size_t cap);
int sgpu_imstats_from_seq_record(const char *line, int *layer, int *image_index, sgpu_imstats *st);

/* host-side scalar work, [nb_layers * nb_frames] arrays */
/* _compute_estimators_for_image's choice of estimators (normalization.c:98-141) */
int sgpu_estimators_from_stats(int normalize, int lite, const sgpu_imstats *st,
		double *offset, double *mul, double *scale);
/* compute_factors_from_estimators (normalization.c:144-181): estimators in, factors out */
int sgpu_compute_factors(int normalize, int nb_layers, int nb_frames, int ref_index, int reglayer,
		int equalizeRGB, double *offset, double *mul, double *scale);
/* compute_noise_weights (median_and_mean.c:1111-1135) */
int sgpu_noise_weights(int nb_layers, int nb_frames, const double *scale, const double *bgnoise,
		double *weights);

/* ---- overlap normalization (compute_normalization_overlaps, normalization.c:666-938; used with -maximize framing) ----
 * Frames that only partly overlap are levelled pair by pair: for every pair (i, j) the statistics of the pixels both
 * frames have (non-zero in both) over their common rectangle, then one small linear system per layer. */
typedef struct { int x, y, w, h; } sgpu_rect;	/* rectangle (core/settings.h:24-26): y counts from the TOP row of the image */
/* compute_overlap (normalization.c:420-456): rx / ry the two frame sizes, dx / dy their translation_from_H
 * (registration.c:301-304); returns the number of pixels of the common area (0: none) and the two rectangles */
size_t sgpu_compute_overlap(int rxi, int ryi, int rxj, int ryj, double dxi, double dyi, double dxj, double dyj,
		sgpu_rect *areai, sgpu_rect *areaj);
/* overlap_stats_t's numbers (core/siril.h:416-424): Xij is frame i's statistic over its overlap with j */
typedef struct { size_t Nij; float medij, medji, madij, madji, locij, locji, scaij, scaji; } sgpu_overlap_stats;
/* one pair on one layer (_compute_estimators_for_images, normalization.c:549-593): plane_i / plane_j are whole planes
 * [ry][rx] in memory order (device pointers when on_device != 0, else host memory: only the rectangle's rows are
 * uploaded); WORD samples are converted to [0,1] floats first, as the reference does.  Nij stays 0 (and the statistics
 * with it) when 3 pixels or fewer are non-zero in both; lite != 0 stops after median / MAD (args->lite_norm). */
int sgpu_overlap_stats_pair(sgpu_ctx *ctx, const void *plane_i, int rxi, int ryi, const void *plane_j, int rxj, int ryj,
		int on_device, int dtype, const sgpu_rect *areai, const sgpu_rect *areaj, int lite, sgpu_overlap_stats *out);
/* the solves of compute_normalization_overlaps (:875-906) for ONE layer: Nij / Mij / Sij are nb_frames x nb_frames
 * row-major (pair counts, locations — median or IKSS location — and scales — MAD or IKSS scale; Mij is modified as the
 * reference modifies it); offset / mul / scale ([nb_frames], initialized 0 / 1 / 1 by init_coeffs) get the coefficients
 * of every frame but the reference one.  solve_overlap_coeffs (:296-355) builds the system; the LU solve behind it is
 * GSL's in the reference (gsl_linalg_LU_decomp / _solve), partial-pivoting Gaussian elimination here. */
int sgpu_overlap_solve(int normalize, int nb_frames, int ref_index, const size_t *Nij, double *Mij, const double *Sij,
		double *offset, double *mul, double *scale);

/* the other weightingTypes (stacking.h:46-52), from registration data / FITS keywords of the SELECTED frames, in stack order:
 * compute_wfwhm_weights (median_and_mean.c:1137-1183; regdata.weighted_fwhm, frames with 0 get weight 0; fails when the
 * weights sum to 0), compute_nbstars_weights (:1185-1230; regdata.number_of_stars), and NBSTACK_WEIGHT (:150-159: the
 * STACKCNT keyword of each frame, not normalized) */
int sgpu_wfwhm_weights(int nb_layers, int nb_frames, const float *weighted_fwhm, double *weights);
int sgpu_nbstars_weights(int nb_layers, int nb_frames, const int *number_of_stars, double *weights);
int sgpu_nbstack_weights(int nb_layers, int nb_frames, const int *stack_count, double *weights);

/* ---- row-block planner (stack_compute_parallel_blocks, median_and_mean.c:295-356) ------------- */
typedef struct { long channel, start_row, end_row, height; } sgpu_image_block;	/* stacking.h:167-169 */
int sgpu_compute_parallel_blocks(sgpu_image_block *blocks, int max_blocks, long max_number_of_rows,
		const long naxes[3], int nb_threads, long *largest_block_height, int *nb_blocks);

/* ---- synthetic frames for the benchmark (device-side generator) ---------------- */
int sgpu_synth_fill(sgpu_ctx *ctx, sgpu_block *blk, uint64_t seed, int row0, int full_height);

#ifdef __cplusplus
}
#endif
#endif /* SGPU_H */
