/*
 * gpu_stack.c — Siril-side glue of the B200 rejection-stacking module (src/gpu in a Siril tree).
 *
 * Two entry points, called from INSIDE the unchanged reference symbols when Siril is built with
 * -Dcuda=true, so that method pointers, logs and the FITS output keep working:
 *   stack_mean_with_rejection / stack_median (stacking/median_and_mean.c:1103-1109)
 *       -> gpu_stack_mean_or_median(args, is_mean)      replaces stack_mean_or_median (:1261-1826)
 *   do_normalization (stacking/normalization.c:46)
 *       -> gpu_do_normalization(args)                   replaces compute_normalization (:217-294)
 * Everything numeric happens behind the C ABI of libsgpu (gpu/sgpu.h); this file does what the
 * reference's orchestration does around it: geometry checks, result / rejection-map allocation,
 * critical values, the row-block loop with cancel polling and progress, the rejection log, and
 * the ownership rules (coeff and weights are freed here, the caller frees critical_value).
 * There is no CPU fallback: without a CUDA device the calls fail with ST_GENERIC_ERROR.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "core/siril.h"
#include "core/proto.h"
#include "core/processing.h"
#include "core/siril_log.h"
#include "io/sequence.h"
#include "io/image_format_fits.h"
#include "algos/statistics.h"
#include "registration/registration.h"
#include "stacking/stacking.h"
#include <gsl/gsl_cdf.h>
#include "gpu/sgpu.h"
#include "gpu/gpu_stack.h"

static sgpu_ctx *gpu_ctx;	/* one per process: Siril stacks on its single processing thread */

static int cancel_cb(void *user) { (void) user; return !processing_should_continue(); }

static sgpu_ctx *get_ctx(void) {
	if (!gpu_ctx && sgpu_ctx_create(0, &gpu_ctx) != SGPU_OK) {
		siril_log_message(_("GPU stacking: no usable CUDA device (there is no CPU fallback in this build)\n"));
		return NULL;
	}
	if (gpu_ctx) sgpu_ctx_set_cancel_callback(gpu_ctx, cancel_cb, NULL);
	return gpu_ctx;
}

/* integer registration shifts of a frame, as the reference derives them (median_and_mean.c:407-418,
 * :1621-1623): x applied per pixel by the kernel, y at load time */
static void frame_shifts(const struct stacking_args *args, int frame, int *sx, int *sy) {
	*sx = *sy = 0;
	if (args->reglayer < 0 || !args->seq->regparam) return;
	const regdata *layerparam = args->seq->regparam[args->reglayer];
	if (!layerparam) return;
	const double scale = args->upscale_at_stacking ? 2. : 1.;
	double dx, dy;
	translation_from_H(layerparam[args->image_indices[frame]].H, &dx, &dy);
	*sx = round_to_int((dx - args->offset[0]) * scale);
	*sy = round_to_int((dy - args->offset[1]) * scale);
}

/* Rows [row0, row0 + h) (memory order: FITS row 0 first) of every layer of one frame into `stage`,
 * laid out [layer][h][W]; what stack_read_block_data does for one block and frame
 * (median_and_mean.c:382-545) WITHOUT the y shift, which sgpu_block_upload_frame applies.
 * seq_opened_read_region counts rows from the top (image_format_fits.c:939-944). */
static int read_rows(struct stacking_args *args, int frame, const long naxes[3], data_type itype, int row0, int h, void *stage) {
	const size_t es = itype == DATA_FLOAT ? sizeof(float) : sizeof(WORD);
	rectangle area = { 0, (int) naxes[1] - row0 - h, (int) naxes[0], h };
	for (long layer = 0; layer < naxes[2]; layer++) {
		void *dst = (char *) stage + (size_t) layer * h * naxes[0] * es;
		if (seq_opened_read_region(args->seq, (int) layer, args->image_indices[frame], dst, &area, 0)) {
			siril_log_error(_("Error reading one of the image areas (%d: %d %d %d %d)\n"), args->image_indices[frame] + 1,
					area.x, area.y, area.w, area.h);
			return ST_SEQUENCE_ERROR;
		}
	}
	return ST_OK;
}

int gpu_stack_mean_or_median(struct stacking_args *args, gboolean is_mean) {
	int retval = ST_OK;
	const int nb_frames = args->nb_images_to_stack;
	fits fit = { 0 }, ref = { 0 };
	fits *fptr = &fit;
	sgpu_block *blk = NULL;
	void *stage = NULL, *d_out = NULL;
	int *shiftx = NULL, *shifty = NULL;
	long naxes[3] = { 0, 0, 0 };
	uint64_t irej[3][2] = { { 0 } };

	if (nb_frames < 2) {	/* median_and_mean.c:1279-1285 */
		siril_log_message(_("Select at least two frames for stacking. Aborting.\n"));
		return ST_GENERIC_ERROR;
	}
	if (nb_frames < 3 && is_mean && args->type_of_rejection == GESDT) {
		siril_log_message(_("The Generalized Extreme Studentized Deviate Test needs at least three frames for stacking. Aborting.\n"));
		return ST_GENERIC_ERROR;
	}
	if (args->feather_dist > 0 || args->drizzle || args->maximize_framing) {
		siril_log_message(_("GPU stacking: feathering, drizzle weights and maximized framing are not available yet\n"));
		return ST_GENERIC_ERROR;
	}
	sgpu_ctx *ctx = get_ctx();
	if (!ctx) return ST_GENERIC_ERROR;

	/* geometry from the first frame (:1290-1316); every frame is opened, closed at the end (:1784-1786) */
	for (int i = 0; i < nb_frames; i++) {
		if (seq_open_image(args->seq, args->image_indices[i])) { retval = ST_SEQUENCE_ERROR; goto free_and_close; }
	}
	naxes[0] = args->seq->rx; naxes[1] = args->seq->ry; naxes[2] = args->seq->nb_layers;
	if (naxes[0] == 0) {
		siril_log_error(_("Rejection stack error: uninitialized sequence\n"));
		retval = ST_SEQUENCE_ERROR;
		goto free_and_close;
	}
	const data_type itype = get_data_type(args->seq->bitpix);
	if (itype == DATA_UNSUPPORTED) { retval = ST_SEQUENCE_ERROR; goto free_and_close; }
	if (itype == DATA_FLOAT && !args->use_32bit_output) { retval = ST_GENERIC_ERROR; goto free_and_close; }
	const size_t es = itype == DATA_FLOAT ? sizeof(float) : sizeof(WORD);
	const size_t npx = (size_t) naxes[0] * naxes[1] * naxes[2];

	/* result image and rejection maps (:1318-1342) */
	if ((retval = new_fit_image(&fptr, naxes[0], naxes[1], naxes[2], args->use_32bit_output ? DATA_FLOAT : DATA_USHORT)))
		goto free_and_close;
	const gboolean maps = is_mean && args->create_rejmaps;
	if (maps) {
		if ((retval = new_fit_image(&args->rejmap_low, naxes[0], naxes[1], naxes[2], DATA_USHORT))) goto free_and_close;
		if (!args->merge_lowhigh_rejmaps &&
		    (retval = new_fit_image(&args->rejmap_high, naxes[0], naxes[1], naxes[2], DATA_USHORT))) goto free_and_close;
	}

	/* critical values of the generalized ESD test (:1479-1486): the caller frees them (command.c:11735) */
	if (is_mean && args->type_of_rejection == GESDT) {
		const int max_outliers = (int) floor(nb_frames * args->sig[0]);
		args->critical_value = malloc((max_outliers > 0 ? max_outliers : 1) * sizeof(float));
		if (!args->critical_value) { retval = ST_ALLOC_ERROR; goto free_and_close; }
		for (int j = 0, size = nb_frames; j < max_outliers; j++, size--) {
			float t_dist = gsl_cdf_tdist_Pinv(1 - args->sig[1] / (2 * size), size - 2);
			float numerator = (size - 1) * t_dist;
			float denominator = sqrtf(size) * sqrtf(size - 2 + (t_dist * t_dist));
			args->critical_value[j] = numerator / denominator;
		}
	}

	/* weights (:1519-1539).  Noise weighting: compute_noise_weights (:1111-1135) from the scale factors and
	 * the noise estimates do_normalization left in the sequence's statistics cache */
	if (is_mean && args->weighting_type != NO_WEIGHT) {
		if (args->weighting_type == WFWHM_WEIGHT || args->weighting_type == NBSTARS_WEIGHT) {
			/* compute_wfwhm_weights / compute_nbstars_weights (:1137-1230) from the registration data of the selected frames */
			const gboolean stars = args->weighting_type == NBSTARS_WEIGHT;
			if (args->reglayer < 0 || !args->seq->regparam || !args->seq->regparam[args->reglayer]) {
				siril_log_error(_("Sequence does not have registration info, cannot use weighing by %s, aborting\n"), stars ? _("number of stars") : "wFWHM");
				retval = ST_GENERIC_ERROR; goto free_and_close;
			}
			const regdata *rd = args->seq->regparam[args->reglayer];
			float *fw = malloc(nb_frames * sizeof(float)); int *ns = malloc(nb_frames * sizeof(int));
			args->weights = malloc((size_t) naxes[2] * nb_frames * sizeof(double));
			if (!fw || !ns || !args->weights) { free(fw); free(ns); retval = ST_ALLOC_ERROR; goto free_and_close; }
			for (int f = 0; f < nb_frames; f++) { fw[f] = rd[args->image_indices[f]].weighted_fwhm; ns[f] = rd[args->image_indices[f]].number_of_stars; }
			siril_log_message(stars ? _("Computing weights based on number of stars...\n") : _("Computing weights based on wFWHM...\n"));
			retval = stars ? sgpu_nbstars_weights((int) naxes[2], nb_frames, ns, args->weights)
				       : sgpu_wfwhm_weights((int) naxes[2], nb_frames, fw, args->weights);
			free(fw); free(ns);
			if (retval) { retval = ST_GENERIC_ERROR; goto free_and_close; }
		} else if (args->weighting_type != NOISE_WEIGHT) {
			/* NBSTACK_WEIGHT reads the STACKCNT keyword of every frame while the files are opened (:150-159): with the
			 * keyword reader of the host build, sgpu_nbstack_weights(nb_layers, nb_frames, stack_count, weights) */
			siril_log_message(_("GPU stacking: weighting by the number of stacked images needs the frames' STACKCNT keywords\n"));
			retval = ST_GENERIC_ERROR; goto free_and_close;
		} else {
		if (args->normalize == NO_NORM || !args->coeff.scale || !args->seq->stats) {
			siril_log_message(_("Weighting by noise requires normalization\n"));	/* command.c:11700-11703 */
			retval = ST_GENERIC_ERROR; goto free_and_close;
		}
		const size_t cn = (size_t) naxes[2] * nb_frames;
		double *bg = malloc(cn * sizeof(double));
		args->weights = malloc(cn * sizeof(double));
		if (!bg || !args->weights) { free(bg); retval = ST_ALLOC_ERROR; goto free_and_close; }
		for (long layer = 0; layer < naxes[2]; layer++)
			for (int f = 0; f < nb_frames; f++) {
				const imstats *st = args->seq->stats[layer] ? args->seq->stats[layer][args->image_indices[f]] : NULL;
				bg[layer * nb_frames + f] = st ? st->bgnoise : 0.0;
			}
		retval = sgpu_noise_weights((int) naxes[2], nb_frames, args->coeff.scale, bg, args->weights);
		free(bg);
		if (retval) goto free_and_close;
		}
	}

	shiftx = calloc(nb_frames, sizeof(int)); shifty = calloc(nb_frames, sizeof(int));
	if (!shiftx || !shifty) { retval = ST_ALLOC_ERROR; goto free_and_close; }
	gboolean any_sx = FALSE;
	for (int f = 0; f < nb_frames; f++) { frame_shifts(args, f, &shiftx[f], &shifty[f]); any_sx |= shiftx[f] != 0; }

	sgpu_stack_args sa;
	memset(&sa, 0, sizeof sa);
	sa.is_mean = is_mean; sa.type_of_rejection = args->type_of_rejection;
	sa.sig[0] = args->sig[0]; sa.sig[1] = args->sig[1];
	sa.critical_value = args->critical_value;
	sa.nb_critical = (is_mean && args->type_of_rejection == GESDT) ? (int) floor(nb_frames * args->sig[0]) : 0;
	sa.normalize = args->normalize;
	sa.scale = args->coeff.scale; sa.offset = args->coeff.offset; sa.mul = args->coeff.mul;
	sa.weights = is_mean ? args->weights : NULL;	/* "Weighting is allowed only with mean stacking" (command.c:11514) */
	sa.shiftx = any_sx ? shiftx : NULL;
	sa.output_norm = args->output_norm; sa.use_32bit_output = args->use_32bit_output;
	sa.create_rejmaps = maps; sa.merge_lowhigh_rejmaps = args->merge_lowhigh_rejmaps;
	sa.source_bitpix = args->seq->bitpix;	/* normalize_to16bit (:547-555) */

	/* Row blocks (the reference: stack_get_max_number_of_rows :1234 from the RAM budget; here from the
	 * HBM the device has left): every frame's rows of the block are read into a pinned staging buffer,
	 * uploaded (y shift and zero fill applied by the upload), stacked, and the block's rows of the
	 * result and of the maps come back.  The whole sequence usually fits: one block. */
	size_t free_b = 0, total_b = 0;
	if (sgpu_mem_info(ctx, &free_b, &total_b)) { retval = ST_GENERIC_ERROR; goto free_and_close; }
	const size_t out_b = npx * (args->use_32bit_output ? 4 : 2) + (maps ? npx * 4 : 0) + ((size_t) 64 << 20);
	const size_t row_b = (size_t) naxes[0] * naxes[2] * es * (size_t) nb_frames;
	size_t budget = free_b > out_b ? (size_t) ((free_b - out_b) * 0.85) : 0;
	long rows = budget / row_b;
	if (rows < 1) { siril_log_error(_("GPU stacking: not enough device memory for one image row of the sequence\n")); retval = ST_ALLOC_ERROR; goto free_and_close; }
	{	/* SIRIL_GPU_MAX_ROWS: caps the block height (the tests use it to force several blocks) */
		const char *cap = getenv("SIRIL_GPU_MAX_ROWS");
		if (cap && atol(cap) > 0 && atol(cap) < rows) rows = atol(cap);
	}
	if (rows > naxes[1]) rows = naxes[1];
	if (rows < naxes[1]) rows = rows >= 4 ? rows & ~3L : rows;	/* keeps block planes 16-byte multiples (TMA tile load) */
	const long nblocks = (naxes[1] + rows - 1) / rows;
	if (sgpu_host_alloc(ctx, (size_t) rows * naxes[0] * naxes[2] * es, &stage)) { retval = ST_ALLOC_ERROR; goto free_and_close; }
	/* device outputs of one block: [48-byte counters, accumulated over the blocks | result | low map | high map], zeroed */
	if (sgpu_device_alloc(ctx, 64 + (size_t) rows * naxes[0] * naxes[2] * (args->use_32bit_output ? 4 : 2) + 16 + 2 * (size_t) rows * naxes[0] * naxes[2] * 2, &d_out)) {
		retval = ST_ALLOC_ERROR; goto free_and_close;
	}

	for (long b = 0; b < nblocks && !retval; b++) {
		if (!processing_should_continue()) { retval = ST_CANCEL; break; }	/* :1563 */
		const int row0 = (int) (b * rows), h = (int) (row0 + rows <= naxes[1] ? rows : naxes[1] - row0);
		sgpu_block_desc d = { nb_frames, (int) naxes[2], (int) naxes[0], h, itype == DATA_FLOAT ? SGPU_F32 : SGPU_U16 };
		if ((retval = sgpu_block_create(ctx, &d, &blk))) break;
		for (int f = 0; f < nb_frames && !retval; f++) {
			/* block row r is image row row0 + r and receives frame row row0 + r - shifty: read the frame rows
			 * that land inside the block, let the upload zero the rest */
			long lo = (long) row0 - shifty[f], hi = lo + h;		/* frame rows wanted */
			long flo = lo < 0 ? 0 : lo, fhi = hi > naxes[1] ? naxes[1] : hi;
			if (fhi <= flo) { flo = 0; fhi = 1 <= naxes[1] ? 1 : 0; }	/* nothing lands: any row will do, it is shifted out */
			const int fh = (int) (fhi - flo);
			if ((retval = read_rows(args, f, naxes, itype, (int) flo, fh, stage))) break;
			/* the staged piece is a "frame" of fh rows whose row 0 belongs at block row flo + shifty - row0 */
			retval = sgpu_block_upload_frame(blk, f, stage, fh, 0, (int) (flo + shifty[f]) - row0);
			if (!retval) retval = sgpu_ctx_synchronize(ctx);	/* one staging buffer: double it for overlap */
		}
		if (retval) break;
		const size_t n = (size_t) h * naxes[0] * naxes[2];
		uint64_t *irej_d = (uint64_t *) d_out;
		char *o = (char *) d_out + 64;
		uint16_t *lo_d = (uint16_t *) (o + ((n * (args->use_32bit_output ? 4 : 2) + 15) & ~(size_t) 15)), *hi_d = lo_d + n;
		if ((retval = sgpu_stack_block(ctx, blk, &sa, o, maps ? lo_d : NULL, (maps && !args->merge_lowhigh_rejmaps) ? hi_d : NULL, irej_d))) break;
		/* rows come back into the result planes: block layout [layer][h][W] -> image [layer][H][W] */
		for (long layer = 0; layer < naxes[2] && !retval; layer++) {
			const size_t src = (size_t) layer * h * naxes[0], dst = ((size_t) layer * naxes[1] + row0) * naxes[0], cnt = (size_t) h * naxes[0];
			if (args->use_32bit_output) retval = sgpu_download(ctx, fit.fdata + dst, (float *) o + src, cnt * 4);
			else retval = sgpu_download(ctx, fit.data + dst, (WORD *) o + src, cnt * 2);
			if (!retval && maps) retval = sgpu_download(ctx, args->rejmap_low->data + dst, lo_d + src, cnt * 2);
			if (!retval && maps && !args->merge_lowhigh_rejmaps) retval = sgpu_download(ctx, args->rejmap_high->data + dst, hi_d + src, cnt * 2);
		}
		if (!retval) retval = sgpu_ctx_synchronize(ctx);
		sgpu_block_destroy(blk); blk = NULL;
		gui_iface.set_progress((double) (b + 1) / (double) nblocks, NULL);	/* :1603-1610 */
	}
	if (retval) goto free_and_close;
	if ((retval = sgpu_download(ctx, &irej[0][0], d_out, sizeof irej)) || (retval = sgpu_ctx_synchronize(ctx))) goto free_and_close;

	gui_iface.set_progress(1.0, _("Finalizing stacking..."));
	if (is_mean) {	/* :1766-1772 */
		double nb_tot = (double) naxes[0] * (double) naxes[1] * (double) nb_frames;
		for (long channel = 0; channel < naxes[2]; channel++) {
			siril_log_message(_("Pixel rejection in channel #%d: %.3lf%% - %.3lf%%\n"),
					channel, (double) irej[channel][0] / nb_tot * 100.0,
					(double) irej[channel][1] / nb_tot * 100.0);
		}
	}
	if (args->use_32bit_output && args->output_norm) {	/* norm_to_0_1_range (:557-582), the scan from index 1 included */
		float mini = FLT_MAX, maxi = -FLT_MAX;
		for (size_t i = 1; i < npx; i++) {
			const float v = fit.fdata[i];
			if (v != 0.f) { if (v < mini) mini = v; if (v > maxi) maxi = v; }
		}
		const float range = maxi - mini;
		for (size_t i = 0; i < npx; i++) if (fit.fdata[i] != 0.f) fit.fdata[i] = (fit.fdata[i] - mini) / range;
	}
	memcpy(&args->result, &fit, sizeof(fits));

free_and_close:
	for (int i = 0; i < nb_frames; ++i) seq_close_image(args->seq, args->image_indices[i]);
	if (blk) sgpu_block_destroy(blk);
	if (stage) sgpu_host_free(ctx, stage);
	if (d_out) sgpu_device_free(ctx, d_out);
	free(shiftx); free(shifty);
	(void) ref;
	if (args->normalize) {	/* :1799-1803: the callee frees the coefficients */
		free(args->coeff.offset); free(args->coeff.scale); free(args->coeff.mul);
		args->coeff.offset = args->coeff.scale = args->coeff.mul = NULL;
	}
	if (args->weights) { free(args->weights); args->weights = NULL; }	/* :1805 */
	if (retval) {	/* :1806-1826: nothing of the result survives an error */
		if (fit.data) free(fit.data);
		if (fit.fdata) free(fit.fdata);
		if (args->rejmap_low) { clearfits(args->rejmap_low); free(args->rejmap_low); args->rejmap_low = NULL; }
		if (args->rejmap_high) { clearfits(args->rejmap_high); free(args->rejmap_high); args->rejmap_high = NULL; }
		if (retval == ST_CANCEL) siril_log_message(_("Stacking cancelled\n"));
		else siril_log_message(_("GPU stacking failed: %s\n"), ctx ? sgpu_last_error(ctx) : "");
	}
	return retval;
}

/* do_normalization's compute_normalization (normalization.c:217-294): statistics of every frame and
 * layer — taken from the sequence's cache when there (statistics.c:506-528), else computed on the
 * GPU and STORED there, so that a second run and the .seq file behave as with the CPU path — then
 * _compute_estimators_for_image (:98-141) and compute_factors_from_estimators (:144-181). */
int gpu_do_normalization(struct stacking_args *args) {
	if (args->normalize == NO_NORM) return ST_OK;
	const int nb_frames = args->nb_images_to_stack, nb_layers = args->seq->nb_layers;
	sgpu_ctx *ctx = get_ctx();
	if (!ctx) return args->retval = ST_GENERIC_ERROR;
	int retval = ST_OK, ref_index = -1;
	const size_t cn = (size_t) nb_layers * nb_frames;
	/* init_coeffs (normalization.c:81-96): one allocation per array, [layer][frame] */
	args->coeff.offset = malloc(cn * sizeof(double));
	args->coeff.mul = malloc(cn * sizeof(double));
	args->coeff.scale = malloc(cn * sizeof(double));
	if (!args->coeff.offset || !args->coeff.mul || !args->coeff.scale) { retval = ST_ALLOC_ERROR; goto done; }
	for (int i = 0; i < nb_frames; i++) if (args->image_indices[i] == args->ref_image) ref_index = i;
	if (ref_index < 0) { siril_log_error(_("The reference image is not in the selected set of images.\n")); retval = ST_GENERIC_ERROR; goto done; }

	for (int i = 0; i < nb_frames && !retval; i++) {
		if (!processing_should_continue()) { retval = ST_CANCEL; break; }
		const int idx = args->image_indices[i];
		fits fit = { 0 };
		gboolean loaded = FALSE;
		for (int layer = 0; layer < nb_layers && !retval; layer++) {
			imstats *st = (args->seq->stats && args->seq->stats[layer] && !args->force_norm) ? args->seq->stats[layer][idx] : NULL;
			/* a cached record serves only if it holds what this normalization needs (statistics.c:506-528) */
			if (st && (st->median < 0. || st->mad < 0. || st->bgnoise < 0. || (!args->lite_norm && (st->location < 0. || st->scale < 0.)))) st = NULL;
			sgpu_imstats g;
			if (!st) {
				if (!loaded) {
					if (seq_read_frame(args->seq, idx, &fit, FALSE, -1)) { retval = ST_SEQUENCE_ERROR; break; }
					loaded = TRUE;
				}
				const void *plane = fit.type == DATA_FLOAT ? (const void *) fit.fpdata[layer] : (const void *) fit.pdata[layer];
				retval = sgpu_image_stats(ctx, plane, 0, fit.type == DATA_FLOAT ? SGPU_F32 : SGPU_U16, fit.rx, fit.ry, args->lite_norm ? 1 : 0, &g);
				if (retval) break;
				st = calloc(1, sizeof(imstats));
				if (!st) { retval = ST_ALLOC_ERROR; break; }
				st->total = (long) g.total; st->ngoodpix = (long) g.ngoodpix;
				st->mean = g.mean; st->median = g.median; st->sigma = g.sigma; st->avgDev = -1.0; st->mad = g.mad;
				st->sqrtbwmv = -1.0; st->location = args->lite_norm ? -1.0 : g.location; st->scale = args->lite_norm ? -1.0 : g.scale;
				st->min = g.min; st->max = g.max; st->normValue = g.normValue; st->bgnoise = g.bgnoise;
				add_stats_to_seq(args->seq, idx, layer, st);	/* the sequence owns it from here */
			}
			memset(&g, 0, sizeof g);
			g.median = st->median; g.mad = st->mad; g.location = st->location; g.scale = st->scale;
			const size_t k = (size_t) layer * nb_frames + i;
			sgpu_estimators_from_stats(args->normalize, args->lite_norm ? 1 : 0, &g, &args->coeff.offset[k], &args->coeff.mul[k], &args->coeff.scale[k]);
		}
		if (loaded) clearfits(&fit);
	}
	if (!retval)
		retval = sgpu_compute_factors(args->normalize, nb_layers, nb_frames, ref_index, args->reglayer, args->equalizeRGB,
				args->coeff.offset, args->coeff.mul, args->coeff.scale);
done:
	if (retval) {
		free(args->coeff.offset); free(args->coeff.mul); free(args->coeff.scale);
		args->coeff.offset = args->coeff.mul = args->coeff.scale = NULL;
	}
	return args->retval = retval;
}
