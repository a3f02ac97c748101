/* TEST HARNESS ONLY — a miniature "Siril" around the glue (integration/siril/src/gpu/gpu_stack.c):
 * the functions of Siril the glue calls, backed by frames in memory, and harness_stack(), which
 * fills the reference's own `struct stacking_args` (src/stacking/stacking.h, compiled unmodified),
 * sets its `method` pointer the way command.c does, and calls do_normalization / args->method
 * through the same two symbols main_stack uses (stacking.c:98,104). */
#include <stdarg.h>
#include <stdio.h>
#include "core/siril.h"
#include "core/proto.h"
#include "core/processing.h"
#include "core/siril_log.h"
#include "io/sequence.h"
#include "io/image_format_fits.h"
#include "algos/statistics.h"
#include "registration/registration.h"
#include "stacking/stacking.h"
#include "gpu/gpu_stack.h"

/* ---- the frame store ---- */
typedef struct { const void *frames; int is_float; int opened; int reads; } store_t;	/* frames: [N][C][H][W], memory (FITS) row order */

static char logbuf[8192]; static size_t loglen; static int cancel_after = -1, polls;
struct gui_iface_t gui_iface;
static double last_progress;
static void progress(double f, const char *t) { (void) t; last_progress = f; }

char *siril_log_message(const char *format, ...) {
	va_list ap; va_start(ap, format);
	loglen += vsnprintf(logbuf + loglen, loglen < sizeof logbuf ? sizeof logbuf - loglen : 0, format, ap);
	va_end(ap);
	if (loglen >= sizeof logbuf) loglen = sizeof logbuf - 1;
	return NULL;
}
gboolean processing_should_continue(void) { polls++; return cancel_after < 0 || polls <= cancel_after; }
data_type get_data_type(int bitpix) { return bitpix == FLOAT_IMG ? DATA_FLOAT : ((bitpix == USHORT_IMG || bitpix == SHORT_IMG || bitpix == BYTE_IMG) ? DATA_USHORT : DATA_UNSUPPORTED); }
void translation_from_H(Homography H, double *dx, double *dy) { *dx = H.h02; *dy = -H.h12; }	/* registration.c:301-304 */

int new_fit_image(fits **fit, int width, int height, int nblayer, data_type type) {
	if (!*fit) { *fit = calloc(1, sizeof(fits)); if (!*fit) return 1; }
	fits *f = *fit;
	const size_t n = (size_t) width * height * nblayer;
	f->rx = width; f->ry = height; f->naxes[0] = width; f->naxes[1] = height; f->naxes[2] = nblayer; f->naxis = nblayer == 1 ? 2 : 3;
	f->type = type; f->bitpix = f->orig_bitpix = type == DATA_FLOAT ? FLOAT_IMG : USHORT_IMG;
	if (type == DATA_FLOAT) { f->fdata = calloc(n, sizeof(float)); if (!f->fdata) return 1; for (int l = 0; l < nblayer; l++) f->fpdata[l] = f->fdata + (size_t) l * width * height; }
	else { f->data = calloc(n, sizeof(WORD)); if (!f->data) return 1; for (int l = 0; l < nblayer; l++) f->pdata[l] = f->data + (size_t) l * width * height; }
	return 0;
}
void clearfits(fits *f) { if (!f) return; free(f->data); free(f->fdata); memset(f, 0, sizeof *f); }

int seq_open_image(sequence *seq, int index) { (void) index; ((store_t *) seq->harness)->opened++; return 0; }
void seq_close_image(sequence *seq, int index) { (void) index; ((store_t *) seq->harness)->opened--; }
int seq_read_frame(sequence *seq, int index, fits *dest, gboolean force_float, int thread_id) {
	(void) force_float; (void) thread_id;
	store_t *s = seq->harness;
	fits *d = dest;
	if (new_fit_image(&d, seq->rx, seq->ry, seq->nb_layers, s->is_float ? DATA_FLOAT : DATA_USHORT)) return 1;
	const size_t n = (size_t) seq->rx * seq->ry * seq->nb_layers, es = s->is_float ? 4 : 2;
	memcpy(s->is_float ? (void *) dest->fdata : (void *) dest->data, (const char *) s->frames + (size_t) index * n * es, n * es);
	s->reads++;
	return 0;
}
/* area in display coordinates: y counted from the top row (image_format_fits.c:939-944); the rows come
 * out in memory (FITS) order */
int seq_opened_read_region(sequence *seq, int layer, int index, void *buffer, const rectangle *area, int thread_id) {
	(void) thread_id;
	store_t *s = seq->harness;
	const size_t es = s->is_float ? 4 : 2, W = seq->rx, H = seq->ry;
	if (area->x != 0 || area->w != (int) W || area->y < 0 || area->h < 1 || area->y + area->h > (int) H) return 1;
	const size_t first = H - area->y - area->h;	/* memory row of the region's first row */
	const char *src = (const char *) s->frames + (((size_t) index * seq->nb_layers + layer) * H + first) * W * es;
	memcpy(buffer, src, (size_t) area->h * W * es);
	s->reads++;
	return 0;
}
void add_stats_to_seq(sequence *seq, int image_index, int layer, imstats *stat) {	/* statistics.c:506-528 */
	if (!seq->stats) seq->stats = calloc(3, sizeof(imstats **));
	if (!seq->stats[layer]) seq->stats[layer] = calloc(seq->number, sizeof(imstats *));
	free(seq->stats[layer][image_index]);
	seq->stats[layer][image_index] = stat;
}

/* ---- the reference's symbols, as Siril would carry them with -Dcuda=true ---- */
int stack_mean_with_rejection(struct stacking_args *args) { return gpu_stack_mean_or_median(args, TRUE); }
int stack_median(struct stacking_args *args) { return gpu_stack_mean_or_median(args, FALSE); }
int do_normalization(struct stacking_args *args) { return gpu_do_normalization(args); }

typedef struct {
	int N, C, H, W, is_float, is_mean, rejection; float sig[2];
	int normalize, lite_norm, weighting, ref_image, reglayer, equalizeRGB;
	const double *dx, *dy;		/* registration shifts per frame (H.h02, -H.h12), or NULL */
	int output_norm, use_32bit_output, create_rejmaps, merge_rejmaps;
	int cancel_after;		/* processing_should_continue() turns false after this many polls; < 0: never */
	int runs;			/* stack this many times on the same sequence (the statistics cache serves from the second) */
	const float *wfwhm; const int *nbstars;	/* regdata.weighted_fwhm / number_of_stars per frame (WFWHM_WEIGHT / NBSTARS_WEIGHT), or NULL */
} harness_params;

/* out: [C][H][W] float or WORD; rej_low / rej_high: WORD or NULL; coeff_out: 3 * C * N doubles (offset, mul, scale) or NULL.
 * info[0] frames still open, [1] frame / region reads of the LAST run, [2] statistics records cached, [3] polls.
 * Returns args.retval. */
int harness_stack(const harness_params *p, const void *frames, void *out, WORD *rej_low, WORD *rej_high,
		double *coeff_out, char *log_out, int log_cap, long info[4]) {
	store_t store = { frames, p->is_float, 0, 0 };
	sequence seq; memset(&seq, 0, sizeof seq);
	seq.number = seq.selnum = p->N; seq.nb_layers = p->C; seq.rx = p->W; seq.ry = p->H;
	seq.bitpix = p->is_float ? FLOAT_IMG : USHORT_IMG; seq.reference_image = p->ref_image; seq.harness = &store;
	regdata *rp = NULL;
	const int has_reg = p->dx || p->wfwhm || p->nbstars;
	if (has_reg) {
		seq.regparam = calloc(3, sizeof(regdata *));
		rp = calloc(p->N, sizeof(regdata));
		for (int i = 0; i < p->N; i++) {
			rp[i].H.h00 = rp[i].H.h11 = rp[i].H.h22 = 1.0;
			if (p->dx) { rp[i].H.h02 = p->dx[i]; rp[i].H.h12 = -p->dy[i]; }
			if (p->wfwhm) rp[i].weighted_fwhm = p->wfwhm[i];
			if (p->nbstars) rp[i].number_of_stars = p->nbstars[i];
		}
		seq.regparam[p->reglayer < 0 ? 0 : p->reglayer] = rp;
	}
	int *indices = malloc(p->N * sizeof(int));
	for (int i = 0; i < p->N; i++) indices[i] = i;
	gui_iface.set_progress = progress;
	int rv = 0;
	loglen = 0; logbuf[0] = 0;
	for (int run = 0; run < (p->runs > 0 ? p->runs : 1); run++) {
		struct stacking_args args; memset(&args, 0, sizeof args);
		args.method = p->is_mean ? stack_mean_with_rejection : stack_median;	/* command.c:11660-11680 */
		args.seq = &seq; args.ref_image = p->ref_image; args.image_indices = indices; args.nb_images_to_stack = p->N;
		args.normalize = p->normalize; args.lite_norm = p->lite_norm; args.equalizeRGB = p->equalizeRGB;
		args.type_of_rejection = p->rejection; args.sig[0] = p->sig[0]; args.sig[1] = p->sig[1];
		args.create_rejmaps = p->create_rejmaps; args.merge_lowhigh_rejmaps = p->merge_rejmaps;
		args.weighting_type = p->weighting; args.output_norm = p->output_norm; args.use_32bit_output = p->use_32bit_output;
		args.reglayer = has_reg ? (p->reglayer < 0 ? 0 : p->reglayer) : -1;
		cancel_after = p->cancel_after; polls = 0; store.reads = 0;
		/* main_stack (stacking.c:76-108): 1. normalization, 3. the method, through its pointer */
		args.retval = do_normalization(&args);
		if (!args.retval) {
			if (coeff_out && args.normalize != NO_NORM) {
				const size_t cn = (size_t) p->C * p->N;
				memcpy(coeff_out, args.coeff.offset, cn * 8); memcpy(coeff_out + cn, args.coeff.mul, cn * 8); memcpy(coeff_out + 2 * cn, args.coeff.scale, cn * 8);
			}
			args.retval = args.method(&args);
		}
		rv = args.retval;
		if (!rv) {
			const size_t n = (size_t) p->C * p->H * p->W;
			memcpy(out, p->use_32bit_output ? (void *) args.result.fdata : (void *) args.result.data, n * (p->use_32bit_output ? 4 : 2));
			if (rej_low && args.rejmap_low) memcpy(rej_low, args.rejmap_low->data, n * 2);
			if (rej_high && args.rejmap_high) memcpy(rej_high, args.rejmap_high->data, n * 2);
			clearfits(&args.result);	/* the caller of main_stack owns the result (command.c:11772,11812) */
			if (args.rejmap_low) { clearfits(args.rejmap_low); free(args.rejmap_low); }
			if (args.rejmap_high) { clearfits(args.rejmap_high); free(args.rejmap_high); }
		}
		free(args.critical_value);	/* command.c:11735 */
		if (args.coeff.offset || args.coeff.mul || args.coeff.scale || args.weights) rv = rv ? rv : -100;	/* the callee frees them */
		if (rv) break;
	}
	long cached = 0;
	if (seq.stats) for (int l = 0; l < 3; l++) if (seq.stats[l]) { for (int i = 0; i < p->N; i++) if (seq.stats[l][i]) { cached++; free(seq.stats[l][i]); } free(seq.stats[l]); }
	free(seq.stats);
	if (info) { info[0] = store.opened; info[1] = store.reads; info[2] = cached; info[3] = polls; }
	if (log_out && log_cap > 0) { strncpy(log_out, logbuf, log_cap - 1); log_out[log_cap - 1] = 0; }
	free(indices); free(rp); free(seq.regparam);
	return rv;
}
