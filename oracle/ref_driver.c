/* TEST INFRASTRUCTURE ONLY — driver around the *reference's own code* (oracle/_ref).
 *
 * Everything algorithmic in this TU is #included from oracle/_ref/gen/*.inc, i.e. line ranges
 * extracted at build time from /root/reference by oracle/build_ref.py (never committed), or
 * linked from the reference's whole TUs (rejection_float.c, sorting.c, siril_fit_linear.c,
 * rt_algo.cc).  What is written here is glue only:
 *   - ref_mean_and_reject : sets up struct _data_block / struct stacking_args and calls the
 *                           reference mean_and_reject (median_and_mean.c:956)
 *   - ref_stack           : restates the block loop median_and_mean.c:1552-1760 (open/read
 *                           replaced by in-RAM frames) around the reference's planner,
 *                           mean_and_reject, quickmedian* and rounding inlines
 *   - ref_image_stats     : restates the sequencing of statistics_internal_float
 *                           (statistics_float.c:281-481) / _ushort (statistics.c:233-443)
 *                           around the reference's own estimators
 *   - ref_compute_factors / ref_noise_weights / ref_planner / ref_norm_to_0_1 : thin wrappers
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product path never does.
 */
#include <string.h>
#include <math.h>
#include "core/siril.h"
#include "core/proto.h"
#include "core/processing.h"
#include <gsl/gsl_statistics_float.h>
#include <gsl/gsl_statistics_ushort.h>
#include <gsl/gsl_linalg.h>
#include <limits.h>
#include <stdio.h>

/* the shim fits/sequence are opaque in core/siril.h; the few members the extracted ranges
 * touch are provided by these layouts (same first members as the shim structs). */
#include "stacking/siril_fit_linear.h"
#include "stacking/stacking.h"
#include "algos/sorting.h"
#include "algos/statistics.h"

struct { int max_thread; } com = { 1 };

static int get_max_memory_in_MB(void) { return 1 << 20; }
static int processing_should_continue(void) { return 1; }

/* ---- reference code, extracted at build time ---- */
/* the planner logs every block to stdout (median_and_mean.c:344-346); keep the harness quiet */
#define fprintf(...) ((void) 0)
#include "gen/mm_planner.inc"        /* refine_blocks_candidate, stack_compute_parallel_blocks */
#undef fprintf
#include "gen/mm_norm01.inc"         /* norm_to_0_1_range */
#include "gen/mm_rejection_u16.inc"  /* u16 clippers, check_G_values, confirm_outliers,
                                        apply_rejection_ushort, mean_and_reject */
#include "gen/mm_noise_weights.inc"  /* compute_noise_weights */
static int layer_has_registration(const sequence *seq, int layer) { return seq->regparam && layer >= 0 && seq->regparam[layer]; }
#include "gen/mm_other_weights.inc"  /* compute_wfwhm_weights, compute_nbstars_weights */
#include "gen/stats_u16_sd_mad.inc"  /* siril_stats_ushort_sd_64/_32/_mad */
#include "gen/statsf_mad_bwmv.inc"   /* siril_stats_float_mad, siril_stats_float_bwmv */
#include "gen/statsf_ikss.inc"       /* IKSSlite */
#include "gen/quant_defs.inc"        /* SIGMA_CLIP, NITER */
static int FnCompare_double(const void *, const void *);
#include "gen/quant_meansigma.inc"   /* FnMeanSigma_ushort, FnMeanSigma_float */
#include "gen/quant_diffmeansigma.inc" /* FnDiffMeanSigma_int, FnDiffMeanSigma_float */
#include "gen/quant_noise1.inc"      /* FnNoise1_ushort, FnNoise1_float, FnCompare_double */
#include "gen/norm_factors.inc"      /* compute_factors_from_estimators */
#include "gen/blend_ramp.inc"        /* RAMP_PACE, init_ramp, get_ramped_value */

/* dead helper symbols sorting.c's strcompare() references */
void g_free(void *p) { free(p); }
int g_strcmp0(const char *a, const char *b) { return strcmp(a ? a : "", b ? b : ""); }
char *g_utf8_collate_key_for_filename(const char *s, long len) { (void)len; return strdup(s); }

/* ------------------------------------------------------------------------------------ */

static void setup_block(struct _data_block *blk, int is_float, int N, int rejtype) {
	size_t es = is_float ? sizeof(float) : sizeof(WORD);
	memset(blk, 0, sizeof(*blk));
	blk->stack = calloc(N + 1, es);
	blk->o_stack = calloc(N + 1, es);
	blk->w_stack = calloc(N + 1, es);
	blk->rejected = calloc(N + 1, sizeof(int));
	blk->xf = calloc(N + 1, sizeof(float));
	blk->yf = calloc(N + 1, sizeof(float));
	blk->mstack = calloc(N + 1, sizeof(float));
	blk->dstack = calloc(N + 1, sizeof(float));
	if (rejtype == LINEARFIT) {
		struct _data_block *data_pool = blk;
		int i = 0, nb_frames = N;
#include "gen/mm_linfit_precalc.inc"
	}
}

static void free_block(struct _data_block *blk) {
	free(blk->stack); free(blk->o_stack); free(blk->w_stack);
	free(blk->rejected); free(blk->xf); free(blk->yf); free(blk->mstack); free(blk->dstack);
}

static void setup_args(struct stacking_args *args, int N, int rejtype, float s0, float s1,
		const float *crit, const double *weights) {
	memset(args, 0, sizeof(*args));
	args->nb_images_to_stack = N;
	args->type_of_rejection = (rejection) rejtype;
	args->sig[0] = s0; args->sig[1] = s1;
	args->critical_value = (float *) crit;
	args->weighting_type = weights ? NOISE_WEIGHT : NO_WEIGHT;
	args->weights = (double *) weights;
	args->sd_calculator = N < 65536 ? siril_stats_ushort_sd_32 : siril_stats_ushort_sd_64;
	args->mad_calculator = siril_stats_ushort_mad;
}

/* one pixel stack through the reference's mean_and_reject (median_and_mean.c:956) */
double ref_mean_and_reject(int is_float, const void *stack, int N, int rejtype, float s0, float s1,
		const float *crit, const double *weights, int layer, int rej[2]) {
	struct _data_block blk;
	struct stacking_args args;
	setup_block(&blk, is_float, N, rejtype);
	setup_args(&args, N, rejtype, s0, s1, crit, weights);
	blk.layer = layer;
	memcpy(blk.stack, stack, (size_t) N * (is_float ? sizeof(float) : sizeof(WORD)));
	rej[0] = rej[1] = 0;
	double m = mean_and_reject(&args, &blk, N, is_float ? DATA_FLOAT : DATA_USHORT, rej);
	free_block(&blk);
	return m;
}

typedef struct {
	int is_float, N, C, H, W;
	int is_mean, rejtype;
	float sig[2];
	const float *crit;
	int normalize;
	const double *scale, *offset, *mul;	/* [C*N], layer-major like coeff.pscale[layer][frame] */
	const double *weights;			/* [C*N] or NULL */
	const int *shiftx, *shifty;		/* [N] or NULL (integer registration shifts) */
	int output_norm, use_32bit_output;
	int create_rejmaps, merge_rejmaps;
	int nthreads;
	long max_rows;				/* planner budget; <=0: whole image */
	int source_bitpix;			/* seq->bitpix; 8 (BYTE_IMG): normalize_to16bit */
	const float *mask;			/* [N][H][W] up-scaled distance masks (data->mask[frame] before :513-521) or NULL */
	const float *drizz;			/* [N][C][H][W] drizzle weights (data->drizz[frame]) or NULL */
	int feather_dist, drizzle;		/* args->feather_dist (> 0: masking), args->drizzle */
} ref_stack_params;

/* the mask normalization stack_read_block_data applies to a block it has just read
 * (median_and_mean.c:513-521, the reference's own lines), here on whole planes */
static void ref_normalize_masks(float **planes, int N, long H, long W, int feather_dist) {
	struct stacking_args a_, *args = &a_;
	struct _image_block b_, *my_block = &b_;
	struct _data_block d_, *data = &d_;
	long naxes[3] = { W, H, 1 };
	a_.feather_dist = feather_dist; b_.height = H; d_.mask = planes;
	init_ramp();
	for (int frame = 0; frame < N; frame++) {
#include "gen/mm_mask_ramp.inc"
	}
}

/* Restatement of the block loop of stack_mean_or_median (median_and_mean.c:1552-1760) on
 * in-RAM frames laid out [N][C][H][W] in memory (FITS, bottom-up) row order; `out` is
 * [C][H][W] in the same order (the reference reads rows top-down and writes them back
 * bottom-up, :1597 — a wash in memory order).  y-shifts are what stack_read_block_data
 * (:407-448) applies at read time, x-shifts as :1617-1636. */
typedef struct {
	const ref_stack_params *p; const void *frames; void *out; WORD *rej_low, *rej_high;
	struct _image_block *blocks; struct _data_block *pool; struct stacking_args *args;
	uint64_t *brej;	/* [nb_blocks][2] */
	float **mask;	/* [N] normalized mask planes, or NULL */
} ref_stack_ctx;

/* provided by ref_omp.c (the only TU of the pinned build compiled with -fopenmp): the
 * reference's `omp parallel for schedule(dynamic)` over blocks (median_and_mean.c:1552) */
void ref_parallel_for(int n, int nthreads, void (*fn)(int, int, void *), void *ctx);

static void ref_stack_block(int b, int tid, void *vctx) {
	ref_stack_ctx *cx = vctx;
	const ref_stack_params *p = cx->p;
	const void *frames = cx->frames; void *out = cx->out;
	WORD *rej_low = cx->rej_low, *rej_high = cx->rej_high;
	struct stacking_args *args = cx->args;
	const int N = p->N, C = p->C;
	const long H = p->H, W = p->W;
	const size_t plane = (size_t) H * W;
	const data_type itype = p->is_float ? DATA_FLOAT : DATA_USHORT;
	{
		struct _image_block *blk = cx->blocks + b;
		struct _data_block *data = &cx->pool[tid];
		const int layer = (int) blk->channel;
		data->layer = layer;
		uint64_t brej[2] = { 0, 0 };
		for (long y = 0; y < blk->height; y++) {
			const long yt = blk->start_row + y;		/* top-down row */
			const long r = H - 1 - yt;			/* memory row of the output */
			for (long x = 0; x < W; x++) {
				for (int f = 0; f < N; f++) {
					long sx = x, sr = r;
					int inside = 1;
					if (p->shiftx) { sx = x - p->shiftx[f]; if (sx < 0 || sx >= W) inside = 0; }
					if (p->shifty) { sr = r - p->shifty[f]; if (sr < 0 || sr >= H) inside = 0; }
					WORD pixel = 0; float fpixel = 0.f;
					if (inside) {	/* outside: pixel stays 0, which every normalization maps to 0 */
						size_t idx = ((size_t) f * C + layer) * plane + (size_t) sr * W + sx;
						if (p->is_float) fpixel = ((const float *) frames)[idx];
						else pixel = ((const WORD *) frames)[idx];
						/* median_and_mean.c:1687-1692 (outside the frame the reference `continue`s before these lines) */
						if (cx->mask) data->mstack[f] = cx->mask[f][(size_t) sr * W + sx];
						if (args->drizzle) data->dstack[f] = p->drizz[idx];
					}
					double tmp;
					const double sc = p->scale ? p->scale[layer * N + f] : 1.0;
					switch (p->normalize) {
					default:
					case NO_NORM:
						if (p->is_float) ((float *) data->stack)[f] = fpixel;
						else ((WORD *) data->stack)[f] = pixel;
						break;
					case ADDITIVE:
					case ADDITIVE_SCALING:
						if (p->is_float) {
							if (fpixel != 0.f) {
								tmp = fpixel * sc;
								((float *) data->stack)[f] = (float) (tmp - p->offset[layer * N + f]);
							} else ((float *) data->stack)[f] = 0.f;
						} else {
							if (pixel > 0) {
								tmp = (double) pixel * sc;
								((WORD *) data->stack)[f] = round_to_WORD(tmp - p->offset[layer * N + f]);
							} else ((WORD *) data->stack)[f] = 0;
						}
						break;
					case MULTIPLICATIVE:
					case MULTIPLICATIVE_SCALING:
						if (p->is_float) {
							tmp = fpixel * sc;
							((float *) data->stack)[f] = (float) (tmp * p->mul[layer * N + f]);
						} else {
							tmp = (double) pixel * sc;
							((WORD *) data->stack)[f] = round_to_WORD(tmp * p->mul[layer * N + f]);
						}
						break;
					}
				}
				double result;
				const size_t o = (size_t) layer * plane + (size_t) r * W + x;
				if (p->is_mean) {
					int rej[2] = { 0, 0 };
					result = mean_and_reject(args, data, N, itype, rej);
					brej[0] += rej[0]; brej[1] += rej[1];
					if (p->create_rejmaps) {
						if (p->merge_rejmaps) rej_low[o] = truncate_to_WORD(rej[0] + rej[1]);
						else { rej_low[o] = truncate_to_WORD(rej[0]); rej_high[o] = truncate_to_WORD(rej[1]); }
					}
				} else {
					result = p->is_float ? quickmedian_float(data->stack, N) : quickmedian(data->stack, N);
				}
				if (p->use_32bit_output) {
					float v;
					if (!p->is_float)
						v = p->output_norm ? double_ushort_to_float_range(result)
							: set_float_in_interval(double_ushort_to_float_range(result), 0.f, 1.f);
					else
						v = p->output_norm ? (float) result : set_float_in_interval((float) result, 0.f, 1.f);
					((float *) out)[o] = v;
				} else {
					if (p->output_norm && p->source_bitpix == 8)	/* normalize_to16bit (median_and_mean.c:547-555, :1729-1732) */
						result *= (USHRT_MAX_DOUBLE / UCHAR_MAX_DOUBLE);
					((WORD *) out)[o] = round_to_WORD(result);
				}
			}
		}
		cx->brej[2 * b] = brej[0]; cx->brej[2 * b + 1] = brej[1];
	}
}

int ref_stack(const ref_stack_params *p, const void *frames, void *out,
		WORD *rej_low, WORD *rej_high, uint64_t irej[6]) {
	const int N = p->N, C = p->C;
	const long H = p->H, W = p->W;
	long naxes[3] = { W, H, C };
	struct _image_block *blocks = NULL;
	long largest; int nb_blocks;
	int nthreads = p->nthreads > 0 ? p->nthreads : 1;
	long max_rows = p->max_rows > 0 ? p->max_rows : H * C;
	if (stack_compute_parallel_blocks(&blocks, max_rows, naxes, nthreads, &largest, &nb_blocks))
		return -1;
	struct _data_block *pool = calloc(nthreads, sizeof(*pool));
	for (int t = 0; t < nthreads; t++) setup_block(&pool[t], p->is_float, N, p->rejtype);
	struct stacking_args args;
	setup_args(&args, N, p->rejtype, p->sig[0], p->sig[1], p->crit, p->weights);
	float **mask = NULL;
	if (p->is_mean && p->feather_dist > 0 && p->mask) {
		args.feather_dist = p->feather_dist;
		mask = calloc(N, sizeof(float *));
		for (int f = 0; f < N; f++) {
			mask[f] = malloc((size_t) H * W * sizeof(float));
			memcpy(mask[f], p->mask + (size_t) f * H * W, (size_t) H * W * sizeof(float));
		}
		ref_normalize_masks(mask, N, H, W, p->feather_dist);
	}
	args.drizzle = p->is_mean && p->drizzle && p->drizz;
	memset(irej, 0, 6 * sizeof(uint64_t));
	ref_stack_ctx cx = { p, frames, out, rej_low, rej_high, blocks, pool, &args,
		calloc(2 * (size_t) nb_blocks, sizeof(uint64_t)), mask };
	ref_parallel_for(nb_blocks, nthreads, ref_stack_block, &cx);
	if (p->is_mean && p->rejtype != NO_REJEC)
		for (int b = 0; b < nb_blocks; b++) {
			irej[blocks[b].channel * 2 + 0] += cx.brej[2 * b];
			irej[blocks[b].channel * 2 + 1] += cx.brej[2 * b + 1];
		}
	free(cx.brej);
	if (p->use_32bit_output && p->output_norm) {
		fits fit; memset(&fit, 0, sizeof(fit));
		fit.naxes[0] = W; fit.naxes[1] = H; fit.naxes[2] = C; fit.fdata = (float *) out;
		norm_to_0_1_range(&fit);
	}
	for (int t = 0; t < nthreads; t++) free_block(&pool[t]);
	free(pool); free(blocks);
	if (mask) { for (int f = 0; f < N; f++) free(mask[f]); free(mask); }
	return 0;
}

/* out[14] = total, ngoodpix, mean, median, sigma, avgDev, mad, sqrtbwmv, location, scale,
 *           min, max, normValue, bgnoise  (imstats order, core/siril.h:407-413).
 * Sequencing restated from statistics_internal_float (statistics_float.c:362-477) and
 * statistics_internal_ushort (statistics.c:311-438) for option STATS_NORM / STATS_LITENORM. */
int ref_image_stats(int is_float, const void *plane, long nx, long ny, int lite, double out[14]) {
	const size_t total = (size_t) nx * ny;
	long ngood = 0; int status = 0;
	double mean = 0, sigma = 0, bgnoise = 0, median, mad, location = 0, scale = 0, minv, maxv, normValue;
	for (int i = 0; i < 14; i++) out[i] = 0;
	if (is_float) {
		const float *data = plane;
		normValue = 1.0;
		float mn = data[0], mx = data[0];
		for (size_t i = 0; i < total; i++) if (!isnan(data[i])) { mn = min(mn, data[i]); mx = max(mx, data[i]); }
		minv = mn; maxv = mx;
		FnMeanSigma_float((float *) data, total, &ngood, &mean, &sigma, &status);
		FnNoise1_float((float *) data, nx, ny, &bgnoise, SINGLE_THREADED, &status);
		if (ngood == 0) return 1;
		float *nz = malloc(ngood * sizeof(float));
		size_t j = 0;
		for (size_t i = 0; i < total; i++) if (data[i] != 0.f && !isnan(data[i])) nz[j++] = data[i];
		median = histogram_median_float(nz, ngood, SINGLE_THREADED);
		mad = siril_stats_float_mad(nz, ngood, median, SINGLE_THREADED, NULL);
		if (!lite) {
			if (IKSSlite(nz, ngood, median, mad, &location, &scale, SINGLE_THREADED)) { free(nz); return 2; }
		}
		free(nz);
	} else {
		const WORD *data = plane;
		normValue = USHRT_MAX_DOUBLE;
		WORD mn = data[0], mx = data[0];
		for (size_t i = 0; i < total; i++) { mn = min(mn, data[i]); mx = max(mx, data[i]); }
		minv = mn; maxv = mx;
		FnMeanSigma_ushort((WORD *) data, total, &ngood, &mean, &sigma, &status);
		FnNoise1_ushort((WORD *) data, nx, ny, &bgnoise, SINGLE_THREADED, &status);
		if (ngood == 0) return 1;
		WORD *nz = malloc(ngood * sizeof(WORD));
		size_t j = 0;
		for (size_t i = 0; i < total; i++) if (data[i] != 0) nz[j++] = data[i];
		median = histogram_median(nz, ngood, SINGLE_THREADED);
		mad = siril_stats_ushort_mad(nz, ngood, median, SINGLE_THREADED);
		if (!lite) {
			float *nd = malloc(ngood * sizeof(float));
			float inv = (float) (1.0 / normValue);
			for (long i = 0; i < ngood; i++) nd[i] = (float) nz[i] * inv;
			float med = (float) median * inv, madf = (float) mad * inv;
			if (IKSSlite(nd, ngood, med, madf, &location, &scale, SINGLE_THREADED)) { free(nd); free(nz); return 2; }
			location *= normValue; scale *= normValue;
			free(nd);
		}
		free(nz);
	}
	out[0] = (double) total; out[1] = (double) ngood; out[2] = mean; out[3] = median; out[4] = sigma;
	out[6] = mad; out[8] = location; out[9] = scale; out[10] = minv; out[11] = maxv;
	out[12] = normValue; out[13] = bgnoise;
	return 0;
}

/* estimators in, factors out (normalization.c:144-181); arrays are [C*N] layer-major */
void ref_compute_factors(int normalize, int nb_layers, int N, int ref_index, int reglayer,
		int equalizeRGB, double *offset, double *mul, double *scale) {
	struct stacking_args args; sequence seq;
	memset(&args, 0, sizeof(args)); memset(&seq, 0, sizeof(seq));
	seq.nb_layers = nb_layers; args.seq = &seq;
	args.nb_images_to_stack = N; args.normalize = (normalization) normalize;
	args.reglayer = reglayer; args.equalizeRGB = equalizeRGB;
	for (int l = 0; l < nb_layers; l++) {
		args.coeff.poffset[l] = offset + l * N;
		args.coeff.pmul[l] = mul + l * N;
		args.coeff.pscale[l] = scale + l * N;
	}
	compute_factors_from_estimators(&args, ref_index);
}

/* median_and_mean.c:1111-1135; scale and bgnoise are [C*N] */
void ref_noise_weights(int nb_layers, int N, const double *scale, const double *bgnoise, double *weights) {
	struct stacking_args args; sequence seq;
	memset(&args, 0, sizeof(args)); memset(&seq, 0, sizeof(seq));
	seq.nb_layers = nb_layers; args.seq = &seq; args.nb_images_to_stack = N;
	int *idx = malloc(N * sizeof(int));
	imstats *st = calloc((size_t) nb_layers * N, sizeof(imstats));
	imstats **pst = malloc((size_t) nb_layers * N * sizeof(imstats *));
	imstats **layers[3] = { 0 };
	for (int i = 0; i < N; i++) idx[i] = i;
	for (int l = 0; l < nb_layers; l++) {
		layers[l] = pst + l * N;
		args.coeff.pscale[l] = (double *) scale + l * N;
		for (int i = 0; i < N; i++) { st[l * N + i].bgnoise = bgnoise[l * N + i]; pst[l * N + i] = &st[l * N + i]; }
	}
	seq.stats = layers; args.image_indices = idx;
	compute_noise_weights(&args);
	memcpy(weights, args.weights, (size_t) nb_layers * N * sizeof(double));
	free(args.weights); free(idx); free(st); free(pst);
}

/* median_and_mean.c:1137-1230 on the selected frames' registration data; returns the reference's ST_* code */
int ref_reg_weights(int use_nbstars, int nb_layers, int N, const float *wfwhm, const int *nbstars, double *weights) {
	struct stacking_args args; sequence seq;
	memset(&args, 0, sizeof(args)); memset(&seq, 0, sizeof(seq));
	regdata *rd = calloc(N, sizeof(regdata)), *layers[1] = { rd };
	int *idx = malloc(N * sizeof(int));
	for (int i = 0; i < N; i++) { idx[i] = i; rd[i].weighted_fwhm = wfwhm ? wfwhm[i] : 0.f; rd[i].number_of_stars = nbstars ? nbstars[i] : 0; }
	seq.nb_layers = nb_layers; seq.regparam = layers; args.seq = &seq; args.reglayer = 0;
	args.nb_images_to_stack = N; args.image_indices = idx;
	int rv = use_nbstars ? compute_nbstars_weights(&args) : compute_wfwhm_weights(&args);
	if (args.weights) memcpy(weights, args.weights, (size_t) nb_layers * N * sizeof(double));
	free(args.weights); free(idx); free(rd);
	return rv;
}

/* ---- SER frames: the reference's own sample conversion (io/ser.c:521-539, enums io/ser.h:70-85) ---- */
#include <endian.h>
#define be16_to_cpu(x) be16toh(x)
#define le16_to_cpu(x) le16toh(x)
#include "gen/ser_enums.inc"
struct ser_struct { ser_pixdepth byte_pixel_depth; ser_endian little_endian; };
#include "gen/ser_endian_depth.inc"
/* data: frame_size samples as read from the file into a WORD buffer (8-bit samples packed at its start) */
void ref_ser_convert(int byte_depth, int little_endian_field, WORD *data, long frame_size) {
	struct ser_struct s = { (ser_pixdepth) byte_depth, (ser_endian) little_endian_field };
	ser_manage_endianess_and_depth(&s, data, frame_size);
}

/* ---- maximized framing: the reference's own lines of stack_open_all_files (median_and_mean.c:161-170 per frame,
 * :182-189 after the loop) around in-RAM registration data ---- */
#include <float.h>
int ref_maximized_framing(int nb_frames, const int *rxs, const int *rys, const double *h02, const double *h12, int upscale,
		long naxes_out[2], int offset_out[2]) {
	struct stacking_args a_, *args = &a_; sequence seq_;
	memset(&a_, 0, sizeof a_); memset(&seq_, 0, sizeof seq_);
	regdata *rd = calloc(nb_frames, sizeof(regdata)); imgdata *im = calloc(nb_frames, sizeof(imgdata));
	for (int i = 0; i < nb_frames; i++) { rd[i].H.h02 = h02[i]; rd[i].H.h12 = h12[i]; im[i].rx = rxs[i]; im[i].ry = rys[i]; }
	regdata *layers[1] = { rd };
	seq_.regparam = layers; seq_.is_variable = TRUE; seq_.imgparam = im; a_.seq = &seq_; a_.reglayer = 0; a_.maximize_framing = TRUE;
	double xmin = DBL_MAX, xmax = -DBL_MAX, ymin = DBL_MAX, ymax = -DBL_MAX;
	double scale = upscale ? 2. : 1.;
	long naxes[3] = { 0, 0, 1 };
	for (int image_index = 0; image_index < nb_frames; image_index++) {
#include "gen/mm_maxframe_accum.inc"
	}
	gboolean update_wcs = TRUE; (void) update_wcs;
#include "gen/mm_maxframe_final.inc"
	}
	naxes_out[0] = naxes[0]; naxes_out[1] = naxes[1]; offset_out[0] = a_.offset[0]; offset_out[1] = a_.offset[1];
	free(rd); free(im);
	return 0;
}

/* ---- the statistics records of a .seq file: the reference's own fprintf (seqfile.c:838-853) and sscanf (:584-600) ---- */
typedef struct { imstats ***stats; int cfa_opened_monochrome; } ref_seq_view;
int ref_seq_stats_write(int layer, int i, const double v[14], char *out, size_t cap) {
	imstats st; memset(&st, 0, sizeof st);
	st.total = (long) v[0]; st.ngoodpix = (long) v[1]; st.mean = v[2]; st.median = v[3]; st.sigma = v[4]; st.avgDev = v[5];
	st.mad = v[6]; st.sqrtbwmv = v[7]; st.location = v[8]; st.scale = v[9]; st.min = v[10]; st.max = v[11]; st.normValue = v[12]; st.bgnoise = v[13];
	imstats *row_[1] = { &st }; imstats **row = row_ - i;
	imstats **layers_[3] = { 0, 0, 0 }; layers_[layer] = row;
	ref_seq_view sv = { layers_, 0 }, *seq = &sv;
	char *buf = NULL; size_t len = 0;
	FILE *seqfile = open_memstream(&buf, &len);
#include "gen/seq_stats_write.inc"
	fclose(seqfile);
	if (len + 1 > cap) { free(buf); return 1; }
	memcpy(out, buf, len + 1); free(buf);
	return 0;
}
int ref_seq_stats_read(const char *line, int *image_out, double v[14]) {
	imstats st, *stats = &st; memset(&st, 0, sizeof st);
	int image = -1, nb_tokens;
#include "gen/seq_stats_read.inc"
	*image_out = image;
	v[0] = (double) st.total; v[1] = (double) st.ngoodpix; v[2] = st.mean; v[3] = st.median; v[4] = st.sigma; v[5] = st.avgDev;
	v[6] = st.mad; v[7] = st.sqrtbwmv; v[8] = st.location; v[9] = st.scale; v[10] = st.min; v[11] = st.max; v[12] = st.normValue; v[13] = st.bgnoise;
	return nb_tokens;
}

/* ---- overlap normalization: the reference's own lines around in-RAM data ---- */
#include "gen/reg_translation_from_H.inc"  /* translation_from_H */
#include "gen/norm_solve_overlap.inc"      /* solve_overlap_coeffs (LU through oracle/shim/gsl/gsl_linalg.h) */
#include "gen/norm_compute_overlap.inc"    /* compute_overlap */

/* normalization.c:420-456 for two frames of sizes (rxi, ryi), (rxj, ryj) shifted by (dx, dy) = translation_from_H */
size_t ref_compute_overlap(int rxi, int ryi, int rxj, int ryj, double dxi, double dyi, double dxj, double dyj,
		int areai_out[4], int areaj_out[4]) {
	struct stacking_args args; sequence seq; regdata rd[2]; imgdata im[2] = { { rxi, ryi }, { rxj, ryj } };
	memset(&args, 0, sizeof args); memset(&seq, 0, sizeof seq); memset(rd, 0, sizeof rd);
	rd[0].H.h02 = dxi; rd[0].H.h12 = -dyi; rd[1].H.h02 = dxj; rd[1].H.h12 = -dyj;
	regdata *layers[1] = { rd };
	seq.regparam = layers; seq.is_variable = TRUE; seq.imgparam = im; seq.number = 2; seq.nb_layers = 1;
	args.seq = &seq; args.reglayer = 0;
	rectangle ai = { 0 }, aj = { 0 };
	size_t n = compute_overlap(&args, 0, 1, &ai, &aj);
	areai_out[0] = ai.x; areai_out[1] = ai.y; areai_out[2] = ai.w; areai_out[3] = ai.h;
	areaj_out[0] = aj.x; areaj_out[1] = aj.y; areaj_out[2] = aj.w; areaj_out[3] = aj.h;
	return n;
}

/* the per-layer statistics of one pair (normalization.c:549-593) on the two overlap regions as
 * readfits_partial_all_layers delivers them ([h][w], nbdata samples each); out[0] = Nij, then
 * medij, medji, madij, madji, locij, locji, scaij, scaji */
int ref_overlap_pair(int is_float, const void *region_i, const void *region_j, size_t nbdata, int lite, double out[9]) {
	struct stacking_args a_, *args = &a_; sequence seq_, *seq = &seq_;
	memset(&a_, 0, sizeof a_); memset(&seq_, 0, sizeof seq_);
	a_.lite_norm = lite; a_.seq = seq;
	overlap_stats_t os; memset(&os, 0, sizeof os);
	overlap_stats_t *ol[1] = { &os };
	seq_.ostats = ol;
	const int nb_layers = 1, ijth = 0, i = 0, j = 1;
	gboolean needs_recalc_lite[3] = { TRUE, 0, 0 }, needs_recalc_detailed[3] = { !lite, 0, 0 };
	fits fiti, fitj; memset(&fiti, 0, sizeof fiti); memset(&fitj, 0, sizeof fitj);
	fiti.type = fitj.type = is_float ? DATA_FLOAT : DATA_USHORT;
	fiti.fpdata[0] = (float *) region_i; fitj.fpdata[0] = (float *) region_j;
	fiti.pdata[0] = (WORD *) region_i; fitj.pdata[0] = (WORD *) region_j;
	float invnorm = 1. / USHRT_MAX;
	float *datai = malloc(nbdata * sizeof(float)), *dataj = malloc(nbdata * sizeof(float));
	size_t Nij = 0;
	(void) i; (void) j;
#include "gen/norm_overlap_layers.inc"
	free(datai); free(dataj);
	out[0] = (double) os.Nij; out[1] = os.medij; out[2] = os.medji; out[3] = os.madij; out[4] = os.madji;
	out[5] = os.locij; out[6] = os.locji; out[7] = os.scaij; out[8] = os.scaji;
	return 0;
}

/* the three solves of compute_normalization_overlaps (normalization.c:875-906) for one layer; Nij_in / Mij_in / Sij_in
 * nb_frames x nb_frames row-major; offset / mul / scale [nb_frames] come in initialized (0 / 1 / 1, init_coeffs) */
int ref_overlap_solve(int normalize, int nb_frames, int index_ref, const size_t *Nij_in, const double *Mij_in, const double *Sij_in,
		double *offset, double *mul, double *scale) {
	struct stacking_args a_, *args = &a_;
	memset(&a_, 0, sizeof a_);
	a_.normalize = (normalization) normalize;
	norm_coeff *coeff = &a_.coeff;
	coeff->poffset[0] = offset; coeff->pmul[0] = mul; coeff->pscale[0] = scale;
	const int nb_layers = 1, N = nb_frames - 1;
	size_t **rowsN = malloc(nb_frames * sizeof(size_t *)); double **rowsM = malloc(nb_frames * sizeof(double *)), **rowsS = malloc(nb_frames * sizeof(double *));
	size_t *nn = malloc((size_t) nb_frames * nb_frames * sizeof(size_t)); double *mm = malloc((size_t) nb_frames * nb_frames * sizeof(double)), *ss = malloc((size_t) nb_frames * nb_frames * sizeof(double));
	memcpy(nn, Nij_in, (size_t) nb_frames * nb_frames * sizeof(size_t)); memcpy(mm, Mij_in, (size_t) nb_frames * nb_frames * sizeof(double)); memcpy(ss, Sij_in, (size_t) nb_frames * nb_frames * sizeof(double));
	for (int k = 0; k < nb_frames; k++) { rowsN[k] = nn + (size_t) k * nb_frames; rowsM[k] = mm + (size_t) k * nb_frames; rowsS[k] = ss + (size_t) k * nb_frames; }
	size_t **Nij[1] = { rowsN }; double **Mij[1] = { rowsM }, **Sij[1] = { rowsS };
	int *index = malloc(nb_frames * sizeof(int)); double *coeffs = calloc(N > 0 ? N : 1, sizeof(double));
	for (int k = 0, c = 0; k < nb_frames; k++) if (k != index_ref) index[c++] = k;	/* :779-782 */
#include "gen/norm_overlap_solves.inc"
	free(rowsN); free(rowsM); free(rowsS); free(nn); free(mm); free(ss); free(index); free(coeffs);
	return 0;
}

/* blocks_out: 4 longs per block (channel, start_row, end_row, height) */
int ref_planner(long max_rows, long W, long H, long C, int nb_threads, long *blocks_out, int max_blocks,
		long *largest, int *nb_blocks) {
	long naxes[3] = { W, H, C };
	struct _image_block *blocks = NULL;
	int rv = stack_compute_parallel_blocks(&blocks, max_rows, naxes, nb_threads, largest, nb_blocks);
	if (!rv) for (int i = 0; i < *nb_blocks && i < max_blocks; i++) {
		blocks_out[4 * i] = blocks[i].channel; blocks_out[4 * i + 1] = blocks[i].start_row;
		blocks_out[4 * i + 2] = blocks[i].end_row; blocks_out[4 * i + 3] = blocks[i].height;
	}
	free(blocks);
	return rv;
}

void ref_norm_to_0_1(float *data, long W, long H, long C) {
	fits fit; memset(&fit, 0, sizeof(fit));
	fit.naxes[0] = W; fit.naxes[1] = H; fit.naxes[2] = C; fit.fdata = data;
	norm_to_0_1_range(&fit);
}

float ref_float_sd(const float *d, int N, float *m) { return siril_stats_float_sd(d, N, m); }
int ref_apply_rejection_float(float *stack, int N, int rejtype, float s0, float s1, const float *crit, int rej[2]) {
	struct _data_block blk; struct stacking_args args;
	setup_block(&blk, 1, N, rejtype);
	setup_args(&args, N, rejtype, s0, s1, crit, NULL);
	memcpy(blk.stack, stack, N * sizeof(float));
	rej[0] = rej[1] = 0;
	int kept = apply_rejection_float(&blk, N, &args, rej);
	memcpy(stack, blk.stack, N * sizeof(float));
	free_block(&blk);
	return kept;
}
