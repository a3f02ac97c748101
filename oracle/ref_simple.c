/* TEST INFRASTRUCTURE ONLY — glue around the reference's sum / min / max stackers
 * (src/stacking/sum.c, src/stacking/stackminmax.c).  The per-image accumulation loops, the
 * search for the maximum and the output conversion are #included from oracle/_ref/gen/*.inc,
 * i.e. extracted at build time from /root/reference by oracle/build_ref.py (never committed).
 * What is written here: stand-in structs with the members those lines touch, the loop over
 * the frames (the reference's generic sequence worker calls the image hook once per frame),
 * and the registration shift handed in as integers (the reference rounds them from the
 * homography with round_to_int, sum.c:160-173).
 * Frames are [N][C][H][W]; outputs [C][H][W].  Frames are accumulated in index order, which is
 * what the reference does with one worker thread (with several, its `omp atomic` additions of
 * doubles happen in an unspecified order).
 */
#include <stdlib.h>
#include <string.h>
#include <limits.h>
#include "core/siril.h"
#include "core/proto.h"
#include "stacking/stacking.h"

#define siril_log_warning(...) do { } while (0)

struct simple_seq { int nb_layers; int is_drizzle; };
struct simple_args { struct simple_seq *seq; };
struct simple_fit { unsigned int rx, ry; float *fpdata[3]; WORD *pdata[3]; };
struct simple_sum {
	guint64 *sum[3]; double *fsum[3]; double *fweight[3];
	int input_32bits, output_32bits; int output_size[2];
};
struct simple_minmax {
	WORD *final_pixel[3]; float *ffinal_pixel[3];
	int input_32bits, ismax; int output_size[2];
};

/* sum.c:54-232 + :236-344 — returns 0, or ST_GENERIC_ERROR when the sum image is all zero */
int ref_stack_sum(int is_float, const void *frames, int N, int C, int H, int W,
		const int *shiftx_, const int *shifty_, int output_32bits, void *out) {
	struct simple_seq seq = { C, 0 };
	struct simple_args args_ = { &seq }, *args = &args_;
	struct simple_sum ss, *ssdata = &ss;
	const size_t np = (size_t) H * W;
	memset(&ss, 0, sizeof ss);
	ss.input_32bits = is_float; ss.output_32bits = output_32bits;
	ss.output_size[0] = W; ss.output_size[1] = H;
	if (is_float) { ss.fsum[0] = calloc(np * C, sizeof(double)); for (int l = 1; l < C; l++) ss.fsum[l] = ss.fsum[0] + np * l; }
	else { ss.sum[0] = calloc(np * C, sizeof(guint64)); for (int l = 1; l < C; l++) ss.sum[l] = ss.sum[0] + np * l; }
	const int is_drizzle = 0;
	float **weights = NULL;
	for (int f = 0; f < N; f++) {
		struct simple_fit fit_, *fit = &fit_;
		fit_.rx = W; fit_.ry = H;
		for (int l = 0; l < C; l++) {
			fit_.fpdata[l] = is_float ? (float *) frames + ((size_t) f * C + l) * np : NULL;
			fit_.pdata[l] = is_float ? NULL : (WORD *) frames + ((size_t) f * C + l) * np;
		}
		const int shiftx = shiftx_ ? shiftx_[f] : 0, shifty = shifty_ ? shifty_[f] : 0;
		int nx, ny, x, y, layer;
		size_t ii, pixel = 0;
		size_t nbdata = np;
#include "gen/sum_pixel_loop.inc"
	}
	int rv = 0;
	{
		guint64 max = 0L;
		double fmax = 0.0;
		size_t i, nbdata = np * C;
		int layer;
		struct simple_fit fit_, *fit = &fit_;
		for (int l = 0; l < C; l++) { fit_.fpdata[l] = (float *) out + np * l; fit_.pdata[l] = (WORD *) out + np * l; }
#define return rv =
#include "gen/sum_find_max.inc"
#undef return
		if (!rv) {
			nbdata = np;
#include "gen/sum_convert.inc"
		}
	}
	free(ss.fsum[0]); free(ss.sum[0]);
	return rv;
}

/* stackminmax.c:49-84 (initial values) + :160-182 */
int ref_stack_minmax(int is_float, const void *frames, int N, int C, int H, int W,
		const int *shiftx_, const int *shifty_, int ismax_, void *out) {
	struct simple_seq seq = { C, 0 };
	struct simple_args args_ = { &seq }, *args = &args_;
	struct simple_minmax mm, *mmdata = &mm;
	const size_t np = (size_t) H * W;
	memset(&mm, 0, sizeof mm);
	mm.input_32bits = is_float; mm.ismax = ismax_;
	mm.output_size[0] = W; mm.output_size[1] = H;
	for (int l = 0; l < C; l++) { mm.ffinal_pixel[l] = (float *) out + np * l; mm.final_pixel[l] = (WORD *) out + np * l; }
	if (is_float) { const float init = ismax_ ? 0.0f : 1.0f; for (size_t k = 0; k < np * C; k++) mm.ffinal_pixel[0][k] = init; }
	else { const WORD init = ismax_ ? 0 : USHRT_MAX; for (size_t k = 0; k < np * C; k++) mm.final_pixel[0][k] = init; }
	const int ismax = ismax_;
	for (int f = 0; f < N; f++) {
		struct simple_fit fit_, *fit = &fit_;
		fit_.rx = W; fit_.ry = H;
		for (int l = 0; l < C; l++) {
			fit_.fpdata[l] = is_float ? (float *) frames + ((size_t) f * C + l) * np : NULL;
			fit_.pdata[l] = is_float ? NULL : (WORD *) frames + ((size_t) f * C + l) * np;
		}
		const int shiftx = shiftx_ ? shiftx_[f] : 0, shifty = shifty_ ? shifty_[f] : 0;
#include "gen/minmax_pixel_loop.inc"
	}
	return 0;
}
