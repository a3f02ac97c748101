/* TEST HARNESS ONLY — stands in for Siril's src/core/siril.h so that the glue
 * (integration/siril/src/gpu/gpu_stack.c) and the reference's own src/stacking/stacking.h compile
 * in this image (no glib / gsl / cfitsio).  Scalar typedefs as oracle/shim; the enums and structs
 * whose content matters are the reference's, extracted at build time into oracle/_ref/gen; `fits`,
 * `sequence`, `regdata` carry the members the glue touches, under the reference's names
 * (core/siril.h:420-500, 590-700). */
#ifndef HARNESS_SIRIL_H
#define HARNESS_SIRIL_H
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <limits.h>
#include <float.h>
#include <math.h>
#include <sys/time.h>

typedef int gboolean;
typedef int gint;
typedef unsigned int guint;
typedef char gchar;
typedef void *gpointer;
typedef const void *gconstpointer;
typedef int64_t gint64;
typedef uint64_t guint64;
typedef uint32_t guint32;
typedef double gdouble;
typedef struct _h_GDateTime GDateTime;
typedef struct _h_GSList GSList;
typedef struct _h_GList GList;
typedef struct _h_gslh gsl_histogram;
#ifndef TRUE
#define TRUE 1
#define FALSE 0
#endif
#define _(s) (s)
typedef void (*destructor)(void *);
typedef unsigned short WORD;
typedef unsigned char BYTE;
#define MEMORY_ALLOCATION 113
#define BYTE_IMG 8
#define SHORT_IMG 16
#define USHORT_IMG 20
#define FLOAT_IMG -32

#include "gen/siril_h_macros.inc"
#include "gen/siril_h_consts.inc"
#include "gen/settings_enums.inc"
#include "gen/siril_h_threading.inc"
#include "gen/siril_h_imstats.inc"

typedef enum { DATA_USHORT, DATA_FLOAT, DATA_UNSUPPORTED } data_type;

typedef struct { double h00, h01, h02, h10, h11, h12, h20, h21, h22; int pair_matched, Inliers; } Homography;
typedef struct registration_data { float roundness, fwhm, weighted_fwhm, background_lvl; int number_of_stars; Homography H; } regdata;
typedef struct imdata { int filenum; gboolean incl; int rx, ry; } imgdata;

typedef struct ffit {
	unsigned int rx, ry;
	int bitpix, orig_bitpix;
	long naxes[3];
	int naxis;
	WORD *data; WORD *pdata[3];
	float *fdata; float *fpdata[3];
	data_type type;
} fits;

typedef struct sequ {
	int number, selnum, nb_layers;
	unsigned int rx, ry;
	gboolean is_variable;
	int bitpix, reference_image;
	imgdata *imgparam;
	regdata **regparam;
	imstats ***stats;
	void *harness;		/* the harness's frame store */
} sequence;

struct stacking_args;	/* core/siril.h forward-declares it for stack_method (stacking/stacking.h:16) */

struct gui_iface_t { void (*set_progress)(double fraction, const char *text); };
extern struct gui_iface_t gui_iface;
#endif
