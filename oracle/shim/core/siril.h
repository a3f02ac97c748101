/* TEST INFRASTRUCTURE ONLY — shim for building the *reference's own* hot-path
 * translation units (oracle/_ref) without glib/gsl/cfitsio.  It supplies only
 * type names; every enum/struct whose *content* matters (rejection,
 * normalization, threading_type, imstats, the rounding inlines) is extracted
 * from /root/reference at build time into oracle/_ref/gen/ by build_ref.py and
 * included below — nothing from the reference is committed here.
 * Stands in for: src/core/siril.h (reference), which pulls in glib/gsl/fitsio. */
#ifndef SHIM_SIRIL_H
#define SHIM_SIRIL_H
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <limits.h>
#include <float.h>
#include <math.h>
#include <sys/time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* glib scalar names used by the compiled TUs */
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
typedef struct _shim_opaque_GDateTime GDateTime;
typedef struct _shim_opaque_GSList GSList;
typedef struct _shim_opaque_GList GList;
typedef struct _shim_opaque_gslh gsl_histogram;
#ifndef TRUE
#define TRUE 1
#define FALSE 0
#endif
#define _(s) (s)
typedef void (*destructor)(void *);

typedef unsigned short WORD;
typedef unsigned char BYTE;

#define siril_log_debug(...) do { } while (0)
#define siril_log_message(...) do { } while (0)
#define siril_log_error(...) do { } while (0)
#define MEMORY_ALLOCATION 113

/* reference macros/consts/enums/structs, extracted verbatim at build time */
#include "gen/siril_h_macros.inc"     /* PRINT_ALLOC_ERR, min/max, SQR        (core/siril.h) */
#include "gen/siril_h_consts.inc"     /* USHRT_MAX_DOUBLE ... BYTES_IN_A_MB   (core/siril.h) */
#include "gen/settings_enums.inc"     /* rectangle, normalization, rejection  (core/settings.h) */
#include "gen/siril_h_threading.inc"  /* threading_type                       (core/siril.h) */
#include "gen/siril_h_imstats.inc"    /* imstats, overlap_stats_t             (core/siril.h) */

typedef enum { DATA_USHORT, DATA_FLOAT, DATA_UNSUPPORTED } data_type;

/* opaque stand-ins: the compiled code only passes these around */
/* regdata: only the two members compute_wfwhm_weights / compute_nbstars_weights read (core/siril.h:438,442) */
/* and the homography compute_overlap reads through translation_from_H (core/siril.h:426-436) */
typedef struct { double h00, h01, h02, h10, h11, h12, h20, h21, h22; int pair_matched, Inliers; } Homography;
typedef struct shim_regdata { float weighted_fwhm; int number_of_stars; Homography H; } regdata;
typedef struct shim_imgdata { int rx, ry; } imgdata;
typedef struct shim_sequence { int nb_layers; int number; imstats ***stats; regdata **regparam;
	unsigned int rx, ry; gboolean is_variable; imgdata *imgparam; overlap_stats_t **ostats; } sequence;
typedef struct shim_fits { long naxes[3]; float *fdata; data_type type; float *fpdata[3]; WORD *pdata[3]; } fits;

void g_free(void *p);
int g_strcmp0(const char *a, const char *b);
char *g_utf8_collate_key_for_filename(const char *s, long len);

#ifdef __cplusplus
}
#endif
#endif
