/* TEST INFRASTRUCTURE ONLY — shim for src/core/processing.h (threading helpers
 * only; the oracle runs the per-pixel code single-threaded inside each call). */
#ifndef SHIM_PROCESSING_H
#define SHIM_PROCESSING_H
#include "core/siril.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef int (*seq_image_filter)(sequence *seq, int nb_img, double any);
struct seq_filter_config { int dummy; };
static inline int check_threading(threading_type *threads) { (void)threads; return 1; }
static inline int limit_threading(threading_type *threads, int min_it, size_t total) {
	(void)threads; (void)min_it; (void)total; return 1;
}
#ifdef __cplusplus
}
#endif
#endif
