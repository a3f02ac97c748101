/* TEST HARNESS ONLY — src/core/processing.h */
#ifndef HARNESS_PROCESSING_H
#define HARNESS_PROCESSING_H
#include "core/siril.h"
typedef int (*seq_image_filter)(sequence *seq, int nb_img, double any);
struct seq_filter_config { int dummy; };
gboolean processing_should_continue(void);	/* core/processing.h:270 */
#endif
