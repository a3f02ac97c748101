/* TEST HARNESS ONLY — src/io/sequence.h:22-33 */
#ifndef HARNESS_SEQUENCE_H
#define HARNESS_SEQUENCE_H
#include "core/siril.h"
int seq_read_frame(sequence *seq, int index, fits *dest, gboolean force_float, int thread_id);
int seq_open_image(sequence *seq, int index);
void seq_close_image(sequence *seq, int index);
int seq_opened_read_region(sequence *seq, int layer, int index, void *buffer, const rectangle *area, int thread_id);
#endif
