/* TEST HARNESS ONLY — src/algos/statistics.h:55 */
#ifndef HARNESS_STATISTICS_H
#define HARNESS_STATISTICS_H
#include "core/siril.h"
void add_stats_to_seq(sequence *seq, int image_index, int layer, imstats *stat);
#endif
