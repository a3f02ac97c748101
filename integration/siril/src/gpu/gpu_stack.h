/* gpu_stack.h — the two calls Siril's stacking code makes into src/gpu (see gpu_stack.c) */
#ifndef SRC_GPU_GPU_STACK_H
#define SRC_GPU_GPU_STACK_H
#include "core/siril.h"
#include "stacking/stacking.h"
int gpu_stack_mean_or_median(struct stacking_args *args, gboolean is_mean);
int gpu_do_normalization(struct stacking_args *args);
#endif
