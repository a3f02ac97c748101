// stats.cu — per-image statistics for normalization (K3).  Placeholder until the kernels land.
#include "common.cuh"
extern "C" int sgpu_image_stats(sgpu_ctx *ctx, const void *plane, int on_device, int dtype, long nx, long ny,
		int lite, sgpu_imstats *out) {
	(void) plane; (void) on_device; (void) dtype; (void) nx; (void) ny; (void) lite; (void) out;
	return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "sgpu_image_stats: not built yet");
}
