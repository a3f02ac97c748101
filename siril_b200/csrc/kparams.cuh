// kparams.cuh — kernel parameter block, normalized sample load and result store shared by the
// stacking engines (stack_exact.cu, stack_sorted.cu).
#pragma once
#include "common.cuh"
#include "exact.cuh"

namespace sgpu {

struct StackParams {
	const void *frames;	// [N][C][H][W]
	int N, C, H, W;
	int normalize;
	const double *scale, *offset, *mul, *weights;	// device, [C*N]
	const int *shiftx;	// device [N] or null
	const float *mask, *drizz;	// feather masks [N][H][W] / drizzle weights [N][C][H][W] of the block, or null (mean stacking only)
	RejCfg rc;
	int is_mean, output_norm, out32, rejmaps;	// rejmaps: 0 none, 1 merged, 2 low+high
	int norm16;		// 8-bit frames, 16-bit output, output_norm: the result is scaled by 65535/255 (normalize_to16bit)
	void *out; u16 *rej_low, *rej_high;
	size_t out_plane;	// elements between the layers of out / rej_low / rej_high (H * W unless the caller's image is taller than the block)
	unsigned long long *irej;	// [3][2]
	void *scratch;		// exact engine: global columns when the tile does not fit in shared memory
	int use_aux;
	// pixels the sorted engine hands to the exact engine: entries (layer << 30) | pixel
	unsigned *list; unsigned *list_count;
	int ld;			// sorted engine: shared-memory row pitch per pixel, in floats
	int is_u16;		// sorted engine: sample type is a run-time switch there (fewer kernel variants)
	int use_tma, tile_rows, box_rows;	// sorted engine: TMA tile load (f32, no x shift, plane a multiple of 16 bytes); rows held / rows per TMA box
	int need_sums;		// sorted engine: phase A leaves sum(x-K), sum((x-K)^2) beside each sorted stack
	// selection engine, split form: the stack kernel ends after the extraction and leaves one record per pixel
	// (PSEL_REC_WORDS words: header + unsorted entries) for psel_finish_kernel; a launch covers pixels [pix_base, ...) of layer layer_base
	float *records; int split; int layer_base; size_t pix_base;
	const double *rcp_tab;	// selection engine: correctly rounded 1 / i, i <= 1280 (device)
	int dbg_stop;		// selection engine, SGPU_PSEL_STOP (timing breakdowns only, results are wrong): 1 return after the tile load, 2 after phase A, 3 after B1
};

// one sample, normalized (median_and_mean.c:1638-1686); x is the pixel's column
template <typename T>
__device__ __forceinline__ T load_norm(const StackParams &p, int f, int layer, size_t pix, int x) {
	T v = (T) 0;
	int sxv = p.shiftx ? p.shiftx[f] : 0;
	const int xs = x - sxv;
	if (xs >= 0 && xs < p.W) {
		const T *plane = (const T *) p.frames + ((size_t) f * p.C + layer) * ((size_t) p.H * p.W);
		v = __ldg(plane + pix - sxv);
	}
	const int ci = layer * p.N + f;
	switch (p.normalize) {
	default: case SGPU_NO_NORM: return v;
	case SGPU_ADDITIVE: case SGPU_ADDITIVE_SCALING:
		if (sizeof(T) == 4) {
			if (v != (T) 0) { const double tmp = (double) v * p.scale[ci]; return (T) (float) (tmp - p.offset[ci]); }
			return (T) 0;
		} else {
			if (v > (T) 0) { const double tmp = (double) v * p.scale[ci]; return (T) round_to_word(tmp - p.offset[ci]); }
			return (T) 0;
		}
	case SGPU_MULTIPLICATIVE: case SGPU_MULTIPLICATIVE_SCALING: {
		const double tmp = (double) v * p.scale[ci];
		if (sizeof(T) == 4) return (T) (float) (tmp * p.mul[ci]);
		return (T) round_to_word(tmp * p.mul[ci]);
	}
	}
}

// result conversion and rejection-map stores (median_and_mean.c:1701-1734)
template <typename T>
__device__ __forceinline__ void store_result(const StackParams &p, size_t o, double result, const int rej[2]) {
	if (p.rejmaps == 1) p.rej_low[o] = truncate_to_word(rej[0] + rej[1]);
	else if (p.rejmaps == 2) { p.rej_low[o] = truncate_to_word(rej[0]); p.rej_high[o] = truncate_to_word(rej[1]); }
	if (p.out32) {
		float v = (sizeof(T) == 4) ? (float) result : (float) result * .000015259022f;
		if (!p.output_norm) { v = v < 0.f ? 0.f : v; v = v > 1.f ? 1.f : v; }
		((float *) p.out)[o] = v;
	} else {
		if (p.norm16) result *= (65535.0 / 255.0);	// normalize_to16bit (median_and_mean.c:547-555)
		((u16 *) p.out)[o] = round_to_word(result);
	}
}

}  // namespace sgpu

void sgpu_fill_stack_params(sgpu::StackParams &p, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej);
// exact engine over the pixels listed in p.list (device-side count); asynchronous
int sgpu_launch_exact_list(sgpu_ctx *ctx, sgpu_block *blk, sgpu::StackParams p);
