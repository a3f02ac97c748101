// common.cuh — internal types shared by the translation units of libsgpu.so
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/sgpu.h"

struct sgpu_ctx {
	int device = 0;
	cudaStream_t own_stream = nullptr;
	cudaStream_t stream = nullptr;
	cudaStream_t copy_stream = nullptr;
	std::string err;
	uint64_t launches = 0;
	int sm_count = 148;
	size_t smem_optin = 0;
	bool stats_full_hist = false;			// SGPU_STATS_FULL_HIST=1: skip the windowed histograms (stats.cu)
	uint64_t stats_window_misses = 0;		// medians that fell outside the window and took the full histogram
	int (*should_cancel)(void *) = nullptr;
	void *cancel_user = nullptr;
	// small device-side parameter arena (coefficients, weights, shifts, critical values)
	void *d_params = nullptr; size_t d_params_cap = 0;
	void *h_params = nullptr;			// pinned mirror
	void *d_scratch = nullptr; size_t d_scratch_cap = 0;	// global-column fallback, stats temporaries
	void *d_arena = nullptr; size_t d_arena_cap = 0;	// sgpu_stack_host / sgpu_stack_fits: row-block buffers and outputs, kept between calls
	void *d_list = nullptr; size_t d_list_cap = 0; bool list_used = false; int last_engine = 0;		// sorted engine: pixels handed to the exact engine
	// raw FITS reader (fitsread.cu): two pinned + two device staging chunks, recycled by events
	void *h_stage[2] = { nullptr, nullptr }, *d_stage[2] = { nullptr, nullptr };
	cudaEvent_t stage_ev[2] = { nullptr, nullptr }; int stage_idx = 0;
};

struct sgpu_block {
	sgpu_ctx *ctx = nullptr;
	sgpu_block_desc d{};
	void *dev = nullptr;
	bool owned = false;
	// optional per-sample weight planes (data->mask[frame] / data->drizz[frame], median_and_mean.c:483-542)
	float *mask = nullptr, *drizz = nullptr;	// [N][H][W] / [N][C][H][W]
	bool mask_owned = false, drizz_owned = false;
	size_t elem() const { return d.dtype == SGPU_F32 ? 4 : 2; }
	size_t plane() const { return (size_t) d.width * d.height; }
	size_t frame_elems() const { return plane() * d.nb_layers; }
	size_t bytes() const { return frame_elems() * d.nb_frames * elem(); }
};

int sgpu_fail(sgpu_ctx *ctx, int code, const char *fmt, ...);
int sgpu_cuda_fail(sgpu_ctx *ctx, cudaError_t e, const char *what);
int sgpu_ensure_scratch(sgpu_ctx *ctx, size_t bytes);
int sgpu_ensure_list(sgpu_ctx *ctx, size_t bytes);
void sgpu_free_stage(sgpu_ctx *ctx);
const double *sgpu_rcp_table(sgpu_ctx *ctx);	// device table of correctly rounded 1.0 / i, i < SGPU_RCP_N (stack_psel.cu); null on failure
int sgpu_launch_mask_ramp(sgpu_ctx *ctx, float *d_mask, size_t n, int feather_dist);

#define SGPU_CUDA(ctx, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return sgpu_cuda_fail((ctx), e__, #call); } while (0)

// device-resident copy of the per-call parameter arrays
struct StackDevParams {
	const double *scale = nullptr, *offset = nullptr, *mul = nullptr, *weights = nullptr;
	const int *shiftx = nullptr;
	const float *crit = nullptr;
};

// engines
bool sgpu_psel_supports(const sgpu_block_desc &d, const sgpu_stack_args *a);
int sgpu_launch_psel(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej);
bool sgpu_lsort_supports(const sgpu_block_desc &d, const sgpu_stack_args *a);
int sgpu_launch_lsort(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej);
bool sgpu_sorted_supports(const sgpu_block_desc &d, const sgpu_stack_args *a);
int sgpu_launch_sorted(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej);
int sgpu_launch_exact(sgpu_ctx *ctx, sgpu_block *blk, const sgpu_stack_args *a, const StackDevParams &dp,
		void *d_out, uint16_t *d_rej_low, uint16_t *d_rej_high, uint64_t *d_irej);
