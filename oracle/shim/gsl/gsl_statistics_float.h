/* TEST INFRASTRUCTURE ONLY — GSL is an un-vendored, unpinned system dependency
 * of the reference (meson.build:284).  The one routine the path calls
 * (rejection_float.c:312) is restated from GSL's published definition:
 * median of sorted data = mean of the two middle elements (lhs=(n-1)/2, rhs=n/2). */
#ifndef SHIM_GSL_STATS_FLOAT_H
#define SHIM_GSL_STATS_FLOAT_H
#include <stddef.h>
static inline double gsl_stats_float_median_from_sorted_data(const float d[], size_t stride, size_t n) {
	if (n == 0) return 0.0;
	size_t lhs = (n - 1) / 2, rhs = n / 2;
	if (lhs == rhs) return d[lhs * stride];
	return (d[lhs * stride] + d[rhs * stride]) / 2.0;
}
#endif
