/* TEST INFRASTRUCTURE ONLY — see gsl_statistics_float.h; call site median_and_mean.c:914 */
#ifndef SHIM_GSL_STATS_USHORT_H
#define SHIM_GSL_STATS_USHORT_H
#include <stddef.h>
static inline double gsl_stats_ushort_median_from_sorted_data(const unsigned short d[], size_t stride, size_t n) {
	if (n == 0) return 0.0;
	size_t lhs = (n - 1) / 2, rhs = n / 2;
	if (lhs == rhs) return d[lhs * stride];
	return (d[lhs * stride] + d[rhs * stride]) / 2.0;
}
#endif
