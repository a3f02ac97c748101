/* TEST INFRASTRUCTURE ONLY — the one TU of the pinned-order reference build that is compiled
 * with -fopenmp: the `omp parallel for schedule(dynamic)` over row blocks of
 * median_and_mean.c:1552, kept apart so that the reference's per-stack `omp simd` reductions in
 * the other TUs stay in the sequential order (see build_ref.py). */
#ifdef _OPENMP
#include <omp.h>
#endif
void ref_parallel_for(int n, int nthreads, void (*fn)(int, int, void *), void *ctx) {
#pragma omp parallel for num_threads(nthreads) schedule(dynamic)
	for (int i = 0; i < n; i++) {
		int tid = 0;
#ifdef _OPENMP
		tid = omp_get_thread_num();
#endif
		fn(i, tid, ctx);
	}
}
