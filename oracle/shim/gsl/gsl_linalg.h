/* TEST INFRASTRUCTURE ONLY — stand-in for the six GSL calls solve_overlap_coeffs makes (normalization.c:340-346).
 * GSL is an un-vendored system dependency of the reference and absent here: the LU solve is RESTATED
 * (Gaussian elimination with partial pivoting, what gsl_linalg_LU_decomp documents), so this one step of
 * the overlap normalization is "parity unpinned"; tests compare it with LAPACK (numpy.linalg.solve) as well. */
#ifndef SHIM_GSL_LINALG_H
#define SHIM_GSL_LINALG_H
#include <stdlib.h>
#include <math.h>
typedef struct { double *data; size_t size1, size2; } gsl_matrix;
typedef struct { double *data; size_t size; } gsl_vector;
typedef struct { gsl_matrix matrix; } gsl_matrix_view;
typedef struct { gsl_vector vector; } gsl_vector_view;
typedef struct { size_t *data; size_t size; } gsl_permutation;
static inline gsl_matrix_view gsl_matrix_view_array(double *a, size_t n1, size_t n2) { gsl_matrix_view v = { { a, n1, n2 } }; return v; }
static inline gsl_vector_view gsl_vector_view_array(double *b, size_t n) { gsl_vector_view v = { { b, n } }; return v; }
static inline gsl_vector *gsl_vector_alloc(size_t n) { gsl_vector *v = malloc(sizeof *v); v->data = calloc(n ? n : 1, sizeof(double)); v->size = n; return v; }
static inline void gsl_vector_free(gsl_vector *v) { if (v) { free(v->data); free(v); } }
static inline gsl_permutation *gsl_permutation_alloc(size_t n) { gsl_permutation *p = malloc(sizeof *p); p->data = malloc((n ? n : 1) * sizeof(size_t)); p->size = n; for (size_t i = 0; i < n; i++) p->data[i] = i; return p; }
static inline void gsl_permutation_free(gsl_permutation *p) { if (p) { free(p->data); free(p); } }
static inline int gsl_linalg_LU_decomp(gsl_matrix *A, gsl_permutation *p, int *signum) {
	const size_t N = A->size1; double *a = A->data;
	*signum = 1;
	for (size_t j = 0; j < N; j++) {
		size_t piv = j; double best = fabs(a[j * N + j]);
		for (size_t i = j + 1; i < N; i++) if (fabs(a[i * N + j]) > best) { best = fabs(a[i * N + j]); piv = i; }
		if (piv != j) {
			for (size_t k = 0; k < N; k++) { double t = a[j * N + k]; a[j * N + k] = a[piv * N + k]; a[piv * N + k] = t; }
			size_t t = p->data[j]; p->data[j] = p->data[piv]; p->data[piv] = t; *signum = -*signum;
		}
		const double ajj = a[j * N + j];
		if (ajj != 0.0)
			for (size_t i = j + 1; i < N; i++) {
				const double l = a[i * N + j] / ajj;
				a[i * N + j] = l;
				for (size_t k = j + 1; k < N; k++) a[i * N + k] -= l * a[j * N + k];
			}
	}
	return 0;
}
static inline int gsl_linalg_LU_solve(const gsl_matrix *LU, const gsl_permutation *p, const gsl_vector *b, gsl_vector *x) {
	const size_t N = LU->size1; const double *a = LU->data;
	for (size_t i = 0; i < N; i++) { double v = b->data[p->data[i]]; for (size_t k = 0; k < i; k++) v -= a[i * N + k] * x->data[k]; x->data[i] = v; }
	for (size_t i = N; i-- > 0;) { double v = x->data[i]; for (size_t k = i + 1; k < N; k++) v -= a[i * N + k] * x->data[k]; x->data[i] = v / a[i * N + i]; }
	return 0;
}
#endif
