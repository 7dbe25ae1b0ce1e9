// TEST INFRASTRUCTURE (oracle side) -- never included by the product.
//
// Stand-in for the slice of GSL 2.7 <gsl/gsl_linalg.h> that nuSIprop.hpp:212-215,
// 296-321 uses: 3x3 gsl_matrix / gsl_vector / gsl_permutation and
// gsl_linalg_LU_decomp / gsl_linalg_LU_solve (Gaussian elimination with partial
// pivoting, unit-lower L stored in place, as GSL documents it).
#ifndef REFSHIM_GSL_LINALG_H
#define REFSHIM_GSL_LINALG_H
#include <cmath>
#include <cstddef>

typedef struct { size_t size1, size2; double *data; } gsl_matrix;
typedef struct { size_t size; double *data; } gsl_vector;
typedef struct { size_t size; size_t *data; } gsl_permutation;

inline gsl_matrix *gsl_matrix_alloc(size_t n1, size_t n2) {
  gsl_matrix *m = new gsl_matrix;
  m->size1 = n1; m->size2 = n2; m->data = new double[n1 * n2]();
  return m;
}
inline void gsl_matrix_free(gsl_matrix *m) { delete[] m->data; delete m; }
inline void gsl_matrix_set(gsl_matrix *m, size_t i, size_t j, double x) { m->data[i * m->size2 + j] = x; }
inline double gsl_matrix_get(const gsl_matrix *m, size_t i, size_t j) { return m->data[i * m->size2 + j]; }
inline gsl_vector *gsl_vector_alloc(size_t n) {
  gsl_vector *v = new gsl_vector;
  v->size = n; v->data = new double[n]();
  return v;
}
inline void gsl_vector_free(gsl_vector *v) { delete[] v->data; delete v; }
inline void gsl_vector_set(gsl_vector *v, size_t i, double x) { v->data[i] = x; }
inline double gsl_vector_get(const gsl_vector *v, size_t i) { return v->data[i]; }
inline gsl_permutation *gsl_permutation_alloc(size_t n) {
  gsl_permutation *p = new gsl_permutation;
  p->size = n; p->data = new size_t[n];
  for (size_t i = 0; i < n; ++i) p->data[i] = i;
  return p;
}
inline void gsl_permutation_free(gsl_permutation *p) { delete[] p->data; delete p; }

inline int gsl_linalg_LU_decomp(gsl_matrix *A, gsl_permutation *p, int *signum) {
  const size_t N = A->size1;
  *signum = 1;
  for (size_t i = 0; i < N; ++i) p->data[i] = i;
  for (size_t j = 0; j + 1 < N; ++j) {
    double amax = std::fabs(gsl_matrix_get(A, j, j));
    size_t ipiv = j;
    for (size_t i = j + 1; i < N; ++i) {
      const double aij = std::fabs(gsl_matrix_get(A, i, j));
      if (aij > amax) { amax = aij; ipiv = i; }
    }
    if (ipiv != j) {
      for (size_t k = 0; k < N; ++k) {
        const double t = gsl_matrix_get(A, j, k);
        gsl_matrix_set(A, j, k, gsl_matrix_get(A, ipiv, k));
        gsl_matrix_set(A, ipiv, k, t);
      }
      const size_t t = p->data[j]; p->data[j] = p->data[ipiv]; p->data[ipiv] = t;
      *signum = -*signum;
    }
    const double ajj = gsl_matrix_get(A, j, j);
    if (ajj != 0.0) {
      for (size_t i = j + 1; i < N; ++i) {
        const double aij = gsl_matrix_get(A, i, j) / ajj;
        gsl_matrix_set(A, i, j, aij);
        for (size_t k = j + 1; k < N; ++k)
          gsl_matrix_set(A, i, k, gsl_matrix_get(A, i, k) - aij * gsl_matrix_get(A, j, k));
      }
    }
  }
  return 0;
}

inline int gsl_linalg_LU_solve(const gsl_matrix *LU, const gsl_permutation *p, const gsl_vector *b,
                               gsl_vector *x) {
  const size_t N = LU->size1;
  for (size_t i = 0; i < N; ++i) x->data[i] = b->data[p->data[i]];
  for (size_t i = 1; i < N; ++i) {  // L y = Pb (unit lower)
    double s = x->data[i];
    for (size_t k = 0; k < i; ++k) s -= gsl_matrix_get(LU, i, k) * x->data[k];
    x->data[i] = s;
  }
  for (size_t ii = N; ii-- > 0;) {  // U x = y
    double s = x->data[ii];
    for (size_t k = ii + 1; k < N; ++k) s -= gsl_matrix_get(LU, ii, k) * x->data[k];
    x->data[ii] = s / gsl_matrix_get(LU, ii, ii);
  }
  return 0;
}
#endif
