// TEST INFRASTRUCTURE (oracle side) -- never included by the product.
//
// Stand-in for GSL 2.7 <gsl/gsl_poly.h> (GSL is not vendored in /root/reference
// and not installed).  Only what aux.hpp:34-36 (nuSIaux::getmL) calls:
// gsl_poly_complex_workspace_{alloc,free} and gsl_poly_complex_solve.  GSL
// finds the roots by QR on the balanced companion matrix; root ORDER is
// irrelevant to getmL (it accepts the unique admissible root), so we use
// Durand-Kerner in long double followed by Newton polishing.
#ifndef REFSHIM_GSL_POLY_H
#define REFSHIM_GSL_POLY_H
#include <complex>
#include <cstddef>

typedef struct { size_t nc; } gsl_poly_complex_workspace;
typedef double *gsl_complex_packed_ptr;

inline gsl_poly_complex_workspace *gsl_poly_complex_workspace_alloc(size_t n) {
  gsl_poly_complex_workspace *w = new gsl_poly_complex_workspace;
  w->nc = n;
  return w;
}
inline void gsl_poly_complex_workspace_free(gsl_poly_complex_workspace *w) { delete w; }

inline int gsl_poly_complex_solve(const double *a, size_t n, gsl_poly_complex_workspace *,
                                  gsl_complex_packed_ptr z) {
  typedef std::complex<long double> C;
  const int deg = (int)n - 1;
  C r[16];
  long double scale = 0;
  for (int i = 0; i < deg; ++i) {
    long double v = std::pow(std::fabs((long double)a[i] / (long double)a[deg]), 1.0L / (deg - i));
    if (v > scale) scale = v;
  }
  if (scale == 0) scale = 1;
  for (int i = 0; i < deg; ++i) r[i] = scale * std::pow(C(0.4L, 0.9L), i);
  for (int it = 0; it < 2000; ++it) {
    long double mv = 0;
    for (int i = 0; i < deg; ++i) {
      C p = (long double)a[deg];
      for (int k = deg - 1; k >= 0; --k) p = p * r[i] + (long double)a[k];
      C d = (long double)a[deg];
      for (int j = 0; j < deg; ++j)
        if (j != i) d *= (r[i] - r[j]);
      C step = p / d;
      r[i] -= step;
      long double m = std::abs(step);
      if (m > mv) mv = m;
    }
    if (mv < 1e-30L * scale) break;
  }
  for (int i = 0; i < deg; ++i) {
    z[2 * i] = (double)r[i].real();
    z[2 * i + 1] = (double)r[i].imag();
  }
  return 0;
}
#endif
