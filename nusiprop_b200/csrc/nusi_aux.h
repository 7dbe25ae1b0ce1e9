// nusi_aux.h -- per-point constants and the "stable difference" helpers of the cross-section
// builder.  Restates nuSIaux:: of the reference (aux.hpp:53-166) for host+device, with the
// same regime thresholds so that the same branch is taken for the same input.
#ifndef NUSI_AUX_H
#define NUSI_AUX_H
#include "nusi_math.h"

namespace nusi {

// 3-point Gauss-Legendre rule (aux.hpp:53-54).  Values are the doubles the reference's
// static initialisers produce: 5./9., 8./9., -sqrt(3./5.), 0, sqrt(3./5.).
#define NUSI_GL_W0 (5. / 9.)
#define NUSI_GL_W1 (8. / 9.)
#define NUSI_GL_X0 0.7745966692414834  // sqrt(3./5.) rounded to nearest (= glibc sqrt)
NUSI_HD double gl_w(int i) { return i == 1 ? NUSI_GL_W1 : NUSI_GL_W0; }
NUSI_HD double gl_x(int i) { return i == 0 ? -NUSI_GL_X0 : (i == 1 ? 0.0 : NUSI_GL_X0); }

// Everything about one (mphi, g, mntot) parameter point that the cross sections need.
// Built on the host (nusi_host.cpp) so that scalar_width (nuSIprop.hpp:748-757), the masses
// (nuSIprop.hpp:184-203) and |U_{flav,k}|^2 (nuSIprop.hpp:130-163) are the same doubles the
// reference uses.
struct PointParams {
  double mphi, g;
  double Ga;      // scalar decay width
  double mn[3];   // neutrino masses m1,m2,m3 [eV]
  double U2f[3];  // |U_{flav,k}|^2
  int majorana, non_resonant, phiphi;
};

// atan(x) - atan(y), series when both are large and of the same sign (aux.hpp:63-75)
NUSI_HD double atandiff(double x, double y) {
  if (fabs(x) < 1e2 || fabs(y) < 1e2 || x * y < 0) return natan(x) - natan(y);
  return -1. / x + cub(1. / x) / 3. - (-1. / y + cub(1. / y) / 3.);
}

// Li2(-x) - Li2(-y)  (aux.hpp:98-113)
NUSI_HD double dilogdiff_large(double x) {
  return -sqr(nlog(x)) / 2. + 1. / x - sqr(1. / x) / 4. + cub(1. / x) / 9. - sqr(sqr(1. / x)) / 16;
}
NUSI_HD double dilogdiff_small(double x) { return -x + sqr(x) / 4. - cub(x) / 9. + sqr(sqr(x)) / 16.; }
NUSI_HD double dilogdiff(double x, double y) {
  if (x > 1e2 && y > 1e2) return dilogdiff_large(x) - dilogdiff_large(y);
  if (x < 1e-2 && y < 1e-2) return dilogdiff_small(x) - dilogdiff_small(y);
  return li2_real(-x) - li2_real(-y);
}

// Li2(-1-x) - Li2(-1-y)  (aux.hpp:115-130)
NUSI_HD double dilog1m_large(double x) {
  return -sqr(nlog(x)) / 2. + (1 - nlog(x)) / x + (-7 + 2 * nlog(x)) / (4. * sqr(x)) +
         (19 - 3 * nlog(x)) / (9. * cub(x)) + (-125 + 12 * nlog(x)) / (48. * sqr(sqr(x)));
}
NUSI_HD double dilog1m_small(double x) {
  return -x * nlog(2.) + (sqr(x) * (-1 + nlog(4.))) / 4. + (cub(x) * (5 - 8 * nlog(2.))) / 24. +
         sqr(sqr(x)) * (-1. / 6. + nlog(2.) / 4.);
}
NUSI_HD double dilog1mdiff(double x, double y) {
  if (x > 1e2 && y > 1e2) return dilog1m_large(x) - dilog1m_large(y);
  if (x < 1e-2 && y < 1e-2) return dilog1m_small(x) - dilog1m_small(y);
  return li2_real(-1 - x) - li2_real(-1 - y);
}

// Li2(1+x) - Li2(1+y), x,y < 0  (aux.hpp:132-148)
NUSI_HD double dilog1p_large(double x) {
  return (-1 - 3 * nlog(-x)) / (9. * cub(x)) + (-1 - nlog(-x)) / x - sqr(nlog(-x)) / 2. +
         (1 + 2 * nlog(-x)) / (4. * sqr(x)) + (1 + 4 * nlog(-x)) / (16. * sqr(sqr(x)));
}
NUSI_HD double dilog1p_small(double x) {
  return x * (1 - nlog(-x)) + (sqr(x) * (-1 + 2 * nlog(-x))) / 4. + (cub(x) * (1 - 3 * nlog(-x))) / 9. +
         (sqr(sqr(x)) * (-1 + 4 * nlog(-x))) / 16.;
}
NUSI_HD double dilog1pdiff(double x, double y) {
  if (-x > 1e2 && -y > 1e2) return dilog1p_large(x) - dilog1p_large(y);
  if (-x < 1e-2 && -y < 1e-2) return dilog1p_small(x) - dilog1p_small(y);
  return li2_real(1 + x) - li2_real(1 + y);
}

// Li2(1/(1-x)) - Li2(1/(1-y)), x,y < 0  (aux.hpp:150-166)
NUSI_HD double dilog1o1m_large(double x) {
  return -25 / (48. * sqr(sqr(x))) - 11 / (18. * cub(x)) - 3 / (4. * sqr(x)) - 1 / x;
}
NUSI_HD double dilog1o1m_small(double x) {
  return (sqr(sqr(x)) * (-19 - 12 * nlog(-x))) / 48. + (cub(x) * (-7 - 6 * nlog(-x))) / 18. +
         (sqr(x) * (-1 - 2 * nlog(-x))) / 4. + x * (1 - nlog(-x));
}
NUSI_HD double dilog1over1mdiff(double x, double y) {
  if (-x > 1e2 && -y > 1e2) return dilog1o1m_large(x) - dilog1o1m_large(y);
  if (-x < 1e-2 && -y < 1e-2) return dilog1o1m_small(x) - dilog1o1m_small(y);
  return li2_real(1 / (1 - x)) - li2_real(1 / (1 - y));
}

// Li2(x) - Li2(y) for complex x, y  (aux.hpp:77-96).  Large-|z| branch: four inverse powers
// plus the analytic continuation term -(i/2)(-sgn(Im z) 2 pi ln z - i ln^2 z).
NUSI_HD cplx dilog_cplx_large(cplx z) {
  const double sg = (z.im >= 0) ? 1.0 : -1.0;
  const cplx z2 = z * z;
  const cplx L = clog_(z);
  const cplx L2 = L * L;
  // inner = (-sg*2*pi)*L - i*L^2 ;  cont = (i/2)*inner
  const double c = -sg * 2 * kPi;
  const cplx inner = mk(c * L.re + L2.im, c * L.im - L2.re);
  const cplx cont = mk(-0.5 * inner.im, 0.5 * inner.re);
  return -(1.0 / (16. * (z2 * z2))) - 1.0 / (9. * (z2 * z)) - 1.0 / (4. * z2) - 1.0 / z - cont;
}
NUSI_HD_NOINLINE cplx dilogdiff_complex(cplx x, cplx y) {
  if (cabs_(x) > 1e2 && cabs_(y) > 1e2) return dilog_cplx_large(x) - dilog_cplx_large(y);
  return cli2(x.re, x.im) - cli2(y.re, y.im);
}
NUSI_HD cplx real_c(double x) { return mk(x, 0.0); }

}  // namespace nusi
#endif
