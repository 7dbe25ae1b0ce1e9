// TEST INFRASTRUCTURE (oracle side) -- never included by the product.
//
// Stand-in for GSL 2.7 <gsl/gsl_sf_dilog.h> (specfunc/dilog.c).  GSL is a
// dynamically linked third-party dependency of the reference (include sites
// nuSIprop.hpp:7-8, aux.hpp:1-2) that is not under /root/reference and not
// installed in this image.  This file restates GSL's published algorithm:
//
//   gsl_sf_dilog(x)                   real part of Li2(x) for any real x
//     x >= 0 : regions (2,inf) (1.01,2] (1,1.01] {1} (0.5,1) (0.25,0.5] (0,0.25]
//              using the plain series  sum x^k/k^2  and the accelerated series
//              Li2(x) = 1 + (1-x)ln(1-x)/x + sum x^k/(k^2 (k+1))
//     x <  0 : Li2(x) = -Li2(-x) + Li2(x^2)/2
//   gsl_sf_complex_dilog_xy_e(x,y)    complex Li2(x + i y)
//     y == 0 exactly : Re = gsl_sf_dilog(x), Im = (x >= 1) ? -pi ln x : 0
//     |z| > 1        : inversion  Li2(z) = -Li2(1/z) - zeta(2) - ln^2(-z)/2
//     Re z > 0.732   : reflection Li2(z) = -Li2(1-z) + zeta(2) - ln z ln(1-z)
//     fundamental region: plain series (r <= 0.25), accelerated series (r <= 0.98)
//   DEVIATION: for 0.98 < r <= 1 inside the fundamental region GSL sums a
//   7-term expansion in ln r around the unit circle built on its Clausen
//   Chebyshev table, and on |z| = 1 it calls gsl_sf_clausen; that table is not
//   reproducible offline, so both cases use the Bernoulli-number series in
//   u = -ln(1-z) instead (same function value to < 2e-16, checked against
//   mpmath in tests/test_oracle_shim.py).
//
// Call sites in the reference: gsl_sf_dilog nuSIprop.hpp:1098,1202,1375-1384,1398,
// aux.hpp:112,129,147,165; gsl_sf_complex_dilog_xy_e nuSIprop.hpp:1444-1451,
// aux.hpp:92-93.  None of the reference's shipped outputs depends on these
// (output/data_massless.txt is resonant-only): parity of this shim versus real
// GSL is pinned only through mpmath.
#ifndef REFSHIM_GSL_SF_DILOG_H
#define REFSHIM_GSL_SF_DILOG_H
#include <cmath>
#include <complex>
#include <cfloat>

typedef struct { double val; double err; } gsl_sf_result;

namespace refshim_gsl {
static const double kEps = DBL_EPSILON;
static const double kPi = 3.14159265358979323846264338328;

inline double series_1(double x) {  // sum_{k>=1} x^k/k^2, |x| small
  double sum = x, term = x;
  for (int k = 2; k < 1000; ++k) {
    const double rk = (k - 1.0) / k;
    term *= x;
    term *= rk * rk;
    sum += term;
    if (std::fabs(term / sum) < kEps) break;
  }
  return sum;
}
inline double series_2_tail(double r) {  // sum_{k>=1} r^k/(k^2 (k+1))
  double rk = r, sum = 0.5 * r;
  int k = 2;
  for (; k < 10; ++k) { rk *= r; sum += rk / (k * k * (k + 1.0)); }
  for (; k < 100; ++k) {
    rk *= r;
    const double ds = rk / (k * k * (k + 1.0));
    sum += ds;
    if (std::fabs(ds / sum) < 0.5 * kEps) break;
  }
  return sum;
}
inline double series_2(double x) {  // accelerated series, -1 < x < 1
  double val = series_2_tail(x);
  double t;
  if (x > 0.01) {
    t = (1.0 - x) * std::log(1.0 - x) / x;
  } else {
    const double c3 = 1.0 / 3.0, c4 = 1.0 / 4.0, c5 = 1.0 / 5.0, c6 = 1.0 / 6.0, c7 = 1.0 / 7.0, c8 = 1.0 / 8.0;
    const double t68 = c6 + x * (c7 + x * c8);
    const double t38 = c3 + x * (c4 + x * (c5 + x * t68));
    t = (x - 1.0) * (1.0 + x * (0.5 + x * t38));
  }
  val += 1.0 + t;
  return val;
}
inline double dilog_xge0(double x) {
  if (x > 2.0) {
    const double ser = series_2(1.0 / x);
    const double lx = std::log(x);
    return kPi * kPi / 3.0 - ser - 0.5 * lx * lx;
  } else if (x > 1.01) {
    const double ser = series_2(1.0 - 1.0 / x);
    const double lx = std::log(x);
    const double lt = lx * (std::log(1.0 - 1.0 / x) + 0.5 * lx);
    return kPi * kPi / 6.0 + ser - lt;
  } else if (x > 1.0) {
    const double e = x - 1.0, lne = std::log(e);
    const double c0 = kPi * kPi / 6.0;
    const double c1 = 1.0 - lne;
    const double c2 = -(1.0 - 2.0 * lne) / 4.0;
    const double c3 = (1.0 - 3.0 * lne) / 9.0;
    const double c4 = -(1.0 - 4.0 * lne) / 16.0;
    const double c5 = (1.0 - 5.0 * lne) / 25.0;
    const double c6 = -(1.0 - 6.0 * lne) / 36.0;
    const double c7 = (1.0 - 7.0 * lne) / 49.0;
    const double c8 = -(1.0 - 8.0 * lne) / 64.0;
    return c0 + e * (c1 + e * (c2 + e * (c3 + e * (c4 + e * (c5 + e * (c6 + e * (c7 + e * c8)))))));
  } else if (x == 1.0) {
    return kPi * kPi / 6.0;
  } else if (x > 0.5) {
    const double ser = series_2(1.0 - x);
    return kPi * kPi / 6.0 - ser - std::log(x) * std::log(1.0 - x);
  } else if (x > 0.25) {
    return series_2(x);
  } else if (x > 0.0) {
    return series_1(x);
  }
  return 0.0;
}
inline double dilog(double x) {
  if (x >= 0.0) return dilog_xge0(x);
  return -dilog_xge0(-x) + 0.5 * dilog_xge0(x * x);
}

typedef std::complex<double> cplx;

// sum_{k>=1} z^k / k^2 with the same trigonometric recurrence GSL uses
inline cplx c_series_1(double r, double x, double y) {
  const double ct = x / r, st = y / r, al = 1.0 - ct, be = st;
  double ck = ct, sk = st, rk = r, sr = r * ck, si = r * sk;
  const int kmax = 50 + (int)(22.0 / (-std::log(r)));
  for (int k = 2; k < kmax; ++k) {
    const double ck0 = ck;
    ck = ck - (al * ck + be * sk);
    sk = sk - (al * sk - be * ck0);
    rk *= r;
    const double dr = rk / ((double)k * k) * ck, di = rk / ((double)k * k) * sk;
    sr += dr; si += di;
    if (std::fabs((dr * dr + di * di) / (sr * sr + si * si)) < kEps * kEps) break;
  }
  return cplx(sr, si);
}
inline cplx c_series_2(double r, double x, double y) {  // 1 + (1-z)ln(1-z)/z + sum z^k/(k^2(k+1))
  if (r == 0.0) return cplx(0.0, 0.0);
  const double ct = x / r, st = y / r, al = 1.0 - ct, be = st;
  double ck = ct, sk = st, rk = r, sr = 0.5 * r * ck, si = 0.5 * r * sk;
  const int kmax = 30 + (int)(18.0 / (-std::log(r)));
  for (int k = 2; k < kmax; ++k) {
    const double ck0 = ck;
    ck = ck - (al * ck + be * sk);
    sk = sk - (al * sk - be * ck0);
    rk *= r;
    const double f = rk / ((double)k * k * (k + 1.0));
    const double dr = f * ck, di = f * sk;
    sr += dr; si += di;
    if (std::fabs((dr * dr + di * di) / (sr * sr + si * si)) < kEps * kEps) break;
  }
  const double lr = std::log(std::hypot(1.0 - x, -y)), lth = std::atan2(-y, 1.0 - x);
  const double tx = (lr * x + lth * y) / (r * r), ty = (-lr * y + lth * x) / (r * r);
  const double rx = (1.0 - x) * tx + y * ty, ry = (1.0 - x) * ty - y * tx;
  return cplx(sr + rx + 1.0, si + ry);
}
// Bernoulli series Li2(z) = u - u^2/4 + sum_k B_2k u^(2k+1)/(2k+1)!, u = -ln(1-z)  (|u| < 2 pi)
inline cplx c_bernoulli(double x, double y) {
  static const double b[] = {1.0 / 36.0, -1.0 / 3600.0, 1.0 / 211680.0, -1.0 / 10886400.0,
                             1.0 / 526901760.0, -4.0647616451442255e-11, 8.9216910204564526e-13,
                             -1.9939295860721076e-14, 4.5189800296199182e-16, -1.0356517612181247e-17,
                             2.3952186210261867e-19, -5.5817858743250093e-21, 1.3091507554183213e-22,
                             -3.0874198024267403e-24, 7.3159756527022034e-26};
  const cplx u = -std::log(cplx(1.0 - x, -y));
  const cplx u2 = u * u;
  cplx s(0.0, 0.0);
  for (int k = 14; k >= 0; --k) s = s * u2 + b[k];
  return u + u2 * (-0.25 + u * s);
}
inline cplx c_fundamental(double r, double x, double y) {
  if (r > 0.98) return c_bernoulli(x, y);
  if (r > 0.25) return c_series_2(r, x, y);
  return c_series_1(r, x, y);
}
inline cplx c_unitdisk(double x, double y) {
  const double split = 0.732, zeta2 = kPi * kPi / 6.0;
  const double r = std::hypot(x, y);
  if (x > split) {
    const double xt = 1.0 - x, yt = -y, rt = std::hypot(xt, yt);
    const cplx t = c_fundamental(rt, xt, yt);
    const double lnz = std::log(r), lnomz = std::log(rt);
    const double argz = std::atan2(y, x), argomz = std::atan2(yt, xt);
    return cplx(-t.real() + zeta2 - lnz * lnomz + argz * argomz, -t.imag() - argz * lnomz - argomz * lnz);
  }
  return c_fundamental(r, x, y);
}
inline cplx cdilog_xy(double x, double y) {
  const double zeta2 = kPi * kPi / 6.0;
  const double r2 = x * x + y * y;
  // GSL's rule for an argument exactly on the real axis.  REFSHIM_GSL_REAL_AXIS_FLIP builds the
  // OTHER sign (the y -> 0+ limit): used only by tests/test_oracle_cpu.py to show that the reference's own positivity
  // check (nuSIprop.hpp:1505) tells the two conventions apart.
#ifdef REFSHIM_GSL_REAL_AXIS_FLIP
  if (y == 0.0) return cplx(dilog(x), (x >= 1.0) ? +kPi * std::log(x) : 0.0);
#else
  if (y == 0.0) return cplx(dilog(x), (x >= 1.0) ? -kPi * std::log(x) : 0.0);
#endif
  if (std::fabs(r2 - 1.0) < kEps) return c_bernoulli(x, y);
  if (r2 < 1.0) return c_unitdisk(x, y);
  const double r = std::sqrt(r2);
  const cplx t = c_unitdisk(x / r2, -y / r2);
  const double th = std::atan2(y, x), tha = std::fabs(th), ths = (th < 0.0 ? -1.0 : 1.0);
  const double lre = std::log(r), lim = ths * (tha - kPi);
  return cplx(-t.real() - 0.5 * (lre * lre - lim * lim) - zeta2, -t.imag() - 0.5 * (2.0 * lre * lim));
}
}  // namespace refshim_gsl

#ifdef REFSHIM_SHARED_LI2
// The libm-locked oracle (oracle/_ref/libnusi_ref_locked.so): the two GSL entry points are served by
// the PRODUCT's own Li2 (nusiprop_b200/csrc/nusi_math.h compiled for the host, its logarithms and arguments from
// nusi_locked_math.h), so that this build and the product's locked build share every transcendental and differ only
// where the product's arithmetic differs from the reference's.  Not an independent check of Li2 (that is the default
// stand-in above, pinned to mpmath); a check of everything else.
#ifndef NUSI_LOCKED_MATH
#define NUSI_LOCKED_MATH
#endif
#include "../../../nusiprop_b200/csrc/nusi_math.h"
inline double gsl_sf_dilog(const double x) { return nusi::li2_real(x); }
inline int gsl_sf_complex_dilog_xy_e(const double x, const double y, gsl_sf_result *re, gsl_sf_result *im) {
  const nusi::cplx v = nusi::cli2(x, y);
  re->val = v.re; re->err = 0.0;
  im->val = v.im; im->err = 0.0;
  return 0;
}
#else
inline double gsl_sf_dilog(const double x) { return refshim_gsl::dilog(x); }
inline int gsl_sf_complex_dilog_xy_e(const double x, const double y, gsl_sf_result *re, gsl_sf_result *im) {
  const refshim_gsl::cplx v = refshim_gsl::cdilog_xy(x, y);
  re->val = v.real(); re->err = 0.0;
  im->val = v.imag(); im->err = 0.0;
  return 0;
}
#endif
#endif
