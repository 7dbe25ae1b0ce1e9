// nusi_math.h -- FP64 special functions and complex helpers for the nuSIprop hot path.
//
// Everything here is `__host__ __device__` so the very same source is (a) compiled by nvcc
// into the sm_100a kernels (the product) and (b) compiled by g++ into the host-side unit
// checks under tests/ (development aid only; the shipped library has NO host compute path).
//
// Functions and the reference behaviour they replace (all in /root/reference):
//   li2_real(x)      Re Li2(x), any real x   <- gsl_sf_dilog            (nuSIprop.hpp:1098,1202,1375-1398; aux.hpp:112-165)
//   cli2(x,y)        complex Li2(x+iy)       <- gsl_sf_complex_dilog_xy_e (nuSIprop.hpp:1444-1451; aux.hpp:92-93)
//                    including GSL's rule for y == 0.0 exactly: Im = (x>=1) ? -pi*ln(x) : 0
//   li2_neg / li3_neg  Li2, Li3 on [-1,0]    <- polylogarithm::Li2/Li3   (nuSIprop.hpp:628-636)
//
// Algorithm (device-friendly, branch-light): every argument is mapped by inversion /
// reflection onto a region where u = -ln(1-z) satisfies |u| <= 1.26 and then
//     Li2(z) = u - u^2/4 + sum_{k>=1} B_2k u^(2k+1)/(2k+1)!
// is summed with a fixed number of terms (Horner in u^2, explicit fma).  One log1p (real)
// or one log+atan2 (complex) per evaluation, no data-dependent loop trip counts.
//
// The translation unit is compiled with -fmad=false (nvcc) / -ffp-contract=off (g++):
// contraction happens only where fma() is written out, so host and device round the
// polynomial parts identically.
#ifndef NUSI_MATH_H
#define NUSI_MATH_H

#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define NUSI_HD __host__ __device__ __forceinline__
#define NUSI_HD_NOINLINE __host__ __device__ __noinline__
#else
#define NUSI_HD inline
#define NUSI_HD_NOINLINE inline
#endif

namespace nusi {

static const double kPi = 3.14159265358979323846264338328;
static const double kZeta2 = 1.6449340668482264364724151666460;  // pi^2/6

NUSI_HD double sqr(double x) { return x * x; }
NUSI_HD double cub(double x) { return x * x * x; }

// ----------------------------------------------------------------------------------------
// complex numbers (plain struct; C99-style semantics for the few operations the path needs)
// ----------------------------------------------------------------------------------------
struct cplx {
  double re, im;
};
NUSI_HD cplx mk(double re, double im) {
  cplx r;
  r.re = re;
  r.im = im;
  return r;
}
NUSI_HD cplx operator+(cplx a, cplx b) { return mk(a.re + b.re, a.im + b.im); }
NUSI_HD cplx operator-(cplx a, cplx b) { return mk(a.re - b.re, a.im - b.im); }
NUSI_HD cplx operator-(cplx a) { return mk(-a.re, -a.im); }
NUSI_HD cplx operator*(cplx a, cplx b) { return mk(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
NUSI_HD cplx operator*(double s, cplx a) { return mk(s * a.re, s * a.im); }
NUSI_HD cplx operator*(cplx a, double s) { return mk(a.re * s, a.im * s); }
NUSI_HD cplx operator/(cplx a, double s) { return mk(a.re / s, a.im / s); }
NUSI_HD cplx operator+(cplx a, double s) { return mk(a.re + s, a.im); }
NUSI_HD cplx operator+(double s, cplx a) { return mk(a.re + s, a.im); }
NUSI_HD cplx operator-(cplx a, double s) { return mk(a.re - s, a.im); }
NUSI_HD cplx operator-(double s, cplx a) { return mk(s - a.re, -a.im); }
// Smith's algorithm (what libgcc's __divdc3 reduces to away from over/underflow)
NUSI_HD cplx operator/(cplx a, cplx b) {
  if (fabs(b.re) >= fabs(b.im)) {
    const double r = b.im / b.re, d = b.re + b.im * r;
    return mk((a.re + a.im * r) / d, (a.im - a.re * r) / d);
  } else {
    const double r = b.re / b.im, d = b.re * r + b.im;
    return mk((a.re * r + a.im) / d, (a.im * r - a.re) / d);
  }
}
NUSI_HD cplx operator/(double s, cplx b) { return mk(s, 0.0) / b; }
NUSI_HD cplx conj(cplx a) { return mk(a.re, -a.im); }
NUSI_HD double cabs_(cplx a) { return hypot(a.re, a.im); }
NUSI_HD double carg_(cplx a) { return atan2(a.im, a.re); }
NUSI_HD cplx clog_(cplx a) { return mk(log(hypot(a.re, a.im)), atan2(a.im, a.re)); }
NUSI_HD cplx csqr(cplx a) { return a * a; }
NUSI_HD cplx ccub(cplx a) { return a * a * a; }

// ----------------------------------------------------------------------------------------
// Bernoulli-series cores
// ----------------------------------------------------------------------------------------
// B_2k/(2k+1)!, k = 1..  (mpmath, 17 digits)
#define NUSI_BERN_1 0.027777777777777778
#define NUSI_BERN_2 -0.00027777777777777778
#define NUSI_BERN_3 4.7241118669690098e-6
#define NUSI_BERN_4 -9.1857730746619636e-8
#define NUSI_BERN_5 1.8978869988970999e-9
#define NUSI_BERN_6 -4.0647616451442255e-11
#define NUSI_BERN_7 8.9216910204564526e-13
#define NUSI_BERN_8 -1.9939295860721076e-14
#define NUSI_BERN_9 4.5189800296199182e-16
#define NUSI_BERN_10 -1.0356517612181247e-17
#define NUSI_BERN_11 2.3952186210261867e-19
#define NUSI_BERN_12 -5.5817858743250093e-21
#define NUSI_BERN_13 1.3091507554183213e-22

// Li2 for real x in [-1, 1/2]:  u = -log1p(-x) in [-0.6932, 0.6932]; (u/2pi)^2 <= 0.0122
// so 9 Bernoulli terms leave a tail < 1e-19 relative.
NUSI_HD double li2_core_real(double x) {
  const double u = -log1p(-x);
  const double w = u * u;
  double p = NUSI_BERN_9;
  p = fma(p, w, NUSI_BERN_8);
  p = fma(p, w, NUSI_BERN_7);
  p = fma(p, w, NUSI_BERN_6);
  p = fma(p, w, NUSI_BERN_5);
  p = fma(p, w, NUSI_BERN_4);
  p = fma(p, w, NUSI_BERN_3);
  p = fma(p, w, NUSI_BERN_2);
  p = fma(p, w, NUSI_BERN_1);
  // u - u^2/4 + u^3 p
  return fma(w, fma(u, p, -0.25), u);
}

// Real part of Li2(x) for any real x (gsl_sf_dilog semantics).  Not inlined: called from dozens of
// sites, and the builder kernels must stay inside the instruction cache.
NUSI_HD_NOINLINE double li2_real(double x) {
  if (x < -1.0) {
    const double l = log(-x);
    return -kZeta2 - 0.5 * l * l - li2_core_real(1.0 / x);
  } else if (x <= 0.5) {
    return li2_core_real(x);
  } else if (x < 1.0) {
    return kZeta2 - log(x) * log1p(-x) - li2_core_real(1.0 - x);
  } else if (x == 1.0) {
    return kZeta2;
  } else if (x <= 2.0) {
    return kZeta2 - log(x) * log(x - 1.0) - li2_core_real(1.0 - x);
  } else {
    const double l = log(x);
    return 2.0 * kZeta2 - 0.5 * l * l - li2_core_real(1.0 / x);
  }
}

// Complex core: |w| <= 1 and Re w <= 1/2  =>  |u| <= 1.256, (|u|/2pi)^2 <= 0.04: 13 terms.
NUSI_HD cplx li2_core_cplx(double x, double y) {
  const cplx u = mk(-0.5 * log1p((x * x - 2.0 * x) + y * y), -atan2(-y, 1.0 - x));
  const cplx w = u * u;
  cplx p = mk(NUSI_BERN_13, 0.0);
  p = p * w + NUSI_BERN_12;
  p = p * w + NUSI_BERN_11;
  p = p * w + NUSI_BERN_10;
  p = p * w + NUSI_BERN_9;
  p = p * w + NUSI_BERN_8;
  p = p * w + NUSI_BERN_7;
  p = p * w + NUSI_BERN_6;
  p = p * w + NUSI_BERN_5;
  p = p * w + NUSI_BERN_4;
  p = p * w + NUSI_BERN_3;
  p = p * w + NUSI_BERN_2;
  p = p * w + NUSI_BERN_1;
  return u + w * ((u * p) - 0.25);
}

// Li2 on the closed unit disk (y != 0 handled by caller for the real axis).
NUSI_HD cplx li2_unitdisk(double x, double y) {
  if (x > 0.5) {
    // reflection  Li2(w) = -Li2(1-w) + zeta2 - ln(w) ln(1-w)
    const cplx t = li2_core_cplx(1.0 - x, -y);
    const double lnz = 0.5 * log(x * x + y * y), argz = atan2(y, x);
    const double lnomz = 0.5 * log((1.0 - x) * (1.0 - x) + y * y), argomz = atan2(-y, 1.0 - x);
    return mk(-t.re + kZeta2 - lnz * lnomz + argz * argomz, -t.im - argz * lnomz - argomz * lnz);
  }
  return li2_core_cplx(x, y);
}

// gsl_sf_complex_dilog_xy_e semantics.
NUSI_HD_NOINLINE cplx cli2(double x, double y) {
  if (y == 0.0) return mk(li2_real(x), (x >= 1.0) ? -kPi * log(x) : 0.0);
  const double r2 = x * x + y * y;
  if (r2 <= 1.0) return li2_unitdisk(x, y);
  // inversion  Li2(z) = -Li2(1/z) - zeta2 - ln^2(-z)/2,  Im ln(-z) = sgn(theta)(|theta|-pi)
  const cplx t = li2_unitdisk(x / r2, -y / r2);
  const double th = atan2(y, x);
  const double lre = 0.5 * log(r2);
  const double lim = (th < 0.0 ? -1.0 : 1.0) * (fabs(th) - kPi);
  return mk(-t.re - 0.5 * (lre * lre - lim * lim) - kZeta2, -t.im - lre * lim);
}

// ----------------------------------------------------------------------------------------
// Li2 / Li3 on [-1, 0] with relative accuracy as x -> 0-  (source term, nuSIprop.hpp:628-646)
// ----------------------------------------------------------------------------------------
#include "polylog_cheb.inc"
#if defined(__CUDA_ARCH__)
#define NUSI_CONST_TABLE __constant__
#else
#define NUSI_CONST_TABLE static const
#endif
}  // namespace nusi

#if defined(__CUDACC__)
__device__ __constant__ double nusi_li2_cheb_dev[NUSI_POLYLOG_NCHEB] = {NUSI_LI2_CHEB};
__device__ __constant__ double nusi_li3_cheb_dev[NUSI_POLYLOG_NCHEB] = {NUSI_LI3_CHEB};
#endif
static const double nusi_li2_cheb_host[NUSI_POLYLOG_NCHEB] = {NUSI_LI2_CHEB};
static const double nusi_li3_cheb_host[NUSI_POLYLOG_NCHEB] = {NUSI_LI3_CHEB};

namespace nusi {
NUSI_HD double cheb_neg(const double *c, double x) {
  // Clenshaw on t = 2x+1; same operation order as the oracle stand-in (no contraction)
  const double t = 2.0 * x + 1.0, t2 = 2.0 * t;
  double b1 = 0.0, b2 = 0.0;
  for (int j = NUSI_POLYLOG_NCHEB - 1; j >= 1; --j) {
    const double b0 = c[j] + (t2 * b1 - b2);
    b2 = b1;
    b1 = b0;
  }
  return x * (c[0] + (t * b1 - b2));
}
NUSI_HD double li2_neg(double x) {
#if defined(__CUDA_ARCH__)
  return cheb_neg(nusi_li2_cheb_dev, x);
#else
  return cheb_neg(nusi_li2_cheb_host, x);
#endif
}
NUSI_HD double li3_neg(double x) {
#if defined(__CUDA_ARCH__)
  return cheb_neg(nusi_li3_cheb_dev, x);
#else
  return cheb_neg(nusi_li3_cheb_host, x);
#endif
}

}  // namespace nusi
#endif  // NUSI_MATH_H
