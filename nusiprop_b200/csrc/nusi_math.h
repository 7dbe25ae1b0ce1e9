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
#define NUSI_HD_NOINLINE static __host__ __device__ __noinline__
#else
#define NUSI_HD inline
#define NUSI_HD_NOINLINE inline
#endif

namespace nusi {

// Elementary functions.  On the device each is ONE out-of-line copy (CUDA inlines ~1-3 KB of SASS
// per call site otherwise, and the builder kernels have ~60 call sites: the instruction cache is
// the first bottleneck of this path).  On the host they are the libm calls.
#if defined(NUSI_LOCKED_MATH)
// libm-locked build (tests only): every elementary function is nusi_locked_math.h -- the same source and
// the same bits on the host and on the device, shared with oracle/_ref/libnusi_ref_locked.so.
}  // namespace nusi
#include "nusi_locked_math.h"
namespace nusi {
NUSI_HD_NOINLINE double nlog(double x) { return nusi_locked::log(x); }
NUSI_HD_NOINLINE double nlog1p(double x) { return nusi_locked::log1p(x); }
NUSI_HD_NOINLINE double natan2(double y, double x) { return nusi_locked::atan2(y, x); }
NUSI_HD_NOINLINE double natan(double x) { return nusi_locked::atan(x); }
NUSI_HD_NOINLINE double natanh(double x) { return nusi_locked::atanh(x); }
NUSI_HD double nhypot(double x, double y) { return nusi_locked::hypot(x, y); }
#else
#if defined(__CUDA_ARCH__)
static __device__ __noinline__ double nlog(double x) { return ::log(x); }
static __device__ __noinline__ double nlog1p(double x) { return ::log1p(x); }
static __device__ __noinline__ double natan2(double y, double x) { return ::atan2(y, x); }
#else
inline double nlog(double x) { return ::log(x); }
inline double nlog1p(double x) { return ::log1p(x); }
inline double natan2(double y, double x) { return ::atan2(y, x); }
#endif
NUSI_HD double natan(double x) { return ::atan(x); }
NUSI_HD double natanh(double x) { return ::atanh(x); }
NUSI_HD double nhypot(double x, double y) { return ::hypot(x, y); }
#endif

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
NUSI_HD double cabs_(cplx a) { return nhypot(a.re, a.im); }
NUSI_HD double carg_(cplx a) { return natan2(a.im, a.re); }
NUSI_HD cplx clog_(cplx a) { return mk(nlog(nhypot(a.re, a.im)), natan2(a.im, a.re)); }
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

}  // namespace nusi
#define NUSI_BERN_LIST NUSI_BERN_1, NUSI_BERN_2, NUSI_BERN_3, NUSI_BERN_4, NUSI_BERN_5, NUSI_BERN_6, NUSI_BERN_7, NUSI_BERN_8, \
                       NUSI_BERN_9, NUSI_BERN_10, NUSI_BERN_11, NUSI_BERN_12, NUSI_BERN_13
#if defined(__CUDACC__)
static __device__ __constant__ double nusi_bern_dev[13] = {NUSI_BERN_LIST};
#endif
static const double nusi_bern_host[13] = {NUSI_BERN_LIST};
// (-1)^(n+1)/n, n = 2..17: Taylor coefficients of log(1+x) beyond the linear term
#define NUSI_LOGC_LIST -1.0 / 2, 1.0 / 3, -1.0 / 4, 1.0 / 5, -1.0 / 6, 1.0 / 7, -1.0 / 8, 1.0 / 9, -1.0 / 10, 1.0 / 11, -1.0 / 12, \
                       1.0 / 13, -1.0 / 14, 1.0 / 15, -1.0 / 16, 1.0 / 17
#if defined(__CUDACC__)
static __device__ __constant__ double nusi_logc_dev[16] = {NUSI_LOGC_LIST};
#endif
static const double nusi_logc_host[16] = {NUSI_LOGC_LIST};
namespace nusi {

// log(r) for an argument near 1 (the bin-to-bin ratios of the compact builder: |r-1| <= 2 (delta-1), delta the
// bin ratio).  x = r - 1 is exact for 1/2 <= r <= 2; for |x| <= 1/8 the series  x + x^2 (c2 + c3 x + ... + c17 x^15)
// is truncated below 2^-55 |x| and the bracket carries <= 1/8 of the result's weight, so the value is within one
// ulp of log(r) -- the accuracy of the libm logarithm it replaces -- in 18 FP64 instructions instead of ~30 plus a call.
#if defined(__CUDA_ARCH__)
#define NUSI_LOGC(n) nusi_logc_dev[(n)-2]
#else
#define NUSI_LOGC(n) nusi_logc_host[(n)-2]
#endif
NUSI_HD double nlog_near1(double r) {
  const double x = r - 1.0;
  if (!(fabs(x) <= 0.125)) return nlog(r);
  double q = NUSI_LOGC(17);
  q = fma(q, x, NUSI_LOGC(16));
  q = fma(q, x, NUSI_LOGC(15));
  q = fma(q, x, NUSI_LOGC(14));
  q = fma(q, x, NUSI_LOGC(13));
  q = fma(q, x, NUSI_LOGC(12));
  q = fma(q, x, NUSI_LOGC(11));
  q = fma(q, x, NUSI_LOGC(10));
  q = fma(q, x, NUSI_LOGC(9));
  q = fma(q, x, NUSI_LOGC(8));
  q = fma(q, x, NUSI_LOGC(7));
  q = fma(q, x, NUSI_LOGC(6));
  q = fma(q, x, NUSI_LOGC(5));
  q = fma(q, x, NUSI_LOGC(4));
  q = fma(q, x, NUSI_LOGC(3));
  q = fma(q, x, NUSI_LOGC(2));
  return fma(x * x, q, x);
}

// Li2(y) from u = -ln(1-y), |u| <= 0.6932: (u/2pi)^2 <= 0.0122, 9 Bernoulli terms leave < 1e-19.
// On the device the coefficients live in constant memory: a DFMA takes a constant-bank operand directly,
// whereas a 64-bit immediate costs two moves per use (8 % of the builder's issue slots before this).
#if defined(__CUDA_ARCH__)
#define NUSI_BERN(k) nusi_bern_dev[(k)-1]
#else
#define NUSI_BERN(k) nusi_bern_host[(k)-1]
#endif
NUSI_HD double li2_from_u(double u) {
  const double w = u * u;
  double p = NUSI_BERN(9);
  p = fma(p, w, NUSI_BERN(8));
  p = fma(p, w, NUSI_BERN(7));
  p = fma(p, w, NUSI_BERN(6));
  p = fma(p, w, NUSI_BERN(5));
  p = fma(p, w, NUSI_BERN(4));
  p = fma(p, w, NUSI_BERN(3));
  p = fma(p, w, NUSI_BERN(2));
  p = fma(p, w, NUSI_BERN(1));
  return fma(w, fma(u, p, -0.25), u);  // u - u^2/4 + u^3 p
}

// Real part of Li2(x) for any real x (gsl_sf_dilog semantics).  Every branch reduces to ONE core
// evaluation; where the reduced argument is 1-x the core's u = -ln(x) is the logarithm the
// reflection term needs anyway, so it is computed once.
NUSI_HD_NOINLINE double li2_real(double x) {
  double u, add, sign;
  if (x < -1.0) {  // inversion: -zeta2 - ln^2(-x)/2 - Li2(1/x)
    const double l = nlog(-x);
    u = -nlog1p(-1.0 / x);
    add = -kZeta2 - 0.5 * l * l;
    sign = -1.0;
  } else if (x <= 0.5) {
    u = -nlog1p(-x);
    add = 0.0;
    sign = 1.0;
  } else if (x < 1.0) {  // reflection: zeta2 - ln(x) ln(1-x) - Li2(1-x)
    const double lx = nlog(x);
    u = -lx;
    add = kZeta2 - lx * nlog1p(-x);
    sign = -1.0;
  } else if (x == 1.0) {
    return kZeta2;
  } else if (x <= 2.0) {  // zeta2 - ln(x) ln(x-1) - Li2(1-x)
    const double lx = nlog(x);
    u = -lx;
    add = kZeta2 - lx * nlog(x - 1.0);
    sign = -1.0;
  } else {  // pi^2/3 - ln^2(x)/2 - Li2(1/x)
    const double l = nlog(x);
    u = -nlog1p(-1.0 / x);
    add = 2.0 * kZeta2 - 0.5 * l * l;
    sign = -1.0;
  }
  return add + sign * li2_from_u(u);
}

// Complex core from u = -ln(1-w): |w| <= 1 and Re w <= 1/2  =>  |u| <= 1.256, (|u|/2pi)^2 <= 0.04.
NUSI_HD cplx li2_from_u_cplx(cplx u) {
  const double wr = u.re * u.re - u.im * u.im, wi = 2.0 * (u.re * u.im);
  double pr = NUSI_BERN(13), pi = 0.0;
#define NUSI_CSTEP(C)                          \
  {                                            \
    const double tr = fma(pr, wr, fma(-pi, wi, C)); \
    pi = fma(pr, wi, pi * wr);                 \
    pr = tr;                                   \
  }
  NUSI_CSTEP(NUSI_BERN(12))
  NUSI_CSTEP(NUSI_BERN(11))
  NUSI_CSTEP(NUSI_BERN(10))
  NUSI_CSTEP(NUSI_BERN(9))
  NUSI_CSTEP(NUSI_BERN(8))
  NUSI_CSTEP(NUSI_BERN(7))
  NUSI_CSTEP(NUSI_BERN(6))
  NUSI_CSTEP(NUSI_BERN(5))
  NUSI_CSTEP(NUSI_BERN(4))
  NUSI_CSTEP(NUSI_BERN(3))
  NUSI_CSTEP(NUSI_BERN(2))
  NUSI_CSTEP(NUSI_BERN(1))
#undef NUSI_CSTEP
  // u + w (u p - 1/4)
  const double qr = fma(u.re, pr, fma(-u.im, pi, -0.25)), qi = fma(u.re, pi, u.im * pr);
  return mk(fma(wr, qr, fma(-wi, qi, u.re)), fma(wr, qi, fma(wi, qr, u.im)));
}

// gsl_sf_complex_dilog_xy_e semantics: y == 0.0 exactly -> (Re Li2(x), x >= 1 ? -pi ln x : 0);
// |z| > 1 -> inversion Li2(z) = -Li2(1/z) - zeta2 - ln^2(-z)/2 with Im ln(-z) = sgn(th)(|th|-pi);
// Re w > 1/2 -> reflection Li2(w) = -Li2(1-w) + zeta2 - ln(w) ln(1-w).  One core evaluation.
NUSI_HD_NOINLINE cplx cli2(double x, double y) {
  if (y == 0.0) return mk(li2_real(x), (x >= 1.0) ? -kPi * nlog(x) : 0.0);
  const double r2 = x * x + y * y;
  const bool inv = r2 > 1.0;
  double wx = x, wy = y;
  if (inv) {
    wx = x / r2;
    wy = -y / r2;
  }
  const bool refl = wx > 0.5;
  // ln(1-w) = (lnomw, argomw) and, if reflected, ln(w) = (lnw, argw)
  const double lnomw = 0.5 * nlog1p((wx * wx - 2.0 * wx) + wy * wy), argomw = natan2(-wy, 1.0 - wx);
  cplx res;
  if (refl) {
    const double lnw = 0.5 * nlog(wx * wx + wy * wy), argw = natan2(wy, wx);
    const cplx t = li2_from_u_cplx(mk(-lnw, -argw));  // Li2(1-w): u = -ln(w)
    res = mk(-t.re + kZeta2 - lnw * lnomw + argw * argomw, -t.im - argw * lnomw - argomw * lnw);
  } else {
    res = li2_from_u_cplx(mk(-lnomw, -argomw));
  }
  if (inv) {
    const double th = natan2(y, x);
    const double lre = 0.5 * nlog(r2);
    const double lim = (th < 0.0 ? -1.0 : 1.0) * (fabs(th) - kPi);
    res = mk(-res.re - 0.5 * (lre * lre - lim * lim) - kZeta2, -res.im - lre * lim);
  }
  return res;
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
static __device__ __constant__ double nusi_li2_cheb_dev[NUSI_POLYLOG_NCHEB] = {NUSI_LI2_CHEB};
static __device__ __constant__ double nusi_li3_cheb_dev[NUSI_POLYLOG_NCHEB] = {NUSI_LI3_CHEB};
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
