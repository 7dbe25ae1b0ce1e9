// nusi_phiphi_gen.h -- generator of the double-scalar (phi phi) cross-section tables that the hot
// path looks up (SURVEY 8(f) "next #1").  Replaces the reference's offline pipeline
//   xsec/funcs.c:12-39          primitive() and dsigma_phiphi_over_tauphi()   (the integrand)
//   xsec/tables_phiphi.py:21-59  scipy dblquad per node of the two grids
//   xsec/text_to_binary.cpp      text -> float32 records
// by fixed Gauss-Legendre quadrature that a B200 evaluates for all 1e8 + 5e5 nodes in seconds
// (dblquad needs days).  The .bin files are "available upon request" only (reference README.md:52).
//
// The integrand F(sbar, tbar) = P(min(tbar, u_hi(s)), s) - P(u_lo(s), s)  (0 if the range is empty),
//   u_hi/lo(s) = -1 - (sqrt(s) -/+ sqrt(s-4))^2 / 4,
// is smooth except on the curve  s = s*(t) = w + 2 + 1/w,  w = -1 - t  (t < -1), where its s-derivative
// jumps (for t < -2 it vanishes identically below the curve), and at the threshold s = 4 where it
// starts like sqrt(s-4).  The quadrature therefore splits the inner s-range at s*(t), substitutes
// s = 4 + x^2 on a piece that starts at the threshold, and splits the outer t-range at t = -1, -2;
// every piece is an 8-point Gauss-Legendre rule.  For t < -1 the integrand also carries log(s - (1 - t))
// whose branch point sits 1/w below the curve: ranges that end near it are covered by panels that
// double in size away from the branch point (and likewise in t for the outer integral).
#ifndef NUSI_PHIPHI_GEN_H
#define NUSI_PHIPHI_GEN_H
#include "nusi_math.h"

namespace nusi {

// \int dtau/(-tau) dsigma/dtau            xsec/funcs.c:12-20
NUSI_HD double pp_primitive(double tau, double s) {
  return (1 / (1 + tau) + 1 / ((-1 + s) * (-1 + s + tau)) +
          (-(sqr(-1 + s) * (4 + (-3 + s) * s) * log(-1 - tau)) + (-2 + s) * cub(s) * log(-tau) +
           (-4 + s * (9 + (-5 + s) * s)) * log(-1 + s + tau)) /
              ((-2 + s) * sqr(-1 + s))) /
         (64. * kPi * sqr(s));
}

// dsigma_phiphi_over_tauphi(sbar, tbar)    xsec/funcs.c:22-39
NUSI_HD double pp_integrand(double sbar, double tbar) {
  const double rs = sqrt(sbar), rs4 = sqrt(sbar - 4);
  const double hi = -1 - 1. / 4. * sqr(rs - rs4);
  const double upperLim = (tbar < hi) ? tbar : hi;
  const double lowerLim = -1 - 1. / 4. * sqr(rs + rs4);
  if (upperLim < lowerLim) return 0;
  return pp_primitive(upperLim, sbar) - pp_primitive(lowerLim, sbar);
}

// 8-point Gauss-Legendre rule on [-1, 1] (positive half; symmetric)
NUSI_HD double pp_gl8_x(int i) {
  return i == 0 ? 0.18343464249564980 : i == 1 ? 0.52553240991632899 : i == 2 ? 0.79666647741362674 : 0.96028985649753623;
}
NUSI_HD double pp_gl8_w(int i) {
  return i == 0 ? 0.36268378337836198 : i == 1 ? 0.31370664587788729 : i == 2 ? 0.22238103445337447 : 0.10122853629037626;
}

// ---- quadrature building blocks ----------------------------------------------------------------
// 8-point rule on [a, b]; sqrt_map: substitute s = 4 + x^2 (removes the sqrt(s-4) threshold behaviour)
template <class F>
NUSI_HD double pp_gl8(const F &f, double a, double b, bool sqrt_map) {
  if (!(b > a)) return 0;
  double acc = 0;
  if (sqrt_map) {
    const double xa = sqrt(fmax(a - 4, 0.0)), xb = sqrt(b - 4), h = 0.5 * (xb - xa), c = 0.5 * (xb + xa);
    for (int i = 0; i < 4; ++i)
      for (int sg = -1; sg <= 1; sg += 2) {
        const double x = c + sg * h * pp_gl8_x(i);
        acc += pp_gl8_w(i) * 2 * x * f(4 + x * x);
      }
    return acc * h;
  }
  const double h = 0.5 * (b - a), c = 0.5 * (b + a);
  for (int i = 0; i < 4; ++i)
    for (int sg = -1; sg <= 1; sg += 2) acc += pp_gl8_w(i) * f(c + sg * h * pp_gl8_x(i));
  return acc * h;
}
// [a, b] with a logarithmic branch point of f at p <= a: panels [p+d, p+2d], d doubling from a-p, so
// that every panel keeps its distance from p (Bernstein parameter 5.8: 8 points give ~1e-12).
template <class F>
NUSI_HD double pp_graded_from_below(const F &f, double a, double b, double p, bool allow_sqrt_map) {
  if (!(b > a)) return 0;
  double d = a - p;
  if (!(d > 0)) d = (b - a) * 9.3132257461547852e-10;  // 2^-30: endpoint singularity, integrable
  double acc = 0, lo = a;
  for (int k = 0; k < 64 && lo < b; ++k) {
    double hi = p + 2 * d;
    if (hi >= b || k == 63) hi = b;
    acc += pp_gl8(f, lo, hi, allow_sqrt_map && lo < 8.0);
    lo = hi;
    d *= 2;
  }
  return acc;
}
// mirror image: branch point at p >= b
template <class F>
NUSI_HD double pp_graded_from_above(const F &f, double a, double b, double p) {
  if (!(b > a)) return 0;
  double d = p - b;
  if (!(d > 0)) d = (b - a) * 9.3132257461547852e-10;
  double acc = 0, hi = b;
  for (int k = 0; k < 64 && hi > a; ++k) {
    double lo = p - 2 * d;
    if (lo <= a || k == 63) lo = a;
    acc += pp_gl8(f, lo, hi, false);
    hi = lo;
    d *= 2;
  }
  return acc;
}

// ---- inner integral I(t) = \int_{s_lo}^{s_hi} ds F(s, t),  s_lo >= 4 ---------------------------------
// For t < -1, F(., t) contains log(s - s0) with s0 = 1 - t = s*(t) - 1/w just below the curve s*(t):
// the part of the range above the curve is graded away from s0.
struct PPInnerF {
  double t;
  NUSI_HD double operator()(double s) const { return pp_integrand(s, t); }
};
NUSI_HD double pp_inner(double t, double s_lo, double s_hi) {
  if (!(s_hi > s_lo)) return 0;
  PPInnerF f;
  f.t = t;
  if (!(t < -1)) return pp_gl8(f, s_lo, s_hi, s_lo < 8.0);
  const double w = -1 - t, kink = w + 2 + 1 / w, s0 = 1 - t;
  double acc = 0;
  if (t < -2) {  // F == 0 below the curve
    if (kink >= s_hi) return 0;
    if (kink > s_lo) s_lo = kink;
  } else if (kink > s_lo) {  // -2 <= t < -1: below the curve F does not depend on t and is smooth
    const double top = fmin(kink, s_hi);
    acc = pp_gl8(f, s_lo, top, s_lo < 8.0);
    if (kink >= s_hi) return acc;
    s_lo = kink;
  }
  return acc + pp_graded_from_below(f, s_lo, s_hi, s0, true);
}

// ---- outer integral over t in [tplus, tminus] ----------------------------------------------------
//   MODE 0 (alpha):      s in [s_lo, s_hi] fixed                                  tables_phiphi.py:53-55
//   MODE 1 (alphaTilde): s in [max(-t, 4, -t^2/(1+t)), s_hi]                     tables_phiphi.py:32-34
template <int MODE>
struct PPOuterF {
  double s_lo, s_hi;
  NUSI_HD double operator()(double t) const {
    double lo = s_lo;
    if (MODE == 1) {
      lo = fmax(-t, 4.0);
      if (t < -1) lo = fmax(lo, -t * t / (1 + t));
    }
    return pp_inner(t, lo, s_hi);
  }
};
// I(t) is smooth in t except at t = -1, -2 and where the curve s*(t) crosses an s-limit S:
// w + 2 + 1/w = S  <=>  w = ((S-2) -/+ sqrt((S-2)^2 - 4))/2,  t = -1 - w.   The outer range is cut there.
NUSI_HD void pp_add_cut(double *cuts, int &n, double t, double lo, double hi) {
  if (!(t > lo && t < hi)) return;
  int i = n;
  while (i > 0 && cuts[i - 1] > t) {
    cuts[i] = cuts[i - 1];
    --i;
  }
  cuts[i] = t;
  ++n;
}
NUSI_HD void pp_curve_roots(double S, double &w_small, double &w_large) {
  const double b = S - 2, disc = b * b - 4;
  if (!(disc > 0)) {
    w_small = w_large = 1.0;
    return;
  }
  w_large = 0.5 * (b + sqrt(disc));
  w_small = 1 / w_large;
}
template <int MODE>
NUSI_HD double pp_outer(double tplus, double tminus, double s_lo, double s_hi) {
  double cuts[8];
  int n = 0;
  double wS1, wS2, wL1 = 1, wL2 = 1;
  pp_curve_roots(s_hi, wS1, wS2);
  pp_add_cut(cuts, n, -2.0, tplus, tminus);
  pp_add_cut(cuts, n, -1.0, tplus, tminus);
  pp_add_cut(cuts, n, -1 - wS2, tplus, tminus);
  pp_add_cut(cuts, n, -1 - wS1, tplus, tminus);
  if (MODE == 0) {
    pp_curve_roots(s_lo, wL1, wL2);
    pp_add_cut(cuts, n, -1 - wL2, tplus, tminus);
    pp_add_cut(cuts, n, -1 - wL1, tplus, tminus);
  }
  cuts[n] = tminus;
  PPOuterF<MODE> f;
  f.s_lo = s_lo;
  f.s_hi = s_hi;
  double acc = 0, a = tplus;
  for (int k = 0; k <= n; ++k) {
    const double b = cuts[k];
    if (b <= -2.0) {
      // w > 1: I(t) inherits the branch points t = 1 - S (and t = 1 - s_lo once the curve is below s_lo)
      const double p = (MODE == 0 && a >= -1 - wL2) ? 1 - s_lo : 1 - s_hi;
      if (a >= -1 - wS2) acc += (p < a) ? pp_graded_from_below(f, a, b, p, false) : pp_gl8(f, a, b, false);
    } else if (a >= -2.0 && b < -1.0) {
      acc += pp_graded_from_above(f, a, b, -1.0);  // s*(t) ~ 1/w as t -> -1
    } else {
      acc += pp_gl8(f, a, b, false);
    }
    a = b;
  }
  return acc;
}

// node of alpha_phiphi: (sbar_plus, n, log10 delta)        tables_phiphi.py:39-59
NUSI_HD double pp_alpha_node(double splus, double n, double l10d) {
  const double delta = pow(10.0, l10d);
  const double smin = splus / delta;
  const double tmin = -smin / pow(delta, n);
  const double tplus = tmin * delta;
  double integral = pp_outer<0>(tplus, tmin, fmax(smin, 4.0), splus);
  if (integral < 1e-37) integral = 0;
  return integral;
}
// node of alphatilde_phiphi: (|tbar_plus|, log10 delta)     tables_phiphi.py:21-37
NUSI_HD double pp_alphatilde_node(double abs_tplus, double l10d) {
  const double delta = pow(10.0, l10d);
  const double tplus = -abs_tplus, tmin = tplus / delta;
  return pp_outer<1>(tplus, tmin, 4.0, -tplus);
}

}  // namespace nusi
#endif
