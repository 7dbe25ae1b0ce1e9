// nusi_interp.h -- device lookup of the double-scalar (phi phi) cross-section tables.
//
// Replaces interp::spline_ND<2>/<3>::f_eval (reference interp.hpp:345-467) and its weight
// table (interp.hpp:576-636) for the two tables the hot path reads:
//   alphatilde_phiphi.bin  5000 x 100       nodes (|tbar_plus|, log10 delta)          nuSIprop.hpp:168,1199
//   alpha_phiphi.bin       1000 x 1000 x 100 nodes (sbar, n, log10 delta)              nuSIprop.hpp:169,1483
// Axis 0 is interpolated in ln(x) (isLog = {true,false,...}), node search is a binary search
// (isRegular = false), evaluation points must be STRICTLY inside the node range (the
// reference prints and exit(1)s otherwise -- here the point's status gets NUSI_WARN_PP_RANGE
// and the value is NaN), and each axis uses the same 4-point (3 at the edges) cubic stencil.
//
// Within one context the binning is fixed, so log10(delta) -- the last axis of both tables --
// is the same number for every lookup.  The host therefore collapses that axis once at
// context creation (tensor-product interpolation is separable, so contracting the last axis
// first is the same sum in a different order) and uploads
//   at_f [n0]        alphaTilde table at this delta        (40 KB)
//   a_f  [n0][n1]    alpha table at this delta             (8 MB, L2 resident)
// together with the node coordinates of the remaining axes.
#ifndef NUSI_INTERP_H
#define NUSI_INTERP_H
#include "nusi_math.h"

namespace nusi {

enum {
  NUSI_WARN_GAMMA = 1,       // reference would print "Negative cross section when computing Gamma"
  NUSI_WARN_ALPHATILDE = 2,  // ... alphaTilde
  NUSI_WARN_ALPHA = 4,       // ... alpha
  NUSI_WARN_PP_RANGE = 8     // phi-phi table lookup outside the tabulated range (reference: exit(1))
};

struct PhiPhiTables {
  // alphaTilde: axis 0 = ln|tbar_plus| (n0 nodes)
  int at_n0;
  const double *at_x0;  // [at_n0]  ln of the node coordinates
  const double *at_f;   // [at_n0]  values, log10(delta) axis already contracted
  // alpha: axis 0 = ln(sbar) (a_n0 nodes), axis 1 = n (a_n1 nodes)
  int a_n0, a_n1;
  const double *a_x0;  // [a_n0]
  const double *a_x1;  // [a_n1]
  const double *a_f;   // [a_n0][a_n1]
  // delta-axis bounds check result computed on the host (1 = log10 delta strictly inside)
  int delta_in_range;
};

// index k with x[k] <= x0 < x[k+1] by the reference's bisection (interp.hpp:375-389)
NUSI_HD int interp_locate(const double *x, int n, double x0) {
  int L = 0, R = n - 1, k = 0;
  while (L <= R) {
    const int m = (L + R) / 2;
    if (x0 < x[m])
      R = m - 1;
    else {
      k = m;
      L = m + 1;
    }
  }
  return k;
}

// The cubic stencil of one axis: first node index i0, number of nodes cnt (3 or 4) and the
// four polynomial weights evaluated at t (interp.hpp:576-636 gives the coefficients; here they
// are formed on the fly from the node coordinates instead of being tabulated).
struct AxisStencil {
  int i0, cnt;
  double c[4];
};
NUSI_HD void interp_axis(const double *x, int n, double x0, AxisStencil &S) {
  const int j = interp_locate(x, n, x0);
  const double t = (x0 - x[j]) / (x[j + 1] - x[j]);
  const double t2 = sqr(t), t3 = cub(t);
  double w[4][4];
  for (int a = 0; a < 4; ++a)
    for (int b = 0; b < 4; ++b) w[a][b] = 0;
  if (j == 0) {
    S.i0 = 0;
    S.cnt = 3;
    w[0][1] = (x[j] - x[j + 1]) / (x[j] - x[j + 2]);
    w[0][2] = (-1 + (x[j + 1] - x[j]) / (x[j] - x[j + 2]));
    w[0][3] = 1;
    w[1][1] = (x[j + 1] - x[j]) / (x[j + 1] - x[j + 2]);
    w[1][2] = (x[j] - x[j + 2]) / (x[j + 1] - x[j + 2]);
    w[2][1] = sqr(x[j + 1] - x[j]) / ((x[j + 2] - x[j + 1]) * (x[j + 2] - x[j]));
    w[2][2] = sqr(x[j + 1] - x[j]) / ((x[j + 2] - x[j + 1]) * (x[j] - x[j + 2]));
  } else if (j == n - 2) {
    S.i0 = j - 1;
    S.cnt = 3;
    w[0][1] = sqr(x[j + 1] - x[j]) / ((x[j - 1] - x[j]) * (x[j - 1] - x[j + 1]));
    w[0][2] = sqr(x[j + 1] - x[j]) / ((x[j] - x[j - 1]) * (x[j - 1] - x[j + 1]));
    w[1][1] = (x[j + 1] - x[j]) / (x[j - 1] - x[j]);
    w[1][2] = (2 * x[j] - x[j + 1] - x[j - 1]) / (x[j - 1] - x[j]);
    w[1][3] = 1;
    w[2][1] = (x[j] - x[j + 1]) / (x[j - 1] - x[j + 1]);
    w[2][2] = (x[j - 1] - x[j]) / (x[j - 1] - x[j + 1]);
  } else {
    S.i0 = j - 1;
    S.cnt = 4;
    w[0][0] = sqr(x[j + 1] - x[j]) / ((x[j] - x[j - 1]) * (x[j - 1] - x[j + 1]));
    w[0][1] = 2 * sqr(x[j + 1] - x[j]) / ((x[j - 1] - x[j]) * (x[j - 1] - x[j + 1]));
    w[0][2] = sqr(x[j + 1] - x[j]) / ((x[j] - x[j - 1]) * (x[j - 1] - x[j + 1]));
    w[1][0] = (x[j] - x[j + 1]) * (1 / (x[j - 1] - x[j]) + 1 / (x[j] - x[j + 2]));
    w[1][1] = (x[j] - x[j + 1]) * (2 / (x[j] - x[j - 1]) + 1 / (x[j + 2] - x[j]));
    w[1][2] = (2 * x[j] - x[j + 1] - x[j - 1]) / (x[j - 1] - x[j]);
    w[1][3] = 1;
    w[2][0] = (x[j + 1] - x[j]) * (1 / (x[j - 1] - x[j + 1]) + 1 / (x[j + 1] - x[j + 2]));
    w[2][1] = (x[j + 1] - x[j]) * (2 / (x[j + 1] - x[j - 1]) + 1 / (x[j + 2] - x[j + 1]));
    w[2][2] = (x[j - 1] - x[j]) / (x[j - 1] - x[j + 1]);
    w[3][0] = sqr(x[j + 1] - x[j]) / ((-x[j + 1] + x[j + 2]) * (-x[j] + x[j + 2]));
    w[3][1] = sqr(x[j + 1] - x[j]) / ((x[j + 1] - x[j + 2]) * (-x[j] + x[j + 2]));
  }
  for (int a = 0; a < 4; ++a) S.c[a] = t3 * w[a][0] + t2 * w[a][1] + t * w[a][2] + w[a][3];
}
NUSI_HD bool interp_inside(const double *x, int n, double x0) { return (x0 > x[0]) && (x0 < x[n - 1]); }

// alphaTilde_phiphi at (|tbar_plus|, log10 delta)   <- spl_alphaTilde_phiphi.f_eval  nuSIprop.hpp:1199
NUSI_HD double phiphi_alphatilde_eval(const PhiPhiTables *T, double abs_tplus, double /*log10_delta*/, int *warn) {
  const double x0 = nlog(abs_tplus);
  if (!T || !T->at_f || !T->delta_in_range || !interp_inside(T->at_x0, T->at_n0, x0)) {
    if (warn) *warn |= NUSI_WARN_PP_RANGE;
    return NAN;
  }
  AxisStencil S;
  interp_axis(T->at_x0, T->at_n0, x0, S);
  double res = 0;
  for (int a = 0; a < S.cnt; ++a) res += T->at_f[S.i0 + a] * S.c[a];
  return res;
}

// alpha_phiphi at (sbar, n, log10 delta)            <- spl_alpha_phiphi.f_eval       nuSIprop.hpp:1483
NUSI_HD double phiphi_alpha_eval(const PhiPhiTables *T, double sbar, double n, double /*log10_delta*/, int *warn) {
  const double x0 = nlog(sbar);
  if (!T || !T->a_f || !T->delta_in_range || !interp_inside(T->a_x0, T->a_n0, x0) ||
      !interp_inside(T->a_x1, T->a_n1, n)) {
    if (warn) *warn |= NUSI_WARN_PP_RANGE;
    return NAN;
  }
  AxisStencil S0, S1;
  interp_axis(T->a_x0, T->a_n0, x0, S0);
  interp_axis(T->a_x1, T->a_n1, n, S1);
  double res = 0;
  // the reference's loop runs axis 0 fastest (interp.hpp:428-458)
  for (int b = 0; b < S1.cnt; ++b)
    for (int a = 0; a < S0.cnt; ++a)
      res += T->a_f[(size_t)(S0.i0 + a) * T->a_n1 + (S1.i0 + b)] * S0.c[a] * S1.c[b];
  return res;
}

}  // namespace nusi
#endif
