// nusi_alpha.h -- the regeneration kernel  alpha(Em, Ep, Em', Ep')  : a neutrino of the source
// bin [Em',Ep'] (variable s' = 2 m_k E'/mphi^2) scatters into the bin [Em,Ep] (t = -2 m_k E/mphi^2).
// Restates calculate_flux::alpha (reference nuSIprop.hpp:1237-1520).  This is the O(NT^2) part
// of the table build and 65-85 % of the reference's evolve() time.
//
// B200-first structure.  All bins are cells of one log-uniform edge grid, so every
// transcendental of the closed forms depends on at most ONE row edge t and ONE column edge s:
//   RowEdge(t)    ln(1-t), ln|1+t|, ln(1+(2+t)^2/ga^2), arg(-1/(1+t))           O(NT) per mass
//   ColEdge(s)    ln s, ln(1+(s-1)^2/ga^2), arg(s-1+i ga), ln((2+s)/s), ...     O(NT) per mass
//   Vertex(s,t)   ln(1+s+t), three real Li2, one complex Li2, one arg           O(NT^2) per mass
// and a cell only combines its 2 row edges, 2 column edges and 4 vertices with + - * / (plus
// four ratio-logarithms of the t channel).  Neighbouring cells share edges and vertices, which
// is what the tiled kernel exploits; the per-cell arithmetic below keeps the reference's term
// order so that the same inputs round the same way.
#ifndef NUSI_ALPHA_H
#define NUSI_ALPHA_H
#include "nusi_aux.h"
#include "nusi_interp.h"

namespace nusi {

struct RowEdge {
  double t;      // t/mphi^2 at the edge, already moved off the pole t = -1 (nuSIprop.hpp:1257-1262)
  double l1m;    // nlog1p(-t)
  double labs;   // ln|1+t| as nuSIprop.hpp:1386-1387 computes it
  double lfabs;  // nlog(fabs(1+t)) as the s-t interference writes it (nuSIprop.hpp:1456)
  double lam2;   // nlog1p((2+t)^2/ga_red^2)
  double argn;   // carg(-(1/(1+t))): pi if 1+t > 0 else 0
  double lnD;    // ln((2+t)^2 + ga_red^2) = lam2 + 2 ln(ga_red)        (shared-logarithm builder)
  double phi;    // arg(2 + t + i ga_red) in (0, pi)                     (shared-logarithm builder)
  double i1p;    // 1/(1+t)                                               (compact builder)
  double i1m;    // 1/(-1+t)                                              (compact builder)
};
struct ColEdge {
  double s;      // s'/mphi^2 at the edge
  double ls;     // nlog(s)
  double lam;    // nlog1p((s-1)^2/ga_red^2)
  double args;   // carg(-1 + i ga_red + s)
  double l2s;    // nlog((2+s)/s)
  double ls2;    // nlog(s/(2+s))
  double l2;     // ln(2+s)                                               (shared-logarithm builder)
  double lnN;    // ln((s-1)^2 + ga_red^2) = lam + 2 ln(ga_red)           (shared-logarithm builder)
  double is;     // 1/s                                                   (compact builder)
};
struct Vertex {
  double l1;     // nlog1p(s+t)
  double lf;     // nlog(1+s+t)
  double dF;     // t < -1 ? Li2((1+s+t)/s) : Li2(s/(1+s+t))
  double lq;     // nlog((1+s+t)/s)          (used only when t >= -1)
  double d2;     // Li2((1+s+t)/(2+s))
  double zr_re, zr_im;  // gsl-style complex Li2 of the REAL number (1+s+t)/(1+t)
  double zc_re, zc_im;  // complex Li2 of (1+s+t)/(2 - i ga_red + t)
  double ca;     // carg(-((-1 + i ga_red + s)/(2 - i ga_red + t)))
};

NUSI_HD double t_of(double mn, double E, double mphi) {
  double t = -2 * mn * E / sqr(mphi);
  if (fabs(t + 1) < 1e-7) t += t * 1e-6;
  return t;
}
NUSI_HD double s_of(double mn, double E, double mphi) { return 2 * mn * E / sqr(mphi); }

NUSI_HD void make_row_edge(double t, double gr, bool full, RowEdge &R) {
  R.t = t;
  R.l1m = nlog1p(-t);
  if (full) {
    R.labs = (t > -1) ? nlog1p(t) : nlog(-1 - t);
    R.lfabs = nlog(fabs(1 + t));
    R.lam2 = nlog1p(sqr(2 + t) / sqr(gr));
    R.argn = (-(1 / (1 + t)) < 0) ? kPi : 0.0;
    R.lnD = R.lam2 + 2 * nlog(gr);
    R.phi = natan2(gr, 2 + t);
    R.i1p = 1 / (1 + t);
    R.i1m = 1 / (-1 + t);
  }
}
NUSI_HD void make_col_edge(double s, double gr, bool full, ColEdge &C) {
  C.s = s;
  C.ls = nlog(s);
  C.lam = nlog1p(sqr(-1 + s) / sqr(gr));
  C.args = natan2(gr, -1 + s);
  if (full) {
    C.l2s = nlog((2 + s) / s);
    C.ls2 = nlog(s / (2 + s));
    C.l2 = nlog(2 + s);
    C.lnN = C.lam + 2 * nlog(gr);
    C.is = 1 / s;
  }
}
// everything of the Majorana closed forms that depends on one (s, t) pair
NUSI_HD void make_vertex(double s, double t, double gr, Vertex &V) {
  const double one_st = 1 + s + t;
  V.l1 = nlog1p(s + t);
  V.lf = nlog(one_st);
  if (t < -1) {
    V.dF = li2_real(one_st / s);
    V.lq = 0.0;
  } else {
    V.dF = li2_real(s / one_st);
    V.lq = nlog(one_st / s);
  }
  V.d2 = li2_real(one_st / (2 + s));
  const double zr = one_st / (1 + t);
  V.zr_re = li2_real(zr);
  V.zr_im = (zr >= 1.0) ? -kPi * nlog(zr) : 0.0;
  const cplx den = mk(2 + t, -gr);
  const cplx zc = one_st / den;
  const cplx dc = cli2(zc.re, zc.im);
  V.zc_re = dc.re;
  V.zc_im = dc.im;
  V.ca = carg_(-(mk(-1 + s, gr) / den));
}

// ---- shared-logarithm vertex ------------------------------------------------------------
// Every logarithm and argument that make_vertex() evaluates is, in exact arithmetic, a signed sum
// of ONE vertex logarithm  l1 = ln(1+s+t)  and logarithms / arguments of the two edges:
//   ln(s/(1+s+t)) = ls - l1            ln((1+t)/(1+s+t)) = labs - l1      (dF, t > -1)
//   ln((1+s+t)/(2+s)) = l1 - l2        ln((1-t)/(2+s))   = l1m - l2       (d2)
//   ln zr = l1 - labs,  ln(zr-1) = ls - labs,  ln(1-zr) = ls - labs        (zr = (1+s+t)/(1+t))
//   ln|zc| = l1 - lnD/2,  arg zc = phi,  ln|1-zc| = (lnN - lnD)/2,  arg(1-zc) = args - pi + phi
//   arg(-(s-1+i ga)/(2+t-i ga)) = args + phi - pi
// so a vertex costs one log1p plus the polynomial cores; the floating-point result differs from
// make_vertex() by a few ulp of the logarithms involved (the size of the glibc-vs-CUDA libm
// difference), which is why both builders exist: opts.strict selects make_vertex().
NUSI_HD double li2_unit_logs(double x, double lnx, double ln1mx) {  // Li2(x), 0 < x < 1
  return (x <= 0.5) ? li2_from_u(-ln1mx) : kZeta2 - lnx * ln1mx - li2_from_u(-lnx);
}
NUSI_HD void make_vertex_shared(const RowEdge &R, const ColEdge &C, double gr, Vertex &V) {
  const double s = C.s, t = R.t;
  const double one_st = 1 + s + t;
  const double l1 = nlog1p(s + t);
  V.l1 = l1;
  V.lf = l1;
  // FCTR dilogarithm (nuSIprop.hpp:1373-1384)
  if (t < -1) {
    V.dF = li2_unit_logs(one_st / s, l1 - C.ls, R.labs - C.ls);
    V.lq = 0.0;
  } else {
    V.dF = li2_unit_logs(s / one_st, C.ls - l1, R.labs - l1);
    V.lq = l1 - C.ls;
  }
  V.d2 = li2_unit_logs(one_st / (2 + s), l1 - C.l2, R.l1m - C.l2);
  // real-axis dilogarithm of zr = (1+s+t)/(1+t) with GSL's imaginary part
  const double zr = one_st / (1 + t);
  const double lnzr = l1 - R.labs;  // ln|zr|
  if (zr > 2.0) {
    V.zr_re = 2.0 * kZeta2 - 0.5 * lnzr * lnzr - li2_from_u(l1 - C.ls);
  } else if (zr > 1.0) {
    V.zr_re = kZeta2 - lnzr * (C.ls - R.labs) - li2_from_u(-lnzr);
  } else if (zr == 1.0) {
    V.zr_re = kZeta2;
  } else if (zr < -1.0) {
    V.zr_re = -kZeta2 - 0.5 * lnzr * lnzr - li2_from_u(l1 - C.ls);
  } else {  // -1 <= zr < 0 (zr in [0,1) cannot occur: s > 0)
    V.zr_re = li2_from_u(R.labs - C.ls);
  }
  V.zr_im = (zr >= 1.0) ? -kPi * lnzr : 0.0;
  // complex dilogarithm of zc = (1+s+t)/(2 - i ga + t), same case split as cli2()
  const double D = sqr(2 + t) + sqr(gr);
  const double x = one_st * (2 + t) / D, y = one_st * gr / D;
  const double r2 = x * x + y * y;
  const bool inv = r2 > 1.0;
  const double wx = inv ? x / r2 : x;
  const bool refl = wx > 0.5;
  const double lnr = l1 - 0.5 * R.lnD;  // ln|zc|
  const double lnw = inv ? -lnr : lnr, argw = inv ? -R.phi : R.phi;
  const double lnomw = inv ? 0.5 * C.lnN - l1 : 0.5 * (C.lnN - R.lnD);
  const double argomw = inv ? C.args : C.args - kPi + R.phi;
  cplx res;
  if (refl) {
    const cplx tt = li2_from_u_cplx(mk(-lnw, -argw));
    res = mk(-tt.re + kZeta2 - lnw * lnomw + argw * argomw, -tt.im - argw * lnomw - argomw * lnw);
  } else {
    res = li2_from_u_cplx(mk(-lnomw, -argomw));
  }
  if (inv) {
    const double lim = R.phi - kPi;  // theta = phi > 0
    res = mk(-res.re - 0.5 * (lnr * lnr - lim * lim) - kZeta2, -res.im - lnr * lim);
  }
  V.zc_re = res.re;
  V.zc_im = res.im;
  V.ca = C.args + R.phi - kPi;
}

// ---- compact vertex -----------------------------------------------------------------------
// The closed forms use the vertex data only through three numbers per vertex:
//   l1 = ln(1+s+t)
//   G  = everything of the s-t interference that belongs to ONE vertex (nuSIprop.hpp:1452-1464: the two
//        dilogarithms, the arg term and the  l1*(...)  term), so that  alpha_st ~ G-- - G+- - G-+ + G++ + edge terms
//   H  = Li2((1+s+t)/(2+s)) - K,  K = dF (t < -1) or -dF - lq^2/2 (t >= -1), so that the dilogarithm part of the
//        t-u interference (nuSIprop.hpp:1373-1399) is  H-- - H+- - H-+ + H++
// (signs: first index s'-/s'+, second t-/t+).  Same logarithms as make_vertex_shared(); the case splits of the
// dilogarithm reductions are decided by products instead of quotients (no division in a vertex).
struct VertexC {
  double l1, G, H;
};
NUSI_HD void make_vertex_compact(const RowEdge &R, const ColEdge &C, double gr, VertexC &V) {
  const double s = C.s, t = R.t;
  const double one_st = 1 + s + t, d = 1 + t;
  const double l1 = nlog1p(s + t);
  V.l1 = l1;
  // FCTR dilogarithm and the squared logarithm that goes with it
  double K;
  if (t < -1) {
    const double lnx = l1 - C.ls, ln1mx = R.labs - C.ls;  // x = (1+s+t)/s
    K = (one_st <= 0.5 * s) ? li2_from_u(-ln1mx) : kZeta2 - lnx * ln1mx - li2_from_u(-lnx);
  } else {
    const double lnx = C.ls - l1, ln1mx = R.labs - l1;  // x = s/(1+s+t)
    const double dF = (s <= 0.5 * one_st) ? li2_from_u(-ln1mx) : kZeta2 - lnx * ln1mx - li2_from_u(-lnx);
    K = -dF - 0.5 * sqr(l1 - C.ls);
  }
  {
    const double lnx = l1 - C.l2, ln1mx = R.l1m - C.l2;  // x = (1+s+t)/(2+s)
    const double d2 = (one_st <= 0.5 * (2 + s)) ? li2_from_u(-ln1mx) : kZeta2 - lnx * ln1mx - li2_from_u(-lnx);
    V.H = d2 - K;
  }
  // real-axis dilogarithm of zr = (1+s+t)/(1+t) = 1 + s/d with GSL's imaginary part
  const double lnzr = l1 - R.labs;
  double zr_re, zr_im;
  if (d > 0) {  // zr > 1
    zr_re = (s > d) ? 2.0 * kZeta2 - 0.5 * lnzr * lnzr - li2_from_u(l1 - C.ls)
                    : kZeta2 - lnzr * (C.ls - R.labs) - li2_from_u(-lnzr);
    zr_im = -kPi * lnzr;
  } else {  // zr < 0 (1+s+t > 0 for every vertex of the table)
    zr_re = (s > -2.0 * d) ? -kZeta2 - 0.5 * lnzr * lnzr - li2_from_u(l1 - C.ls) : li2_from_u(R.labs - C.ls);
    zr_im = 0.0;
  }
  // complex dilogarithm of zc = (1+s+t)/(2 - i ga + t): |zc|^2 = (1+s+t)^2/D, Re zc = (1+s+t)(2+t)/D
  const double D = sqr(2 + t) + sqr(gr);
  const bool inv = sqr(one_st) > D;
  const bool refl = inv ? ((2 + t) > 0.5 * one_st) : (one_st * (2 + t) > 0.5 * D);
  const double lnr = l1 - 0.5 * R.lnD;
  const double lnw = inv ? -lnr : lnr, argw = inv ? -R.phi : R.phi;
  const double lnomw = inv ? 0.5 * C.lnN - l1 : 0.5 * (C.lnN - R.lnD);
  const double argomw = inv ? C.args : C.args - kPi + R.phi;
  cplx res;
  if (refl) {
    const cplx tt = li2_from_u_cplx(mk(-lnw, -argw));
    res = mk(-tt.re + kZeta2 - lnw * lnomw + argw * argomw, -tt.im - argw * lnomw - argomw * lnw);
  } else {
    res = li2_from_u_cplx(mk(-lnomw, -argomw));
  }
  if (inv) {
    const double lim = R.phi - kPi;
    res = mk(-res.re - 0.5 * (lnr * lnr - lim * lim) - kZeta2, -res.im - lnr * lim);
  }
  const double ca = C.args + R.phi - kPi;
  V.G = 2 * gr * (zr_im - res.im) - 2 * (zr_re - res.re) + 2 * gr * (R.argn - ca) * l1 +
        l1 * (C.lam - R.lam2 - 2 * (C.ls - R.lfabs));
}

// ---- compact vertex, split for parameter scans ------------------------------------------------
// Points of a scan that share (mphi, m_nu) differ only in the coupling g, which enters a vertex through
// ga_red alone (the complex dilogarithm and the arg / Breit-Wigner-logarithm terms of G).  The two functions
// below are make_vertex_compact() cut at that line: the SAME operations on the SAME operands in the same
// order, so their composition equals make_vertex_compact() bit for bit (tests/test_host_math.py asserts it on
// the host build, tests/test_gpu_parity.py on the device).
struct VertexS {
  double l1, H, zr_re;  // everything of a vertex that does not depend on g
};
NUSI_HD void make_vertex_compact_shared(const RowEdge &R, const ColEdge &C, VertexS &V) {
  const double s = C.s, t = R.t;
  const double one_st = 1 + s + t, d = 1 + t;
  const double l1 = nlog1p(s + t);
  V.l1 = l1;
  double K;
  if (t < -1) {
    const double lnx = l1 - C.ls, ln1mx = R.labs - C.ls;  // x = (1+s+t)/s
    K = (one_st <= 0.5 * s) ? li2_from_u(-ln1mx) : kZeta2 - lnx * ln1mx - li2_from_u(-lnx);
  } else {
    const double lnx = C.ls - l1, ln1mx = R.labs - l1;  // x = s/(1+s+t)
    const double dF = (s <= 0.5 * one_st) ? li2_from_u(-ln1mx) : kZeta2 - lnx * ln1mx - li2_from_u(-lnx);
    K = -dF - 0.5 * sqr(l1 - C.ls);
  }
  {
    const double lnx = l1 - C.l2, ln1mx = R.l1m - C.l2;  // x = (1+s+t)/(2+s)
    const double d2 = (one_st <= 0.5 * (2 + s)) ? li2_from_u(-ln1mx) : kZeta2 - lnx * ln1mx - li2_from_u(-lnx);
    V.H = d2 - K;
  }
  const double lnzr = l1 - R.labs;
  if (d > 0) {
    V.zr_re = (s > d) ? 2.0 * kZeta2 - 0.5 * lnzr * lnzr - li2_from_u(l1 - C.ls)
                      : kZeta2 - lnzr * (C.ls - R.labs) - li2_from_u(-lnzr);
  } else {
    V.zr_re = (s > -2.0 * d) ? -kZeta2 - 0.5 * lnzr * lnzr - li2_from_u(l1 - C.ls) : li2_from_u(R.labs - C.ls);
  }
}
// the g-dependent rest: G of make_vertex_compact() from the shared l1 and zr_re
NUSI_HD double make_vertex_compact_G(const RowEdge &R, const ColEdge &C, double gr, double l1, double zr_re) {
  const double s = C.s, t = R.t;
  const double one_st = 1 + s + t, d = 1 + t;
  const double lnzr = l1 - R.labs;
  const double zr_im = (d > 0) ? -kPi * lnzr : 0.0;
  const double D = sqr(2 + t) + sqr(gr);
  const bool inv = sqr(one_st) > D;
  const bool refl = inv ? ((2 + t) > 0.5 * one_st) : (one_st * (2 + t) > 0.5 * D);
  const double lnr = l1 - 0.5 * R.lnD;
  const double lnw = inv ? -lnr : lnr, argw = inv ? -R.phi : R.phi;
  const double lnomw = inv ? 0.5 * C.lnN - l1 : 0.5 * (C.lnN - R.lnD);
  const double argomw = inv ? C.args : C.args - kPi + R.phi;
  cplx res;
  if (refl) {
    const cplx tt = li2_from_u_cplx(mk(-lnw, -argw));
    res = mk(-tt.re + kZeta2 - lnw * lnomw + argw * argomw, -tt.im - argw * lnomw - argomw * lnw);
  } else {
    res = li2_from_u_cplx(mk(-lnomw, -argomw));
  }
  if (inv) {
    const double lim = R.phi - kPi;
    res = mk(-res.re - 0.5 * (lnr * lnr - lim * lim) - kZeta2, -res.im - lnr * lim);
  }
  const double ca = C.args + R.phi - kPi;
  return 2 * gr * (zr_im - res.im) - 2 * (zr_re - res.re) + 2 * gr * (R.argn - ca) * l1 +
         l1 * (C.lam - R.lam2 - 2 * (C.ls - R.lfabs));
}

// 3x3 Gauss-Legendre fallback on  y in [t+, t-], x in [s-, s+]  (nuSIprop.hpp:1288-1303 etc.)
// KIND 0: t and u propagators (Majorana t), 1: single propagator (Dirac t, u), 2: t-u interference
template <int KIND>
NUSI_HD double alpha_gl2d(double tplus, double tminus, double sminus, double splus) {
  const double a_y = tplus, b_y = tminus, a_x = sminus, b_x = splus;
  double acc = 0;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const double y = (b_y - a_y) / 2. * gl_x(i) + (b_y + a_y) / 2.;
      const double x = (b_x - a_x) / 2. * gl_x(j) + (b_x + a_x) / 2.;
      double F;
      if (KIND == 0)
        F = sqr(y / x) / sqr(y - 1) + sqr((-x - y) / x) / sqr((-x - y) - 1);
      else if (KIND == 1)
        F = sqr(y / x) / sqr(y - 1);
      else
        F = 2 * y * (-y - x) / sqr(x) / ((y - 1) * (-y - x - 1));
      acc += gl_w(i) * gl_w(j) * F;
    }
  acc *= 1. / 4. * (b_y - a_y) * (b_x - a_x);
  return acc;
}

// s-channel factor of a source bin: everything of alpha_s except (t- - t+)   (nuSIprop.hpp:1265-1269)
NUSI_HD double alpha_s_colfactor(const PointParams &P, double sminus, double splus) {
  const double mphi = P.mphi, Ga = P.Ga;
  if (splus < 1e-5) {
    const double gr = Ga / mphi;
    return (gr * (1 + sqr(gr) + 2 * sminus)) / sqr(1 + sqr(gr)) * (splus - sminus) +
           (gr) / sqr(1 + sqr(gr)) * sqr(splus - sminus);
  }
  return atandiff(mphi * (splus - 1) / Ga, mphi * (sminus - 1) / Ga);
}

// double-scalar production, asymptotic forms above the tabulated range (nuSIprop.hpp:1487-1492)
NUSI_HD double alpha_pp_asymptotic(double sm, double sp, double tm, double tp) {
  const double lsm = nlog(sm), lsp = nlog(sp);
  const double tail = (-6 * sm + 6 * sp - 2 * (-2 + sm) * sp * lsm + sm * sp * sqr(lsm) + 2 * sm * (-2 + sp) * lsp -
                       sm * sp * sqr(lsp));
  if (tm < -1) {
    const double a = nlog(-1 - tm), b = nlog(-tm), c = nlog(-1 - tp), d = nlog(-tp);
    return ((-sm + sp) * ((tm - tp) * (sp * (-2 + tm + tp) + sm * (-2 - 24 * sp + tm + tp)) +
                          4 * (-(sp * (1 + tm)) + sm * (-1 + 2 * sp + (-1 + sp) * tm)) * a +
                          2 * (3 * sp + sm * (3 + 4 * sp)) * tm * b +
                          4 * (sp + sp * tp + sm * (1 + tp - sp * (2 + tp))) * c -
                          2 * (3 * sp + sm * (3 + 4 * sp)) * tp * d) +
            2 * sqr(sm) * lsp *
                ((3 + 2 * sp) * (tm - tp) + 2 * sqr(sp) * ((-1 - tm) * a + tm * b + (1 + tp) * c - tp * d)) +
            2 * sqr(sp) * lsm *
                ((-3 - 2 * sm) * (tm - tp) + 2 * sqr(sm) * ((1 + tm) * a - tm * b - (1 + tp) * c + tp * d))) /
           (256. * kPi * sqr(sm) * sqr(sp));
  } else if (tp < -1) {
    const double c = nlog(-1 - tp), d = nlog(-tp);
    return ((2 * sqr(sm) * lsp * ((1 + tp) * (-3 - 2 * sp + 2 * sqr(sp) * c) - 2 * sqr(sp) * tp * d) +
             (sm - sp) * ((1 + tp) * (-3 * (sm + sp + 8 * sm * sp) + (sm + sp) * tp) +
                          4 * (-(sp * (1 + tp)) + sm * (-1 + 2 * sp + (-1 + sp) * tp)) * c +
                          2 * (3 * sp + sm * (3 + 4 * sp)) * tp * d) +
             2 * sqr(sp) * lsm * ((3 + 2 * sm) * (1 + tp) + 2 * sqr(sm) * (-((1 + tp) * c) + tp * d))) /
                (256. * kPi * sqr(sm) * sqr(sp)) +
            (-1 - tm) * tail / (128. * kPi * sm * sp));
  }
  return (tp - tm) * tail / (128. * kPi * sm * sp);
}

// Contribution of target mass k to one table entry, added channel by channel to `tot`.
//   Rm,Rp  row edges t-, t+      Cm,Cp  column edges s'-, s'+
//   Vmm = (s-,t-)  Vpm = (s+,t-)  Vmp = (s-,t+)  Vpp = (s+,t+)   (vertices: only Majorana non-resonant)
//   fs = alpha_s_colfactor of the source bin
// Logarithms of the t channel that belong to a cell EDGE rather than to the cell (two neighbouring
// entries evaluate the same expression): computed here, or handed in by the tiled kernel.
NUSI_HD double alpha_log_LR(double s, double tm, double tp) {  // nuSIprop.hpp:1284, "log(((1+s+t-)(-1+t+))/((-1+t-)(1+s+t+)))"
  return nlog(((1 + s + tm) * (-1 + tp)) / ((-1 + tm) * (1 + s + tp)));
}
NUSI_HD double alpha_log_LY(double sm, double sp, double t) {  // nuSIprop.hpp:1284, "log((s-(1+s++t))/(s+(1+s-+t)))"
  return nlog((sm * (1 + sp + t)) / (sp * (1 + sm + t)));
}
// the same two logarithms for the compact builder: arguments are bin-to-bin ratios near 1
NUSI_HD double alpha_log_LR_c(double s, double tm, double tp) {
  return nlog_near1(((1 + s + tm) * (-1 + tp)) / ((-1 + tm) * (1 + s + tp)));
}
NUSI_HD double alpha_log_LY_c(double sm, double sp, double t) {
  return nlog_near1((sm * (1 + sp + t)) / (sp * (1 + sm + t)));
}
struct CellLogs {
  double lmp, lpm;            // ln(s-/s+), ln(s+/s-)
  double LR1, LR2, LYm, LYp;  // LR at s-, s+ ; LY at t-, t+
};
// HAVE_LOGS: take the six logarithms above from `L` instead of evaluating them.
template <bool HAVE_LOGS>
NUSI_HD void alpha_cell_mass_t(const PointParams &P, const PhiPhiTables *pp, int k, double fs, const CellLogs &L,
                               const RowEdge &Rm,
                             const RowEdge &Rp, const ColEdge &Cm, const ColEdge &Cp, const Vertex &Vmm,
                             const Vertex &Vpm, const Vertex &Vmp, const Vertex &Vpp, double &tot, int *warn) {
  const double mphi = P.mphi, g = P.g, Ga = P.Ga;
  const double g4 = sqr(sqr(g)), m4 = sqr(sqr(mphi));
  const double U2 = P.U2f[k];
  const double wgt = m4 / (2 * P.mn[k]);
  const double tm = Rm.t, tp = Rp.t, sm = Cm.s, sp = Cp.s;

  /* s channel */
  double a_s = g4 / (8 * kPi * Ga * cub(mphi)) * (tm - tp) * fs;
  a_s *= U2;
  if (!P.majorana) a_s /= 2.;
  tot += wgt * a_s;
  if (!P.non_resonant) return;

  double a_t, a_u, a_tu, a_st;
  const double gr = Ga / mphi;
  if (P.majorana) {
    /* t channel (u is identical for Majorana) */
    double LR1, LR2, LYm, LYp;
    if (HAVE_LOGS) {
      LR1 = L.LR1;
      LR2 = L.LR2;
      LYm = L.LYm;
      LYp = L.LYp;
    } else {
      LR1 = alpha_log_LR(sm, tm, tp);
      LR2 = alpha_log_LR(sp, tm, tp);
      LYm = alpha_log_LY(sm, sp, tm);
      LYp = alpha_log_LY(sm, sp, tp);
    }
    const double N1 = sm * sp * (-tm + tp) * Cm.ls + sm * sp * (tm - tp) * Cp.ls - sm * sp * Vmm.l1 -
                      sm * sp * tp * Vmm.l1 + sm * sp * Vpm.l1 + sm * sp * tp * Vpm.l1 - sp * LR1 - sp * tm * LR1 -
                      sp * tp * LR1 - sp * tm * tp * LR1 + sm * sp * Vmp.lf + sm * sp * tm * Vmp.l1 + sm * LR2 +
                      sm * tm * LR2 + sm * tp * LR2 + sm * tm * tp * LR2 - sm * sp * Vpp.lf - sm * sp * tm * Vpp.l1;
    a_t = g4 / (sm * sp * 16 * kPi * m4) *
          (-((sm - sp) * (3 + 2 * tm * (-1 + tp) - 2 * tp) * (tm - tp)) / ((-1 + tm) * (-1 + tp)) +
           2 * (N1) / ((1 + tm) * (1 + tp)) -
           ((sm * sp * LYm) / sqr(1 + tm) +
            (((sm - sp) * (tm - tp) * (1 + tp)) / (1 + tm) - sm * sp * LYp) / sqr(1 + tp)));
    if (a_t < 0) a_t = alpha_gl2d<0>(tp, tm, sm, sp) * (g4 / (16 * kPi * m4));
    a_t *= U2;
    a_u = a_t;
  } else {
    const double core = (-((tm - tp) * (2 + tm * (-1 + tp) - tp)) - 2 * (-1 + tm) * (-1 + tp) * (Rm.l1m - Rp.l1m));
    a_t = 3. / 2. * g4 / (32 * kPi * m4 * sm * sp * (-1 + tm) * (-1 + tp)) * (sm - sp) * core;
    if (a_t < 0) a_t = alpha_gl2d<1>(tp, tm, sm, sp) * (3. / 2. * g4 / (32 * kPi * m4));
    a_t *= U2;
    a_u = 1. / 2. * g4 / (32 * kPi * m4 * sm * sp * (-1 + tm) * (-1 + tp)) * (sm - sp) * core;
    if (a_u < 0) a_u = alpha_gl2d<1>(tp, tm, sm, sp) * (1. / 2. * g4 / (32 * kPi * m4));
    a_u *= U2;
  }
  tot += wgt * a_t;
  tot += wgt * a_u;

  /* t-u interference (Majorana only) */
  if (P.majorana) {
    double F_tp, F_tm;
    if (tp < -1)
      F_tp = Vmp.dF - Vpp.dF;
    else
      F_tp = -Vmp.dF + Vpp.dF - 0.5 * (sqr(Vmp.lq) - sqr(Vpp.lq));
    if (tm < -1)
      F_tm = -Vmm.dF + Vpm.dF;
    else
      F_tm = Vmm.dF - Vpm.dF + 0.5 * (sqr(Vmm.lq) - sqr(Vpm.lq));
    const double lmp = HAVE_LOGS ? L.lmp : nlog(sm / sp);
    const double lpm = HAVE_LOGS ? L.lpm : nlog(sp / sm);
    a_tu = g4 / (32 * kPi * m4 * sm * sp * (1 + tm) * (1 + tp)) *
           (-4 * (sm - sp) * (1 + tm) * (tm - tp) * (1 + tp) + 2 * sm * sp * tp * (lmp - Vmm.l1 + Vpm.l1) +
            2 * sp * (1 + tm) * (1 + tp) * (Rm.l1m - Vmm.l1 - Rp.l1m + Vmp.l1) -
            2 * sm * (1 + tm) * (1 + tp) * (Rm.l1m - Vpm.l1 - Rp.l1m + Vpp.l1) +
            2 * sm * sp * (-Vmm.l1 + Vpm.l1 + Vmp.l1 - Vpp.l1) +
            sm * sp * (1 + tm) * (1 + tp) *
                (Cm.l2s * (Cp.ls + Vmp.l1) - Cp.l2s * (Cm.ls + Vpp.l1) + Rp.l1m * (lmp - Vmp.l1 + Vpp.l1)) +
            sm * sp * (1 + tm) * (1 + tp) *
                ((Cp.ls + Vmm.l1) * (Cm.ls2 + Rm.l1m - Rm.labs) + (Cm.ls + Vpm.l1) * (Cp.l2s - Rm.l1m + Rm.labs)) +
            sm * sp * (lpm + Vmp.l1 - Vpp.l1) * (2 * tm + (1 + tm) * (1 + tp) * Rp.labs) +
            sm * sp * (1 + tm) * (1 + tp) * (Vmm.d2 - Vpm.d2 - Vmp.d2 + Vpp.d2) +
            sm * sp * (1 + tm) * (1 + tp) * (F_tp + F_tm));
    // NB: the reference's quadrature fallback for a negative alpha_tu assigns to a shadowing local
    // (nuSIprop.hpp:1406) and therefore has no effect; a negative value is kept.
  } else {
    a_tu = 0.;
  }
  a_tu *= U2;
  tot += wgt * a_tu;

  /* s-t interference */
  if (P.majorana) {
    const double dS = 2 * (tm - tp) + (Rm.l1m - Rp.l1m);
    a_st = g4 / (32 * kPi * (1 + sqr(gr)) * m4) *
           (2 * gr * (Vmm.zr_im - Vmm.zc_im - Vpm.zr_im + Vpm.zc_im - Vmp.zr_im + Vmp.zc_im + Vpp.zr_im - Vpp.zc_im) -
            2 * (Vmm.zr_re - Vmm.zc_re - Vpm.zr_re + Vpm.zc_re - Vmp.zr_re + Vmp.zc_re + Vpp.zr_re - Vpp.zc_re) +
            2 * gr * (Rm.argn - Vmm.ca) * Vmm.l1 - 2 * gr * (Rm.argn - Vpm.ca) * Vpm.l1 +
            2 * gr * (Rp.argn - Vpp.ca) * Vpp.l1 - 2 * gr * (Rp.argn - Vmp.ca) * Vmp.l1 +
            2 * (gr * Cm.args - gr * Cp.args + Cp.lam / 2. - Cm.lam / 2. + Cm.ls - Cp.ls) * dS +
            Vmm.l1 * (Cm.lam - Rm.lam2 - 2 * (Cm.ls - Rm.lfabs)) - Vpm.l1 * (Cp.lam - Rm.lam2 - 2 * (Cp.ls - Rm.lfabs)) -
            Vmp.l1 * (Cm.lam - Rp.lam2 - 2 * (Cm.ls - Rp.lfabs)) + Vpp.l1 * (Cp.lam - Rp.lam2 - 2 * (Cp.ls - Rp.lfabs)));
  } else {
    a_st = g4 / (32 * kPi * (1 + sqr(gr)) * m4) *
           ((2 * gr * Cm.args - 2 * gr * Cp.args + 2 * Cm.ls - 2 * Cp.ls + Cp.lam - Cm.lam) *
            (tm - tp + Rm.l1m - Rp.l1m));
  }
  a_st *= U2;
  tot += wgt * a_st;

  /* s-u interference: Majorana only */
  const double a_su = P.majorana ? a_st : 0.;
  tot += wgt * a_su;

  /* double-scalar production */
  double a_pp = 0;
  if ((sm > 4) && P.phiphi) {
    if (sm < 1e4) {
      const double delta = sp / sm;
      a_pp = g4 / m4 * fabs(phiphi_alpha_eval(pp, sm, nlog(-sm / tm) / nlog(delta) * 1.0001, log10(delta), warn));
    } else {
      a_pp = g4 / m4 * alpha_pp_asymptotic(sm, sp, tm, tp);
    }
    a_pp *= U2;
    if (P.majorana) a_pp *= 2;
    a_pp *= 2;
    if (P.majorana) a_pp *= 2;
  }
  tot += wgt * a_pp;

  // the reference tests x/(g/mphi)^4 < -1e-11 (nuSIprop.hpp:1505); same test without the divisions
  const double lim = -1e-11 * sqr(sqr(g / mphi));
  if (warn && (a_s < 0 || a_t < lim || a_u < lim || a_tu < lim || (a_st + a_s + a_t) < lim)) *warn |= NUSI_WARN_ALPHA;
}

// Majorana non-resonant table entry from compact vertices.  Same closed forms and term order as
// alpha_cell_mass_t<true>; the vertex sums Gs = G-- - G+- - G-+ + G++ and Hs (same signs) replace the
// per-field sums.  l1xy: x = s'-/s'+ (m/p), y = t-/t+ (m/p).
// Per-point factors of the closed forms (same expressions as in alpha_cell_mass_t, evaluated once per tile)
struct CellConst {
  double g4, m4, gr, c_s, c_st, lim;
  double c_t, c_tu;  // g4/(16 pi m4), g4/(32 pi m4)
  int phiphi;
};
NUSI_HD void make_cell_const(const PointParams &P, CellConst &K) {
  K.g4 = sqr(sqr(P.g));
  K.m4 = sqr(sqr(P.mphi));
  K.gr = P.Ga / P.mphi;
  K.c_s = K.g4 / (8 * kPi * P.Ga * cub(P.mphi));
  K.c_st = K.g4 / (32 * kPi * (1 + sqr(K.gr)) * K.m4);
  K.lim = -1e-11 * sqr(sqr(P.g / P.mphi));
  K.c_t = K.g4 / (16 * kPi * K.m4);
  K.c_tu = K.g4 / (32 * kPi * K.m4);
  K.phiphi = P.phiphi;
}
// wgt = mphi^4/(2 m_k), U2 = |U_fk|^2 of the target eigenstate.  RECIP: the quotients by edge quantities (1+t, -1+t, s) are
// taken as products with the per-edge reciprocals (8 divisions per entry and mass become multiplications).
template <bool RECIP>
NUSI_HD void alpha_cell_mass_compact(const CellConst &K, const PhiPhiTables *pp, double wgt, double U2, double fs,
                                     const CellLogs &L, const RowEdge &Rm, const RowEdge &Rp, const ColEdge &Cm,
                                     const ColEdge &Cp, double l1mm, double l1pm, double l1mp, double l1pp, double Gs,
                                     double Hs, double &tot, int *warn) {
  const double g4 = K.g4, m4 = K.m4, gr = K.gr;
  const double tm = Rm.t, tp = Rp.t, sm = Cm.s, sp = Cp.s;

  double a_s = K.c_s * (tm - tp) * fs;
  a_s *= U2;
  tot += wgt * a_s;

  const double LR1 = L.LR1, LR2 = L.LR2, LYm = L.LYm, LYp = L.LYp;
  const double N1 = sm * sp * (-tm + tp) * Cm.ls + sm * sp * (tm - tp) * Cp.ls - sm * sp * l1mm - sm * sp * tp * l1mm +
                    sm * sp * l1pm + sm * sp * tp * l1pm - sp * LR1 - sp * tm * LR1 - sp * tp * LR1 - sp * tm * tp * LR1 +
                    sm * sp * l1mp + sm * sp * tm * l1mp + sm * LR2 + sm * tm * LR2 + sm * tp * LR2 + sm * tm * tp * LR2 -
                    sm * sp * l1pp - sm * sp * tm * l1pp;
  double a_t;
  if (RECIP)
    a_t = K.c_t * (Cm.is * Cp.is) *
          (-((sm - sp) * (3 + 2 * tm * (-1 + tp) - 2 * tp) * (tm - tp)) * (Rm.i1m * Rp.i1m) + 2 * (N1) * (Rm.i1p * Rp.i1p) -
           ((sm * sp * LYm) * sqr(Rm.i1p) + (((sm - sp) * (tm - tp) * (1 + tp)) * Rm.i1p - sm * sp * LYp) * sqr(Rp.i1p)));
  else
    a_t = g4 / (sm * sp * 16 * kPi * m4) *
          (-((sm - sp) * (3 + 2 * tm * (-1 + tp) - 2 * tp) * (tm - tp)) / ((-1 + tm) * (-1 + tp)) +
           2 * (N1) / ((1 + tm) * (1 + tp)) -
           ((sm * sp * LYm) / sqr(1 + tm) + (((sm - sp) * (tm - tp) * (1 + tp)) / (1 + tm) - sm * sp * LYp) / sqr(1 + tp)));
  if (a_t < 0) a_t = alpha_gl2d<0>(tp, tm, sm, sp) * (g4 / (16 * kPi * m4));
  a_t *= U2;
  tot += wgt * a_t;
  tot += wgt * a_t;  // u channel = t channel for Majorana

  const double lmp = L.lmp, lpm = L.lpm;
  const double pre_tu = RECIP ? K.c_tu * (Cm.is * Cp.is) * (Rm.i1p * Rp.i1p) : g4 / (32 * kPi * m4 * sm * sp * (1 + tm) * (1 + tp));
  double a_tu = pre_tu *
                (-4 * (sm - sp) * (1 + tm) * (tm - tp) * (1 + tp) + 2 * sm * sp * tp * (lmp - l1mm + l1pm) +
                 2 * sp * (1 + tm) * (1 + tp) * (Rm.l1m - l1mm - Rp.l1m + l1mp) -
                 2 * sm * (1 + tm) * (1 + tp) * (Rm.l1m - l1pm - Rp.l1m + l1pp) + 2 * sm * sp * (-l1mm + l1pm + l1mp - l1pp) +
                 sm * sp * (1 + tm) * (1 + tp) *
                     (Cm.l2s * (Cp.ls + l1mp) - Cp.l2s * (Cm.ls + l1pp) + Rp.l1m * (lmp - l1mp + l1pp)) +
                 sm * sp * (1 + tm) * (1 + tp) *
                     ((Cp.ls + l1mm) * (Cm.ls2 + Rm.l1m - Rm.labs) + (Cm.ls + l1pm) * (Cp.l2s - Rm.l1m + Rm.labs)) +
                 sm * sp * (lpm + l1mp - l1pp) * (2 * tm + (1 + tm) * (1 + tp) * Rp.labs) + sm * sp * (1 + tm) * (1 + tp) * Hs);
  a_tu *= U2;
  tot += wgt * a_tu;

  const double dS = 2 * (tm - tp) + (Rm.l1m - Rp.l1m);
  double a_st = K.c_st * (Gs + 2 * (gr * Cm.args - gr * Cp.args + Cp.lam / 2. - Cm.lam / 2. + Cm.ls - Cp.ls) * dS);
  a_st *= U2;
  tot += wgt * a_st;
  tot += wgt * a_st;  // s-u interference = s-t for Majorana

  double a_pp = 0;
  if ((sm > 4) && K.phiphi) {
    if (sm < 1e4) {
      const double delta = sp / sm;
      a_pp = g4 / m4 * fabs(phiphi_alpha_eval(pp, sm, nlog(-sm / tm) / nlog(delta) * 1.0001, log10(delta), warn));
    } else {
      a_pp = g4 / m4 * alpha_pp_asymptotic(sm, sp, tm, tp);
    }
    a_pp *= U2;
    a_pp *= 8;  // x2 targets, x2 neutrinos, x2 Majorana (nuSIprop.hpp:1494-1501)
  }
  tot += wgt * a_pp;

  const double lim = K.lim;
  if (warn && (a_s < 0 || a_t < lim || a_tu < lim || (a_st + a_s + a_t) < lim)) *warn |= NUSI_WARN_ALPHA;
}

// ---- compact cell, split for parameter scans --------------------------------------------------
// alpha_cell_mass_compact<true>() cut the same way: the brackets of the t channel (X) and of the t-u
// interference (Y) do not contain g; the prefactors, the s channel and the s-t interference do.  Same
// operations, operands and order as alpha_cell_mass_compact<true> (phiphi off), so X/Y computed once per
// (mphi, m_nu) group and combined per g reproduce it bit for bit.
struct CellShared {
  double X, Y;
  double GL;  // the t channel's Gauss-Legendre fallback integral (nuSIprop.hpp:1288-1303), evaluated only where X < 0:
              // a_t = prefactor * X can be negative (prefactor >= 0) only there
};
NUSI_HD void alpha_cell_shared(const CellLogs &L, const RowEdge &Rm, const RowEdge &Rp, const ColEdge &Cm,
                               const ColEdge &Cp, double l1mm, double l1pm, double l1mp, double l1pp, double Hs,
                               CellShared &O) {
  const double tm = Rm.t, tp = Rp.t, sm = Cm.s, sp = Cp.s;
  const double LR1 = L.LR1, LR2 = L.LR2, LYm = L.LYm, LYp = L.LYp;
  const double N1 = sm * sp * (-tm + tp) * Cm.ls + sm * sp * (tm - tp) * Cp.ls - sm * sp * l1mm - sm * sp * tp * l1mm +
                    sm * sp * l1pm + sm * sp * tp * l1pm - sp * LR1 - sp * tm * LR1 - sp * tp * LR1 - sp * tm * tp * LR1 +
                    sm * sp * l1mp + sm * sp * tm * l1mp + sm * LR2 + sm * tm * LR2 + sm * tp * LR2 + sm * tm * tp * LR2 -
                    sm * sp * l1pp - sm * sp * tm * l1pp;
  O.X = (-((sm - sp) * (3 + 2 * tm * (-1 + tp) - 2 * tp) * (tm - tp)) * (Rm.i1m * Rp.i1m) + 2 * (N1) * (Rm.i1p * Rp.i1p) -
         ((sm * sp * LYm) * sqr(Rm.i1p) + (((sm - sp) * (tm - tp) * (1 + tp)) * Rm.i1p - sm * sp * LYp) * sqr(Rp.i1p)));
  const double lmp = L.lmp, lpm = L.lpm;
  O.Y = (-4 * (sm - sp) * (1 + tm) * (tm - tp) * (1 + tp) + 2 * sm * sp * tp * (lmp - l1mm + l1pm) +
         2 * sp * (1 + tm) * (1 + tp) * (Rm.l1m - l1mm - Rp.l1m + l1mp) -
         2 * sm * (1 + tm) * (1 + tp) * (Rm.l1m - l1pm - Rp.l1m + l1pp) + 2 * sm * sp * (-l1mm + l1pm + l1mp - l1pp) +
         sm * sp * (1 + tm) * (1 + tp) *
             (Cm.l2s * (Cp.ls + l1mp) - Cp.l2s * (Cm.ls + l1pp) + Rp.l1m * (lmp - l1mp + l1pp)) +
         sm * sp * (1 + tm) * (1 + tp) *
             ((Cp.ls + l1mm) * (Cm.ls2 + Rm.l1m - Rm.labs) + (Cm.ls + l1pm) * (Cp.l2s - Rm.l1m + Rm.labs)) +
         sm * sp * (lpm + l1mp - l1pp) * (2 * tm + (1 + tm) * (1 + tp) * Rp.labs) + sm * sp * (1 + tm) * (1 + tp) * Hs);
  O.GL = (O.X < 0) ? alpha_gl2d<0>(tp, tm, sm, sp) : 0.0;
}
NUSI_HD void alpha_cell_combine(const CellConst &K, double wgt, double U2, double fs, const RowEdge &Rm,
                                const RowEdge &Rp, const ColEdge &Cm, const ColEdge &Cp, double X, double Y, double GL,
                                double Gs, double &tot, int *warn) {
  const double g4 = K.g4, m4 = K.m4, gr = K.gr;
  const double tm = Rm.t, tp = Rp.t;

  double a_s = K.c_s * (tm - tp) * fs;
  a_s *= U2;
  tot += wgt * a_s;

  double a_t = K.c_t * (Cm.is * Cp.is) * X;
  if (a_t < 0) a_t = GL * (g4 / (16 * kPi * m4));
  a_t *= U2;
  tot += wgt * a_t;
  tot += wgt * a_t;  // u channel = t channel for Majorana

  const double pre_tu = K.c_tu * (Cm.is * Cp.is) * (Rm.i1p * Rp.i1p);
  double a_tu = pre_tu * Y;
  a_tu *= U2;
  tot += wgt * a_tu;

  const double dS = 2 * (tm - tp) + (Rm.l1m - Rp.l1m);
  double a_st = K.c_st * (Gs + 2 * (gr * Cm.args - gr * Cp.args + Cp.lam / 2. - Cm.lam / 2. + Cm.ls - Cp.ls) * dS);
  a_st *= U2;
  tot += wgt * a_st;
  tot += wgt * a_st;  // s-u interference = s-t for Majorana

  const double a_pp = 0;  // grouped builder runs with phiphi off only
  tot += wgt * a_pp;

  const double lim = K.lim;
  if (warn && (a_s < 0 || a_t < lim || a_tu < lim || (a_st + a_s + a_t) < lim)) *warn |= NUSI_WARN_ALPHA;
}

NUSI_HD void alpha_cell_mass(const PointParams &P, const PhiPhiTables *pp, int k, double fs, const RowEdge &Rm,
                             const RowEdge &Rp, const ColEdge &Cm, const ColEdge &Cp, const Vertex &Vmm,
                             const Vertex &Vpm, const Vertex &Vmp, const Vertex &Vpp, double &tot, int *warn) {
  CellLogs L;
  alpha_cell_mass_t<false>(P, pp, k, fs, L, Rm, Rp, Cm, Cp, Vmm, Vpm, Vmp, Vpp, tot, warn);
}

// One table entry computed from scratch (no sharing): used by the straightforward kernel and
// by the host-side checks.  Em..Ep is the target bin (row), Emp..Epp the source bin (column).
NUSI_HD double alpha_entry(const PointParams &P, const PhiPhiTables *pp, double Em, double Ep, double Emp, double Epp,
                           int *warn) {
  double tot = 0;
  const double gr = P.Ga / P.mphi;
  const bool full = P.non_resonant && P.majorana;
  for (int k = 0; k < 3; ++k) {
    RowEdge Rm, Rp;
    ColEdge Cm, Cp;
    Vertex Vmm, Vpm, Vmp, Vpp;
    const double tm = t_of(P.mn[k], Em, P.mphi), tp = t_of(P.mn[k], Ep, P.mphi);
    const double sm = s_of(P.mn[k], Emp, P.mphi), sp = s_of(P.mn[k], Epp, P.mphi);
    const double fs = alpha_s_colfactor(P, sm, sp);
    if (P.non_resonant) {
      make_row_edge(tm, gr, full, Rm);
      make_row_edge(tp, gr, full, Rp);
      make_col_edge(sm, gr, full, Cm);
      make_col_edge(sp, gr, full, Cp);
      if (full) {
        make_vertex(sm, tm, gr, Vmm);
        make_vertex(sp, tm, gr, Vpm);
        make_vertex(sm, tp, gr, Vmp);
        make_vertex(sp, tp, gr, Vpp);
      }
    } else {
      Rm.t = tm;
      Rp.t = tp;
      Cm.s = sm;
      Cp.s = sp;
    }
    alpha_cell_mass(P, pp, k, fs, Rm, Rp, Cm, Cp, Vmm, Vpm, Vmp, Vpp, tot, warn);
  }
  return tot;
}

// same, through the shared-logarithm vertex (what the default tiled kernel computes)
NUSI_HD double alpha_entry_shared(const PointParams &P, const PhiPhiTables *pp, double Em, double Ep, double Emp,
                                  double Epp, int *warn) {
  double tot = 0;
  const double gr = P.Ga / P.mphi;
  for (int k = 0; k < 3; ++k) {
    RowEdge Rm, Rp;
    ColEdge Cm, Cp;
    Vertex Vmm, Vpm, Vmp, Vpp;
    const double tm = t_of(P.mn[k], Em, P.mphi), tp = t_of(P.mn[k], Ep, P.mphi);
    const double sm = s_of(P.mn[k], Emp, P.mphi), sp = s_of(P.mn[k], Epp, P.mphi);
    const double fs = alpha_s_colfactor(P, sm, sp);
    make_row_edge(tm, gr, true, Rm);
    make_row_edge(tp, gr, true, Rp);
    make_col_edge(sm, gr, true, Cm);
    make_col_edge(sp, gr, true, Cp);
    make_vertex_shared(Rm, Cm, gr, Vmm);
    make_vertex_shared(Rm, Cp, gr, Vpm);
    make_vertex_shared(Rp, Cm, gr, Vmp);
    make_vertex_shared(Rp, Cp, gr, Vpp);
    CellLogs L;
    alpha_cell_mass_t<false>(P, pp, k, fs, L, Rm, Rp, Cm, Cp, Vmm, Vpm, Vmp, Vpp, tot, warn);
  }
  return tot;
}

// same, through compact vertices (what the default tiled kernel computes)
template <bool RECIP>
NUSI_HD double alpha_entry_compact(const PointParams &P, const PhiPhiTables *pp, double Em, double Ep, double Emp,
                                   double Epp, int *warn) {
  double tot = 0;
  const double gr = P.Ga / P.mphi;
  CellConst K;
  make_cell_const(P, K);
  for (int k = 0; k < 3; ++k) {
    RowEdge Rm, Rp;
    ColEdge Cm, Cp;
    VertexC Vmm, Vpm, Vmp, Vpp;
    const double tm = t_of(P.mn[k], Em, P.mphi), tp = t_of(P.mn[k], Ep, P.mphi);
    const double sm = s_of(P.mn[k], Emp, P.mphi), sp = s_of(P.mn[k], Epp, P.mphi);
    make_row_edge(tm, gr, true, Rm);
    make_row_edge(tp, gr, true, Rp);
    make_col_edge(sm, gr, true, Cm);
    make_col_edge(sp, gr, true, Cp);
    make_vertex_compact(Rm, Cm, gr, Vmm);
    make_vertex_compact(Rm, Cp, gr, Vpm);
    make_vertex_compact(Rp, Cm, gr, Vmp);
    make_vertex_compact(Rp, Cp, gr, Vpp);
    CellLogs L;
    L.lmp = nlog(sm / sp);
    L.lpm = nlog(sp / sm);
    L.LR1 = alpha_log_LR_c(sm, tm, tp);
    L.LR2 = alpha_log_LR_c(sp, tm, tp);
    L.LYm = alpha_log_LY_c(sm, sp, tm);
    L.LYp = alpha_log_LY_c(sm, sp, tp);
    alpha_cell_mass_compact<RECIP>(K, pp, K.m4 / (2 * P.mn[k]), P.U2f[k], alpha_s_colfactor(P, sm, sp), L, Rm, Rp, Cm, Cp, Vmm.l1, Vpm.l1, Vmp.l1, Vpp.l1,
                            Vmm.G - Vpm.G - Vmp.G + Vpp.G, Vmm.H - Vpm.H - Vmp.H + Vpp.H, tot, warn);
  }
  return tot;
}

// same entry through the split (shared / per-g) vertex and cell functions: what the grouped scan builder computes
NUSI_HD double alpha_entry_grouped(const PointParams &P, double Em, double Ep, double Emp, double Epp, int *warn) {
  double tot = 0;
  const double gr = P.Ga / P.mphi;
  CellConst K;
  make_cell_const(P, K);
  for (int k = 0; k < 3; ++k) {
    RowEdge Rm, Rp;
    ColEdge Cm, Cp;
    VertexS Vmm, Vpm, Vmp, Vpp;
    const double tm = t_of(P.mn[k], Em, P.mphi), tp = t_of(P.mn[k], Ep, P.mphi);
    const double sm = s_of(P.mn[k], Emp, P.mphi), sp = s_of(P.mn[k], Epp, P.mphi);
    make_row_edge(tm, gr, true, Rm);
    make_row_edge(tp, gr, true, Rp);
    make_col_edge(sm, gr, true, Cm);
    make_col_edge(sp, gr, true, Cp);
    make_vertex_compact_shared(Rm, Cm, Vmm);
    make_vertex_compact_shared(Rm, Cp, Vpm);
    make_vertex_compact_shared(Rp, Cm, Vmp);
    make_vertex_compact_shared(Rp, Cp, Vpp);
    CellLogs L;
    L.lmp = nlog(sm / sp);
    L.lpm = nlog(sp / sm);
    L.LR1 = alpha_log_LR_c(sm, tm, tp);
    L.LR2 = alpha_log_LR_c(sp, tm, tp);
    L.LYm = alpha_log_LY_c(sm, sp, tm);
    L.LYp = alpha_log_LY_c(sm, sp, tp);
    CellShared O;
    alpha_cell_shared(L, Rm, Rp, Cm, Cp, Vmm.l1, Vpm.l1, Vmp.l1, Vpp.l1, Vmm.H - Vpm.H - Vmp.H + Vpp.H, O);
    const double Gmm = make_vertex_compact_G(Rm, Cm, gr, Vmm.l1, Vmm.zr_re), Gpm = make_vertex_compact_G(Rm, Cp, gr, Vpm.l1, Vpm.zr_re);
    const double Gmp = make_vertex_compact_G(Rp, Cm, gr, Vmp.l1, Vmp.zr_re), Gpp = make_vertex_compact_G(Rp, Cp, gr, Vpp.l1, Vpp.zr_re);
    alpha_cell_combine(K, K.m4 / (2 * P.mn[k]), P.U2f[k], alpha_s_colfactor(P, sm, sp), Rm, Rp, Cm, Cp, O.X, O.Y, O.GL,
                       Gmm - Gpm - Gmp + Gpp, tot, warn);
  }
  return tot;
}

}  // namespace nusi
#endif
