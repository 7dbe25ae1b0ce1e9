// nusi_alphatilde.h -- self-regeneration  alphaTilde(Em, Ep)  (incoming and outgoing energy in
// the same bin).  Restates calculate_flux::alphaTilde (reference nuSIprop.hpp:924-1235): per
// target mass k the s, t, u channels, t-u / s-t / s-u interferences and double-scalar
// production, with the reference's regime switches and its 3x3
// Gauss-Legendre fallbacks for closed forms that come out negative.
#ifndef NUSI_ALPHATILDE_H
#define NUSI_ALPHATILDE_H
#include "nusi_aux.h"
#include "nusi_interp.h"

namespace nusi {

// 2-D fallback quadrature over  y in [tplus, tminus],  x in [-y, -tplus]
// (nuSIprop.hpp:987-1003, 1014-1029, 1109-1124).  KIND: 0 = t+u (Majorana t), 1 = t only (Dirac), 2 = t-u.
template <int KIND>
NUSI_HD double atilde_gl2d(double tplus, double tminus) {
  const double a_y = tplus, b_y = tminus;
  double acc = 0;
  for (int i = 0; i < 3; ++i) {
    const double y = (b_y - a_y) / 2. * gl_x(i) + (b_y + a_y) / 2.;
    const double a_x = -y, b_x = -tplus;
    for (int j = 0; j < 3; ++j) {
      const double x = (b_x - a_x) / 2. * gl_x(j) + (b_x + a_x) / 2.;
      double F;
      if (KIND == 0)
        F = sqr(y / x) / sqr(y - 1) + sqr((-x - y) / x) / sqr((-x - y) - 1);
      else if (KIND == 1)
        F = sqr(y / x) / sqr(y - 1);
      else
        F = 2 * y * (-y - x) / sqr(x) / ((y - 1) * (-y - x - 1));
      acc += 1. / 4. * (b_y - a_y) * (b_x - a_x) * gl_w(i) * gl_w(j) * F;
    }
  }
  return acc;
}

// the one-propagator t (or u) piece shared by Majorana and Dirac (nuSIprop.hpp:980-982, 1009-1011)
NUSI_HD double atilde_t_core(double tminus, double tplus, double lm, double lp) {
  return (-2 + tminus) * (tminus - tplus) - (-1 + tminus) * (-2 + tplus) * (lm - lp);
}

// combination of four dilogarithms in the t-u interference, three regimes (nuSIprop.hpp:1075-1098)
NUSI_HD double atilde_dilog_combi(double tminus, double tplus) {
  if (-tplus < 1e-2 && -tminus < 1e-2) {
    const double d = tplus / tminus;
    return -(((-1 + d) * tplus * nlog(-2 * tplus)) / d) -
           ((-1 + d) * sqr(tplus) * (-2 + d + d * nlog(2.) + nlog(-2 / tplus) - d * nlog(-tplus))) / (2. * sqr(d)) +
           (cub(tplus) * (8 - 30 * d + 21 * sqr(d) + cub(d) - 8 * cub(d) * nlog(2.) + nlog(256.) + 8 * nlog(-tplus) -
                          8 * cub(d) * nlog(-tplus))) /
               (24. * cub(d)) +
           (sqr(sqr(tplus)) * (-32 + 56 * d - 51 * sqr(d) + 30 * cub(d) - 3 * sqr(sqr(d)) + nlog(4096.) -
                               sqr(sqr(d)) * nlog(4096.) - 12 * nlog(-tplus) + 12 * sqr(sqr(d)) * nlog(-tplus))) /
               (48. * sqr(sqr(d)));
  } else if (-tplus > 1e2 && -tminus > 1e2) {
    const double d = tplus / tminus;
    const double lr = nlog((-1 + d) / d), lt = nlog(-tplus);
    return (-2 * (-1 + d) * lr) / tplus - (2 * (-1 + nlog(-(d / ((-1 + d) * tplus))))) / sqr(tplus) +
           (-6 + 4 * d + sqr(d) - 2 * cub(d) - 8 * lr + 8 * d * lr + 2 * cub(d) * lr - 2 * sqr(sqr(d)) * lr - 6 * lt +
            6 * d * lt) /
               (3. * (-1 + d) * cub(tplus)) +
           (8 - 12 * d + 3 * sqr(d) + 12 * lr - 24 * d * lr + 12 * sqr(d) * lr + 12 * lt - 24 * d * lt +
            12 * sqr(d) * lt) /
               (3. * sqr(-1 + d) * sqr(sqr(tplus)));
  }
  return li2_real(1 + 1 / (-2 + tplus)) - li2_real((-1 + tminus) / (-2 + tplus)) +
         li2_real(1 + (1 + tminus - tplus) / tplus) - li2_real(1 + 1 / tplus);
}

// Channel contributions of target mass k to alphaTilde(Em, Ep) in the reference's order of
// accumulation: ch[0..6] = s, t, u, tu, st, su, phiphi (weighted with mphi^4/(2 m_k)); returns
// how many are filled (1 for resonant-only).
#define NUSI_ATILDE_NCH 7
NUSI_HD int alphatilde_mass(const PointParams &P, const PhiPhiTables *pp, int k, double Em, double Ep,
                            double *ch, int *warn) {
  const double mphi = P.mphi, g = P.g, Ga = P.Ga;
  const double g4 = sqr(sqr(g)), m4 = sqr(sqr(mphi));
  const double U2 = P.U2f[k];
  const double wgt = m4 / (2 * P.mn[k]);
  double tplus = -2 * P.mn[k] * Ep / sqr(mphi);
  double tminus = -2 * P.mn[k] * Em / sqr(mphi);
  // keep away from the pole at t = -1 exactly (nuSIprop.hpp:949-954)
  if (fabs(tminus + 1) < 1e-7) tminus += tminus * 1e-6;
  if (fabs(tplus + 1) < 1e-7) tplus += tplus * 1e-6;

  /* s channel */
  const double bw = sqr(mphi) / (sqr(mphi) + sqr(Ga));
  const double dlog = nlog1p(bw * tplus * (tplus + 2)) - nlog1p(bw * tminus * (tminus + 2));
  double As;
  if (fabs(tplus) < 1e-5) {
    const double gr = Ga / mphi;
    As = g4 / (16 * kPi * Ga * m4) *
         (2 * mphi * (1 + tminus) *
              (-((gr * (1 + sqr(gr) - 2 * tminus) * (-tminus + tplus)) / sqr(1 + sqr(gr))) +
               (gr * sqr(-tminus + tplus)) / sqr(1 + sqr(gr))) +
          Ga * dlog);
  } else {
    As = g4 / (16 * kPi * Ga * m4) *
         (2 * mphi * (1 + tminus) * atandiff(mphi * (1 + tminus) / Ga, mphi * (1 + tplus) / Ga) + Ga * dlog);
  }
  As *= U2;
  if (!P.majorana) As /= 2.;
  ch[0] = wgt * As;
  if (!P.non_resonant) return 1;

  const double lm = nlog1p(-tminus), lp = nlog1p(-tplus);  // ln(1-t-), ln(1-t+)
  const double ld = nlog1p(tminus - tplus);               // ln(1 + t- - t+)
  const double lrat = nlog(tminus / tplus);

  /* t channel, u channel */
  double At, Au;
  if (P.majorana) {
    At = g4 * (1 / (16 * m4 * kPi * (-1 + tminus) * tplus) * atilde_t_core(tminus, tplus, lm, lp) +
               1 / (16 * m4 * kPi * sqr(1 + tminus) * tplus) *
                   ((1 + tminus) * (2 + tminus) * (tminus - tplus) +
                    (-2 * sqr(1 + tminus) + tplus + 2 * tminus * tplus) * ld - sqr(tminus) * tplus * lrat));
    if (At < 0) At = atilde_gl2d<0>(tplus, tminus) * (g4 / (16 * kPi * m4));
    At *= U2;
    Au = At;
  } else {
    At = 3. / 2. * g4 / (32 * m4 * kPi * (-1 + tminus) * tplus) * atilde_t_core(tminus, tplus, lm, lp);
    if (At < 0) At = atilde_gl2d<1>(tplus, tminus) * (3. / 2. * g4 / (32 * kPi * m4));
    At *= U2;
    Au = 1. / 2. * g4 / (32 * m4 * kPi * (-1 + tminus) * tplus) * atilde_t_core(tminus, tplus, lm, lp);
    if (Au < 0) Au = atilde_gl2d<1>(tplus, tminus) * (1. / 2. * g4 / (32 * kPi * m4));
    Au *= U2;
  }
  ch[1] = wgt * At;
  ch[2] = wgt * Au;

  /* t-u interference (Majorana only) */
  double Atu;
  if (P.majorana) {
    const double combi = atilde_dilog_combi(tminus, tplus);
    Atu = g4 / (32 * kPi * m4 * (1 + tminus) * tplus) *
          (2 * (2 * (1 + tminus) * (tminus - tplus) -
                2 * (1 + tminus) * tplus * natanh(1 / (1 - tplus)) * natanh((tminus - tplus) / (-2 + tminus + tplus)) +
                tminus * tplus * (-lm + lp) + (1 + tminus) * (lm - lp - ld) + tplus * (-lm + lp + ld) -
                tminus * tplus * lrat) +
           (1 + tminus) * tplus * ((-sqr(lm) + sqr(lp)) / 2. + dilog1over1mdiff(tplus, tminus)) -
           (1 + tminus) * tplus * (dilog1pdiff(tminus, tplus) + combi));
    if (Atu < 0) Atu = atilde_gl2d<2>(tplus, tminus) * (g4 / (16 * kPi * m4));
  } else {
    Atu = 0;
  }
  Atu *= U2;
  ch[3] = wgt * Atu;

  /* s-t interference */
  const double gr = Ga / mphi;
  const cplx den = mk(gr, 2.0);  // 2i + ga_red
  cplx dd78, dd51, dd26, dd43;
  if (-tplus < 1e-5) {
    const double d = tplus / tminus;
    const cplx I = mk(0.0, 1.0);
    const cplx Lm = clog_(real_c(tminus)), Lp = clog_(real_c(tplus)), Ld = clog_(real_c(d));
    dd78 = tminus * (-1 + Lm) + (sqr(tminus) * (-1 + 2 * Lm)) / 4. -
           (tplus * (-1 + Lp) + (sqr(tplus) * (-1 + 2 * Lp)) / 4.);
    const cplx cL = clog_(1.0 - I / den);
    dd51 = (-tminus + tplus) * cL +
           ((-sqr(tminus) + sqr(tplus)) * (I * (1 + cL) + cL * gr)) / (2. * mk(gr, 1.0));
    dd26 = (tplus * (-1 + d - Ld + Lp - d * Lp)) / d +
           (sqr(tplus) * (-1 + sqr(d) + 2 * Ld - 2 * Lp + 4 * d * Lp - 2 * sqr(d) * Lp)) / (4. * sqr(d)) +
           (cub(tplus) * (7 - 9 * d + 2 * cub(d) - 6 * Ld + 6 * Lp - 18 * d * Lp + 18 * sqr(d) * Lp - 6 * cub(d) * Lp)) /
               (18. * cub(d));
    const cplx cM = clog_(mk(gr, 1.0) / den);
    dd43 = ((-1 + d) * tplus * cM) / d +
           ((-1 + d) * sqr(tplus) * (I * ((1 + d) / mk(gr, 1.0) - 2 / den) + (-1 + d) * cM)) / (2. * sqr(d));
  } else {
    const cplx z1 = mk(0.0, -(-1 + tminus)) / den;
    const double z2 = 1 / (1 + tminus);
    const cplx z3 = 1.0 / mk(2 + tminus, -gr);
    const cplx z4 = (1 + tminus - tplus) / mk(2 + tminus, -gr);
    const cplx z5 = mk(0.0, -(-1 + tplus)) / den;
    const double z6 = 1 - tplus / (1 + tminus);
    const double z7 = 1 - tminus, z8 = 1 - tplus;
    dd78 = dilogdiff_complex(real_c(z7), real_c(z8));
    dd51 = dilogdiff_complex(z5, z1);
    dd26 = dilogdiff_complex(real_c(z2), real_c(z6));
    dd43 = dilogdiff_complex(z4, z3);
  }
  const double arg_m = natan2(gr, -1 - tminus), arg_p = natan2(gr, -1 - tplus);  // carg(-1 + i ga_red - t)
  const double arg_qm = carg_(mk(gr, 1 + tminus) / den), arg_qp = carg_(mk(gr, 1 + tplus) / den);
  const double lam_p = nlog1p(sqr(1 + tplus) / sqr(gr)), lam_m = nlog1p(sqr(1 + tminus) / sqr(gr));
  double Ast;
  if (P.majorana) {
    const double arg_1m = (1 + tminus < 0) ? kPi : 0.0;  // carg of the real number 1 + tminus
    Ast = g4 / (32 * kPi * (1 + sqr(gr)) * m4) *
          (2 * kPi * arg_m - 2 * kPi * arg_p + 2 * gr * (dd51.im + dd26.im + dd43.im) -
           2 * (dd51.re + dd26.re + dd43.re + dd78.re) - arg_qm * (2 * kPi + 2 * gr * lm) +
           arg_qp * (2 * kPi + 2 * gr * lp) + (arg_m - arg_p) * (4 * gr * tminus + 2 * gr * lm) +
           2 * gr * (arg_1m - natan2(-gr, 2 + tminus) + natan2(-gr, 1 + tplus)) * ld +
           nlog(4 + sqr(gr)) * (lp - lm) + nlog(sqr(gr) + sqr(2 + tminus)) * ld - 2 * lm * nlog(-tplus) -
           2 * gr * kPi * (nlog(sqr(tplus)) + ld) + 2 * gr * kPi * nlog(sqr(tplus)) + 4 * tminus * lrat +
           (-lp + lm - ld) * (lam_p + 2 * nlog(gr)) - ld * nlog1p(sqr(tminus) + 2 * tminus) +
           2 * (sqr(gr) + tminus) * (lam_p - lam_m) + 2 * (nlog(-tplus) * (lp + ld) + (lam_p - lam_m)));
  } else {
    const cplx s58 = dd51 + dd78;
    Ast = g4 / (32 * kPi * (1 + sqr(gr)) * m4) *
          (gr * dd51.im - 2 * (s58.re) + 2 * arg_qm * (-kPi - gr * lm) + 2 * arg_m * (kPi + gr * tminus + gr * lm) -
           2 * arg_p * (kPi + gr * tminus + gr * lm) + 2 * arg_qp * (kPi + gr * lp) - 2 * lm * nlog(-tplus) +
           2 * tminus * lrat + 2 * lp * nlog(-tplus) + (lp - lm) * (nlog(4 + sqr(gr)) - 2 * nlog(gr) - lam_p) +
           (1 + tminus + sqr(gr)) * (lam_p - lam_m));
  }
  Ast *= U2;
  ch[4] = wgt * Ast;

  /* s-u interference: Majorana only */
  const double Asu = P.majorana ? Ast : 0.0;
  ch[5] = wgt * Asu;

  /* double-scalar production: table lookup inside 4 < -t+ < 1e4, asymptotic form above */
  double App = 0;
  if ((-tplus > 4) && P.phiphi) {
    if (-tplus < 1e4) {
      App = g4 / m4 * phiphi_alphatilde_eval(pp, -tplus, log10(tplus / tminus), warn);
    } else {
      App = g4 / m4 *
            (6 * tminus * nlog(-tminus) - tplus * sqr(nlog(-tminus)) +
             2 * (-8 * tminus + 8 * tplus + 4 * tplus * nlog(-tminus) +
                  nlog(tminus - tplus) * (tminus - tplus - tplus * lrat)) -
             2 * (2 * tminus + 5 * tplus) * nlog(-tplus) + tplus * sqr(nlog(-tplus)) -
             2 * tplus * li2_real(1 - tminus / tplus)) /
            (128. * kPi * tplus);
    }
    App *= U2;
    if (P.majorana) App *= 2;
    App *= 2;
    if (P.majorana) App *= 2;
  }
  ch[6] = wgt * App;

  const double nrm = sqr(sqr(g / mphi));
  if (warn && (As < 0 || At < 0 || Au < 0 || Atu / nrm < -1e-11 || (Ast + At + As) / nrm < -1e-11 ||
               (Asu + Au + As) / nrm < -1e-11))
    *warn |= 2;
  return NUSI_ATILDE_NCH;
}

// ---- split for parameter scans (Majorana, non-resonant, phiphi off) ---------------------------------
// alphatilde_mass() cut into the part without g and the per-point rest, as gamma_mass_shared/_point: same
// operations, operands and order, so the composition equals alphatilde_mass() bit for bit.
struct ATildeShared {
  double tminus, tplus, lm, lp, ld, lrat;
  double Bt, GLt;            // (t-channel closed form)/g4 and its fallback integral (where Bt < 0)
  double den_tu, Btu, GLtu;  // t-u interference: denominator of the prefactor, bracket, fallback integral
  double dd78_re, dd26_re, dd26_im;
  double lnmt;   // nlog(-tplus)
  double lt2;    // nlog(sqr(tplus))
  double l1p2;   // nlog1p(sqr(tminus) + 2 tminus)
};
NUSI_HD void alphatilde_mass_shared(double mphi, double mnk, double Em, double Ep, ATildeShared &S) {
  const double m4 = sqr(sqr(mphi));
  double tplus = -2 * mnk * Ep / sqr(mphi);
  double tminus = -2 * mnk * Em / sqr(mphi);
  if (fabs(tminus + 1) < 1e-7) tminus += tminus * 1e-6;
  if (fabs(tplus + 1) < 1e-7) tplus += tplus * 1e-6;
  S.tminus = tminus;
  S.tplus = tplus;
  const double lm = nlog1p(-tminus), lp = nlog1p(-tplus);
  const double ld = nlog1p(tminus - tplus);
  const double lrat = nlog(tminus / tplus);
  S.lm = lm; S.lp = lp; S.ld = ld; S.lrat = lrat;

  S.Bt = (1 / (16 * m4 * kPi * (-1 + tminus) * tplus) * atilde_t_core(tminus, tplus, lm, lp) +
          1 / (16 * m4 * kPi * sqr(1 + tminus) * tplus) *
              ((1 + tminus) * (2 + tminus) * (tminus - tplus) +
               (-2 * sqr(1 + tminus) + tplus + 2 * tminus * tplus) * ld - sqr(tminus) * tplus * lrat));
  S.GLt = (S.Bt < 0) ? atilde_gl2d<0>(tplus, tminus) : 0.0;

  const double combi = atilde_dilog_combi(tminus, tplus);
  S.den_tu = (32 * kPi * m4 * (1 + tminus) * tplus);
  S.Btu = (2 * (2 * (1 + tminus) * (tminus - tplus) -
                2 * (1 + tminus) * tplus * natanh(1 / (1 - tplus)) * natanh((tminus - tplus) / (-2 + tminus + tplus)) +
                tminus * tplus * (-lm + lp) + (1 + tminus) * (lm - lp - ld) + tplus * (-lm + lp + ld) -
                tminus * tplus * lrat) +
           (1 + tminus) * tplus * ((-sqr(lm) + sqr(lp)) / 2. + dilog1over1mdiff(tplus, tminus)) -
           (1 + tminus) * tplus * (dilog1pdiff(tminus, tplus) + combi));
  // g4/den_tu * Btu < 0 needs den_tu and Btu of opposite sign (g4 >= 0)
  S.GLtu = ((S.Btu >= 0 && S.den_tu > 0) || (S.Btu <= 0 && S.den_tu < 0)) ? 0.0 : atilde_gl2d<2>(tplus, tminus);

  cplx dd78, dd26;
  if (-tplus < 1e-5) {
    const double d = tplus / tminus;
    const cplx Lm = clog_(real_c(tminus)), Lp = clog_(real_c(tplus)), Ld = clog_(real_c(d));
    dd78 = tminus * (-1 + Lm) + (sqr(tminus) * (-1 + 2 * Lm)) / 4. -
           (tplus * (-1 + Lp) + (sqr(tplus) * (-1 + 2 * Lp)) / 4.);
    dd26 = (tplus * (-1 + d - Ld + Lp - d * Lp)) / d +
           (sqr(tplus) * (-1 + sqr(d) + 2 * Ld - 2 * Lp + 4 * d * Lp - 2 * sqr(d) * Lp)) / (4. * sqr(d)) +
           (cub(tplus) * (7 - 9 * d + 2 * cub(d) - 6 * Ld + 6 * Lp - 18 * d * Lp + 18 * sqr(d) * Lp - 6 * cub(d) * Lp)) /
               (18. * cub(d));
  } else {
    const double z2 = 1 / (1 + tminus);
    const double z6 = 1 - tplus / (1 + tminus);
    const double z7 = 1 - tminus, z8 = 1 - tplus;
    dd78 = dilogdiff_complex(real_c(z7), real_c(z8));
    dd26 = dilogdiff_complex(real_c(z2), real_c(z6));
  }
  S.dd78_re = dd78.re;
  S.dd26_re = dd26.re;
  S.dd26_im = dd26.im;
  S.lnmt = nlog(-tplus);
  S.lt2 = nlog(sqr(tplus));
  S.l1p2 = nlog1p(sqr(tminus) + 2 * tminus);
}
NUSI_HD int alphatilde_mass_point(const PointParams &P, int k, const ATildeShared &S, double *ch, int *warn) {
  const double mphi = P.mphi, g = P.g, Ga = P.Ga;
  const double g4 = sqr(sqr(g)), m4 = sqr(sqr(mphi));
  const double U2 = P.U2f[k];
  const double wgt = m4 / (2 * P.mn[k]);
  const double tplus = S.tplus, tminus = S.tminus;

  const double bw = sqr(mphi) / (sqr(mphi) + sqr(Ga));
  const double dlog = nlog1p(bw * tplus * (tplus + 2)) - nlog1p(bw * tminus * (tminus + 2));
  double As;
  if (fabs(tplus) < 1e-5) {
    const double gr = Ga / mphi;
    As = g4 / (16 * kPi * Ga * m4) *
         (2 * mphi * (1 + tminus) *
              (-((gr * (1 + sqr(gr) - 2 * tminus) * (-tminus + tplus)) / sqr(1 + sqr(gr))) +
               (gr * sqr(-tminus + tplus)) / sqr(1 + sqr(gr))) +
          Ga * dlog);
  } else {
    As = g4 / (16 * kPi * Ga * m4) *
         (2 * mphi * (1 + tminus) * atandiff(mphi * (1 + tminus) / Ga, mphi * (1 + tplus) / Ga) + Ga * dlog);
  }
  As *= U2;
  ch[0] = wgt * As;

  const double lm = S.lm, lp = S.lp, ld = S.ld, lrat = S.lrat;

  double At = g4 * S.Bt;
  if (At < 0) At = S.GLt * (g4 / (16 * kPi * m4));
  At *= U2;
  const double Au = At;
  ch[1] = wgt * At;
  ch[2] = wgt * Au;

  double Atu = g4 / S.den_tu * S.Btu;
  if (Atu < 0) Atu = S.GLtu * (g4 / (16 * kPi * m4));
  Atu *= U2;
  ch[3] = wgt * Atu;

  const double gr = Ga / mphi;
  const cplx den = mk(gr, 2.0);
  cplx dd51, dd43;
  if (-tplus < 1e-5) {
    const double d = tplus / tminus;
    const cplx I = mk(0.0, 1.0);
    const cplx cL = clog_(1.0 - I / den);
    dd51 = (-tminus + tplus) * cL +
           ((-sqr(tminus) + sqr(tplus)) * (I * (1 + cL) + cL * gr)) / (2. * mk(gr, 1.0));
    const cplx cM = clog_(mk(gr, 1.0) / den);
    dd43 = ((-1 + d) * tplus * cM) / d +
           ((-1 + d) * sqr(tplus) * (I * ((1 + d) / mk(gr, 1.0) - 2 / den) + (-1 + d) * cM)) / (2. * sqr(d));
  } else {
    const cplx z1 = mk(0.0, -(-1 + tminus)) / den;
    const cplx z3 = 1.0 / mk(2 + tminus, -gr);
    const cplx z4 = (1 + tminus - tplus) / mk(2 + tminus, -gr);
    const cplx z5 = mk(0.0, -(-1 + tplus)) / den;
    dd51 = dilogdiff_complex(z5, z1);
    dd43 = dilogdiff_complex(z4, z3);
  }
  const double arg_m = natan2(gr, -1 - tminus), arg_p = natan2(gr, -1 - tplus);
  const double arg_qm = carg_(mk(gr, 1 + tminus) / den), arg_qp = carg_(mk(gr, 1 + tplus) / den);
  const double lam_p = nlog1p(sqr(1 + tplus) / sqr(gr)), lam_m = nlog1p(sqr(1 + tminus) / sqr(gr));
  const double arg_1m = (1 + tminus < 0) ? kPi : 0.0;
  double Ast = g4 / (32 * kPi * (1 + sqr(gr)) * m4) *
               (2 * kPi * arg_m - 2 * kPi * arg_p + 2 * gr * (dd51.im + S.dd26_im + dd43.im) -
                2 * (dd51.re + S.dd26_re + dd43.re + S.dd78_re) - arg_qm * (2 * kPi + 2 * gr * lm) +
                arg_qp * (2 * kPi + 2 * gr * lp) + (arg_m - arg_p) * (4 * gr * tminus + 2 * gr * lm) +
                2 * gr * (arg_1m - natan2(-gr, 2 + tminus) + natan2(-gr, 1 + tplus)) * ld +
                nlog(4 + sqr(gr)) * (lp - lm) + nlog(sqr(gr) + sqr(2 + tminus)) * ld - 2 * lm * S.lnmt -
                2 * gr * kPi * (S.lt2 + ld) + 2 * gr * kPi * S.lt2 + 4 * tminus * lrat +
                (-lp + lm - ld) * (lam_p + 2 * nlog(gr)) - ld * S.l1p2 +
                2 * (sqr(gr) + tminus) * (lam_p - lam_m) + 2 * (S.lnmt * (lp + ld) + (lam_p - lam_m)));
  Ast *= U2;
  ch[4] = wgt * Ast;
  const double Asu = Ast;
  ch[5] = wgt * Asu;
  const double App = 0;
  ch[6] = wgt * App;

  const double nrm = sqr(sqr(g / mphi));
  if (warn && (As < 0 || At < 0 || Au < 0 || Atu / nrm < -1e-11 || (Ast + At + As) / nrm < -1e-11 ||
               (Asu + Au + As) / nrm < -1e-11))
    *warn |= 2;
  return NUSI_ATILDE_NCH;
}
NUSI_HD double alphatilde_row_split(const PointParams &P, double Em, double Ep, int *warn) {
  double tot = 0;
  for (int k = 0; k < 3; ++k) {
    double ch[NUSI_ATILDE_NCH];
    ATildeShared S;
    alphatilde_mass_shared(P.mphi, P.mn[k], Em, Ep, S);
    const int n = alphatilde_mass_point(P, k, S, ch, warn);
    for (int c = 0; c < n; ++c) tot += ch[c];
  }
  return tot;
}

NUSI_HD double alphatilde_row(const PointParams &P, const PhiPhiTables *pp, double Em, double Ep, int *warn) {
  double tot = 0;
  for (int k = 0; k < 3; ++k) {
    double ch[NUSI_ATILDE_NCH];
    const int n = alphatilde_mass(P, pp, k, Em, Ep, ch, warn);
    for (int c = 0; c < n; ++c) tot += ch[c];
  }
  return tot;
}

}  // namespace nusi
#endif
