// nusi_gamma.h -- bin-integrated absorption rate  Gamma(Em, Ep)  of one table row.
// Restates calculate_flux::Gamma (reference nuSIprop.hpp:759-922): per target mass j the
// s channel, t+u, the t-u, s-t and s-u interferences and double-scalar production, with the
// reference's Taylor switches (splus < 1e-5) and its Gauss-Legendre fallbacks when a closed
// form comes out negative.
#ifndef NUSI_GAMMA_H
#define NUSI_GAMMA_H
#include "nusi_aux.h"

namespace nusi {

// integrands of the three fallbacks (nuSIprop.hpp:809, 833, 900)
NUSI_HD double gamma_tu_sum_integrand(double z) { return (z + 2) / (z * (z + 1)) - 2 / sqr(z) * nlog1p(z); }
NUSI_HD double gamma_tu_intf_integrand(double z) {
  return 1 / z - 2 * (1 + z) / (sqr(z) * (2 + z)) * nlog1p(z);
}
NUSI_HD double gamma_pp_integrand(double z) {
  const double q = sqrt(z * (z - 4));
  return (sqr(z) - 4 * z + 6) / (sqr(z) * (z - 2)) * nlog(sqr((q + z - 2) / (q - z + 2))) - 6 * q / sqr(z);
}
template <double (*F)(double)>
NUSI_HD double gl3(double a, double b) {
  const double h = (b - a) / 2., c = (b + a) / 2.;
  const double z1 = h * gl_x(0) + c, z2 = h * gl_x(1) + c, z3 = h * gl_x(2) + c;
  return h * (gl_w(0) * F(z1) + gl_w(1) * F(z2) + gl_w(2) * F(z3));
}

// closed form of the double-scalar absorption between a (>= 4) and b  (nuSIprop.hpp:885-887;
// the reference's second copy is the same expression with the literal 4 in place of sminus)
NUSI_HD double gamma_pp_closed(double a, double b) {
  const double qa = sqrt((-4 + a) * a), qb = sqrt((-4 + b) * b);
  const double ra = sqrt((-4 + a) / a), rb = sqrt((-4 + b) / b);
  const double sa = sqrt(-4 + a), sb = sqrt(-4 + b), wa = sqrt(a), wb = sqrt(b);
  return 12 * ra - 12 * rb - 2 * nlog(sqr(sa - wa) / 4.) * nlog(sqr(-2 + a + qa) / 4.) -
         ((6 + a * nlog((-2 + a) * a)) * nlog(sqr(-2 + a + qa) / sqr(2 - a + qa))) / a -
         24 * (ra - rb - nlog(sa + wa) + nlog(sb + wb)) +
         2 * nlog(sqr(sb - wb) / 4.) * nlog(sqr(-2 + b + qb) / 4.) +
         ((6 + b * nlog((-2 + b) * b)) * nlog(sqr(-2 + b + qb) / sqr(2 - b + qb))) / b +
         8 * dilogdiff(4 / sqr(sa + wa), 4 / sqr(sb + wb)) +
         2 * dilogdiff(4 / sqr(-2 + a + qa), 4 / sqr(-2 + b + qb));
}

// Channel contributions of target mass j to Gamma(Em, Ep), in the order the reference adds them
// to its running total: ch[0..5] = s, t+u, tu, st, su, phiphi (each already weighted with
// mphi^2/(2 m_j)); returns how many are filled (1 for resonant-only).  *warn is set when the
// reference would print its "Negative cross section when computing Gamma" message.
#define NUSI_GAMMA_NCH 6
NUSI_HD int gamma_mass(const PointParams &P, int j, double Em, double Ep, double *ch, int *warn) {
  const double mphi = P.mphi, g = P.g, Ga = P.Ga;
  const double g4 = sqr(sqr(g));
  const double U2 = P.U2f[j];
  const double wgt = sqr(mphi) / (2 * P.mn[j]);
  const double splus = 2 * P.mn[j] * Ep / sqr(mphi);
  const double sminus = 2 * P.mn[j] * Em / sqr(mphi);

  /* s channel: Breit-Wigner integrated over the bin */
  const double bw = sqr(mphi) / (sqr(mphi) + sqr(Ga));
  const double dlog = nlog1p(bw * splus * (splus - 2)) - nlog1p(bw * sminus * (sminus - 2));
  double Gs;
  if (splus < 1e-5) {
    const double gr = Ga / mphi;
    Gs = g4 / (32 * kPi * sqr(mphi) * Ga) *
         (2 * mphi *
              ((gr * (1 + sqr(gr) + 2 * sminus)) / sqr(1 + sqr(gr)) * (splus - sminus) +
               (gr) / sqr(1 + sqr(gr)) * sqr(splus - sminus)) +
          Ga * dlog);
  } else {
    Gs = g4 / (32 * kPi * sqr(mphi) * Ga) *
         (2 * mphi * atandiff(mphi * (splus - 1) / Ga, mphi * (sminus - 1) / Ga) + Ga * dlog);
  }
  Gs *= U2;
  ch[0] = wgt * Gs;
  if (!P.non_resonant) return 1;

  const double l1p = nlog1p(splus), l1m = nlog1p(sminus);

  /* t + u without interference */
  double Gtu_sum = g4 / (16 * kPi * sqr(mphi)) * (2 * l1p / splus - 2 * l1m / sminus + l1p - l1m);
  if (Gtu_sum < 0) Gtu_sum = g4 / (16 * kPi * sqr(mphi)) * gl3<gamma_tu_sum_integrand>(sminus, splus);
  Gtu_sum *= 2 * U2;  // same factor for Majorana and Dirac (nuSIprop.hpp:812-815)
  ch[1] = wgt * Gtu_sum;

  /* t-u interference */
  double Gtu = g4 / (32 * kPi * sqr(mphi) * sminus * splus) *
               (sminus * l1p * (2 + 2 * splus + splus * nlog(2 + splus)) -
                splus * l1m * (2 + 2 * sminus + sminus * nlog(2 + sminus)) +
                sminus * splus * (dilog1mdiff(splus, sminus) + dilogdiff(splus, sminus)));
  if (Gtu < 0) Gtu = g4 / (16 * kPi * sqr(mphi)) * gl3<gamma_tu_intf_integrand>(sminus, splus);
  Gtu *= P.majorana ? U2 : 0.5 * U2;
  ch[2] = wgt * Gtu;

  /* s-t interference */
  const double gr = Ga / mphi;
  const cplx den = mk(gr, 2.0);                       // 2i + ga_red
  const cplx z1p = mk(0.0, 1 + splus) / den;          // i(1+s+)/(2i+ga_red)
  const cplx z1m = mk(0.0, 1 + sminus) / den;
  const cplx z2p = conj(z1p), z2m = conj(z1m);
  cplx dd1, dd2;
  if (splus < 1e-5) {
    const cplx ip = mk(gr, 1.0), im = mk(gr, -1.0);   // i+ga_red, -i+ga_red
    const cplx c1 = clog_(ip / den);
    const cplx c2 = clog_(im / mk(gr, -2.0));
    const cplx I = mk(0.0, 1.0);
    dd1 = sqr(sminus) * ((-(I / 2.0)) / ip - c1 / 2.) + sminus * c1 - splus * c1 +
          (sqr(splus) * (I / ip + c1)) / 2.;
    dd2 = sqr(sminus) * ((I / 2.0) / im - c2 / 2.) + sminus * c2 - splus * c2 +
          (sqr(splus) * ((-I) / im + c2)) / 2.;
  } else {
    dd1 = dilogdiff_complex(z1p, z1m);
    dd2 = dilogdiff_complex(z2p, z2m);
  }
  const double lam_p = nlog1p(sqr(-1 + splus) / sqr(gr)), lam_m = nlog1p(sqr(-1 + sminus) / sqr(gr));
  double Gst = -g4 / (32 * kPi * sqr(mphi) * (1 + sqr(gr))) *
               (dd1.re + dd2.re + gr * (dd2.im - dd1.im) + 2 * gr * carg_(1.0 - z2p) * l1p -
                2 * gr * carg_(1.0 - z2m) * l1m + nlog1p(4 / sqr(gr)) * (l1m - l1p) + lam_p * l1p -
                lam_m * l1m + (1 + sqr(gr)) * (lam_m - lam_p) + 2 * dilogdiff(splus, sminus));
  Gst *= U2;
  ch[3] = wgt * Gst;

  /* s-u interference: Majorana only */
  const double Gsu = P.majorana ? Gst : 0.0;
  ch[4] = wgt * Gsu;

  /* double-scalar production above threshold s = 4 mphi^2 */
  double Gpp = 0;
  if ((splus > 4) && P.phiphi) {
    const double a = (sminus > 4) ? sminus : 4.0;
    Gpp = g4 / (128. * kPi * sqr(mphi)) * gamma_pp_closed(a, splus);
    if (Gpp < 0) Gpp = g4 / (64 * kPi * sqr(mphi)) * gl3<gamma_pp_integrand>((sminus < 4) ? 4 : sminus, splus);
    Gpp *= U2;
    if (P.majorana) Gpp *= 2;
  }
  ch[5] = wgt * Gpp;

  if (warn && (Gs < 0 || Gtu_sum < 0 || Gtu < 0 || (Gs + Gtu_sum + Gst + Gsu) < 0)) *warn |= 1;
  return NUSI_GAMMA_NCH;
}

// ---- split for parameter scans (Majorana, non-resonant, phiphi off) ---------------------------------
// gamma_mass() cut into the part that does not contain g (shared by all points of a scan row with the same
// mphi and neutrino masses) and the per-point rest.  Same operations on the same operands in the same order,
// so gamma_mass_point(gamma_mass_shared()) reproduces gamma_mass() bit for bit (tests/test_host_math.py, and
// tests/test_gpu_parity.py for the kernels).  Prefactors are >= 0, so a closed form can only come out negative
// where its g-independent bracket is negative: the Gauss-Legendre fallbacks are evaluated there, once.
struct GammaShared {
  double splus, sminus, l1p, l1m;
  double B1, GL1;  // t+u bracket, its fallback integral (where B1 < 0)
  double B2, GL2;  // t-u interference bracket, its fallback
  double dd;       // dilogdiff(splus, sminus)
};
NUSI_HD void gamma_mass_shared(double mphi, double mnj, double Em, double Ep, GammaShared &S) {
  const double splus = 2 * mnj * Ep / sqr(mphi);
  const double sminus = 2 * mnj * Em / sqr(mphi);
  S.splus = splus;
  S.sminus = sminus;
  const double l1p = nlog1p(splus), l1m = nlog1p(sminus);
  S.l1p = l1p;
  S.l1m = l1m;
  S.B1 = (2 * l1p / splus - 2 * l1m / sminus + l1p - l1m);
  S.GL1 = (S.B1 < 0) ? gl3<gamma_tu_sum_integrand>(sminus, splus) : 0.0;
  S.dd = dilogdiff(splus, sminus);
  S.B2 = (sminus * l1p * (2 + 2 * splus + splus * nlog(2 + splus)) -
          splus * l1m * (2 + 2 * sminus + sminus * nlog(2 + sminus)) +
          sminus * splus * (dilog1mdiff(splus, sminus) + S.dd));
  S.GL2 = (S.B2 < 0) ? gl3<gamma_tu_intf_integrand>(sminus, splus) : 0.0;
}
NUSI_HD int gamma_mass_point(const PointParams &P, int j, const GammaShared &S, double *ch, int *warn) {
  const double mphi = P.mphi, g = P.g, Ga = P.Ga;
  const double g4 = sqr(sqr(g));
  const double U2 = P.U2f[j];
  const double wgt = sqr(mphi) / (2 * P.mn[j]);
  const double splus = S.splus, sminus = S.sminus;

  const double bw = sqr(mphi) / (sqr(mphi) + sqr(Ga));
  const double dlog = nlog1p(bw * splus * (splus - 2)) - nlog1p(bw * sminus * (sminus - 2));
  double Gs;
  if (splus < 1e-5) {
    const double gr = Ga / mphi;
    Gs = g4 / (32 * kPi * sqr(mphi) * Ga) *
         (2 * mphi *
              ((gr * (1 + sqr(gr) + 2 * sminus)) / sqr(1 + sqr(gr)) * (splus - sminus) +
               (gr) / sqr(1 + sqr(gr)) * sqr(splus - sminus)) +
          Ga * dlog);
  } else {
    Gs = g4 / (32 * kPi * sqr(mphi) * Ga) *
         (2 * mphi * atandiff(mphi * (splus - 1) / Ga, mphi * (sminus - 1) / Ga) + Ga * dlog);
  }
  Gs *= U2;
  ch[0] = wgt * Gs;

  const double l1p = S.l1p, l1m = S.l1m;

  double Gtu_sum = g4 / (16 * kPi * sqr(mphi)) * S.B1;
  if (Gtu_sum < 0) Gtu_sum = g4 / (16 * kPi * sqr(mphi)) * S.GL1;
  Gtu_sum *= 2 * U2;
  ch[1] = wgt * Gtu_sum;

  double Gtu = g4 / (32 * kPi * sqr(mphi) * sminus * splus) * S.B2;
  if (Gtu < 0) Gtu = g4 / (16 * kPi * sqr(mphi)) * S.GL2;
  Gtu *= U2;  // Majorana
  ch[2] = wgt * Gtu;

  const double gr = Ga / mphi;
  const cplx den = mk(gr, 2.0);
  const cplx z1p = mk(0.0, 1 + splus) / den;
  const cplx z1m = mk(0.0, 1 + sminus) / den;
  const cplx z2p = conj(z1p), z2m = conj(z1m);
  cplx dd1, dd2;
  if (splus < 1e-5) {
    const cplx ip = mk(gr, 1.0), im = mk(gr, -1.0);
    const cplx c1 = clog_(ip / den);
    const cplx c2 = clog_(im / mk(gr, -2.0));
    const cplx I = mk(0.0, 1.0);
    dd1 = sqr(sminus) * ((-(I / 2.0)) / ip - c1 / 2.) + sminus * c1 - splus * c1 +
          (sqr(splus) * (I / ip + c1)) / 2.;
    dd2 = sqr(sminus) * ((I / 2.0) / im - c2 / 2.) + sminus * c2 - splus * c2 +
          (sqr(splus) * ((-I) / im + c2)) / 2.;
  } else {
    dd1 = dilogdiff_complex(z1p, z1m);
    // z2 = conj(z1): every operation of dilogdiff_complex is sign-symmetric in the imaginary part (odd atan2, the
    // sgn(Im z) continuation term), so Li2(conj z) comes out as the exact conjugate -- except on the real axis, where
    // the GSL convention (Im = -pi ln x for x >= 1) is not conjugation-symmetric: evaluated there as written.
    dd2 = (z1p.im != 0.0 && z1m.im != 0.0) ? conj(dd1) : dilogdiff_complex(z2p, z2m);
  }
  const double lam_p = nlog1p(sqr(-1 + splus) / sqr(gr)), lam_m = nlog1p(sqr(-1 + sminus) / sqr(gr));
  double Gst = -g4 / (32 * kPi * sqr(mphi) * (1 + sqr(gr))) *
               (dd1.re + dd2.re + gr * (dd2.im - dd1.im) + 2 * gr * carg_(1.0 - z2p) * l1p -
                2 * gr * carg_(1.0 - z2m) * l1m + nlog1p(4 / sqr(gr)) * (l1m - l1p) + lam_p * l1p -
                lam_m * l1m + (1 + sqr(gr)) * (lam_m - lam_p) + 2 * S.dd);
  Gst *= U2;
  ch[3] = wgt * Gst;
  const double Gsu = Gst;  // Majorana
  ch[4] = wgt * Gsu;
  const double Gpp = 0;  // phiphi off
  ch[5] = wgt * Gpp;

  if (warn && (Gs < 0 || Gtu_sum < 0 || Gtu < 0 || (Gs + Gtu_sum + Gst + Gsu) < 0)) *warn |= 1;
  return NUSI_GAMMA_NCH;
}
NUSI_HD double gamma_row_split(const PointParams &P, double Em, double Ep, int *warn) {
  double tot = 0;
  for (int j = 0; j < 3; ++j) {
    double ch[NUSI_GAMMA_NCH];
    GammaShared S;
    gamma_mass_shared(P.mphi, P.mn[j], Em, Ep, S);
    const int n = gamma_mass_point(P, j, S, ch, warn);
    for (int c = 0; c < n; ++c) tot += ch[c];
  }
  return tot;
}

// Gamma(Em, Ep) = sum over the three target masses, accumulated in mass order like the reference
NUSI_HD double gamma_row(const PointParams &P, double Em, double Ep, int *warn) {
  double tot = 0;
  for (int j = 0; j < 3; ++j) {
    double ch[NUSI_GAMMA_NCH];
    const int n = gamma_mass(P, j, Em, Ep, ch, warn);
    for (int c = 0; c < n; ++c) tot += ch[c];
  }
  return tot;
}

}  // namespace nusi
#endif
