// nusi_host.cpp -- see nusi_host.hpp.  Compiled with -ffp-contract=off so every expression
// rounds exactly as written (the oracle build uses the same flag).
// The one elementary function the host side shares with the reference's constructor / source / table-loader code is log():
// the libm-locked test build (NUSI_LOCKED_MATH) routes it to the same source as the locked oracle.
#ifdef NUSI_LOCKED_MATH
#include "nusi_locked_math.h"
#define NUSI_HOST_LOG(x) nusi_locked::log(x)
#else
#define NUSI_HOST_LOG(x) std::log(x)
#endif
#include "nusi_host.hpp"
#include "nusi_interp.h"
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstring>

namespace nusi {

static double cosmo_nd(double z) { return 4.3528e-13 * cub(1 + z); }                              // nuSIprop.hpp:573-580
static double cosmo_H(double z) { return 1.5e-33 * std::pow(0.692 + 0.308 * cub(1 + z), 0.5); }   // :582-589
static double cosmo_SFR(double z) {                                                                // :591-605
  return std::pow(std::pow(1 + z, -3.4 * 10) + std::pow((1 + z) / 5161, 0.3 * 10) + std::pow((1 + z) / 9.06, 3.5 * 10),
                  -1. / 10.);
}
static double cosmo_RSN(double z) {  // :607-616
  const double m_solar = 1.989 * 56.1;
  return cosmo_SFR(z) * 0.01 / (m_solar);
}

void build_grid(int N, double lEmin, double lEmax, double zmax, bool normal_ordering, HostGrid &G) {
  G.N = N;
  G.Emin.resize(N);
  G.Ecen.resize(N);
  G.Emax.resize(N);
  G.dE.resize(N);
  for (int i = 0; i < N; ++i) {
    G.Emin[i] = std::pow(10, lEmin + (lEmax - lEmin) * (i * 1.0) / N);
    G.Ecen[i] = std::pow(10, lEmin + (lEmax - lEmin) * (i + 0.5) / N);
    G.Emax[i] = std::pow(10, lEmin + (lEmax - lEmin) * (i + 1.0) / N);
    G.dE[i] = G.Emax[i] - G.Emin[i];
  }
  // redshift step = energy step (nuSIprop.hpp:120-128); the (int) truncation is the reference's
  G.Nz = (int)(NUSI_HOST_LOG((1 + zmax) / (1 + 0)) / NUSI_HOST_LOG(G.Emax[0] / G.Emin[0]) + 2);
  G.z.resize(G.Nz);
  for (int i = 0; i < G.Nz; ++i) G.z[i] = (1 + 0) * std::pow(G.Emax[0] / G.Emin[0], i) - 1;
  G.zmax_eff = G.z[G.Nz - 1];
  G.NT = N + G.Nz - 2;

  G.row_lo.resize(G.NT);
  G.row_hi.resize(G.NT);
  for (int i = 0; i < G.NT; ++i) {
    if (i < N) {
      G.row_lo[i] = G.Emin[i];
      G.row_hi[i] = G.Emax[i];
    } else {  // top bin blue-shifted to higher redshift
      G.row_lo[i] = G.Emin[N - 1] * (1 + G.z[i - N + 1]);
      G.row_hi[i] = G.Emax[N - 1] * (1 + G.z[i - N + 1]);
    }
  }

  const double dlogz = NUSI_HOST_LOG(1 + G.z[1]) - NUSI_HOST_LOG(1 + G.z[0]);
  const int S = G.Nz - 1;
  G.cz.resize(S);
  G.ndz.resize(S);
  G.rsn.resize(S);
  G.zsrc.resize(S);
  for (int i = 1; i < G.Nz; ++i) {
    const double zr = G.z[i - 1];
    G.cz[i - 1] = (1 + zr) * dlogz / cosmo_H(zr);
    G.ndz[i - 1] = cosmo_nd(zr) / sqr(1 + zr);
    G.rsn[i - 1] = cosmo_RSN(G.z[i]);
    G.zsrc[i - 1] = G.z[i];
  }
  {
    const double Etot = 3 * 6.24, Tnue = 6e6;
    G.lum_pref = (Etot * 120 / (6 * 7 * std::pow(M_PI, 4) * std::pow(Tnue, 2)));
  }

  // NuFIT 5.0 mixing (nuSIprop.hpp:130-163); only |U|^2 is ever used
  double t12, t13, t23, dcp;
  if (normal_ordering) {
    t12 = 33.44 * (M_PI / 180);
    t13 = 8.57 * (M_PI / 180);
    t23 = 49.0 * (M_PI / 180);
    dcp = 195.0 * (M_PI / 180);
  } else {
    t12 = 33.45 * (M_PI / 180);
    t13 = 8.61 * (M_PI / 180);
    t23 = 49.3 * (M_PI / 180);
    dcp = 286.0 * (M_PI / 180);
  }
  const double c12 = std::cos(t12), c13 = std::cos(t13), c23 = std::cos(t23);
  const double s12 = std::sin(t12), s13 = std::sin(t13), s23 = std::sin(t23);
  const std::complex<double> del(std::cos(dcp), std::sin(dcp));
  std::complex<double> U[3][3];
  U[0][0] = c12 * c13;
  U[0][1] = s12 * c13;
  U[0][2] = s13 * 1.0 / del;
  U[1][0] = -s12 * c23 - c12 * s23 * s13 * del;
  U[1][1] = c12 * c23 - s12 * s23 * s13 * del;
  U[1][2] = s23 * c13;
  U[2][0] = s12 * s23 - c12 * c23 * s13 * del;
  U[2][1] = -c12 * s23 - s12 * c23 * s13 * del;
  U[2][2] = c23 * c13;
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) G.U2[a * 3 + b] = std::norm(U[a][b]);
}

// ---- free-streaming diagnostics (nuSIprop.hpp:666-744) -----------------------------------------
// 100 panels x 3 Gauss-Legendre nodes in z.  flux_FS_E0: integrand (1+z)^-si SFR/H.  energy_FS:
// integrand Lum_times_E(z, 10^lEmin, 10^lEmax)/H with the power-law source normalised by norm_total.
template <typename F>
static double gl3_panels(double z_min, double z_max, F integrand) {
  const int n_panels = 100;  // N_integ_z, nuSIprop.hpp:550
  double res = 0;
  for (int f = 0; f < n_panels; f++) {
    const double a = z_min + f * (z_max - z_min) / n_panels;
    const double b = z_min + (f + 1.0) * (z_max - z_min) / n_panels;
    const double z1 = (b - a) / 2. * gl_x(0) + (b + a) / 2., z2 = (b - a) / 2. * gl_x(1) + (b + a) / 2.,
                 z3 = (b - a) / 2. * gl_x(2) + (b + a) / 2.;
    res += (b - a) / 2. * (gl_w(0) * integrand(z1) + gl_w(1) * integrand(z2) + gl_w(2) * integrand(z3));
  }
  return res;
}
// Power-law source of the upstream code (kept as a comment in the reference, nuSIprop.hpp:648-657), bin-integrated:
//   Lum(z, Em, Ep) = norm_total/3 SFR(z) (Ep (Ep/E0 (1+z))^-si - Em (Em/E0 (1+z))^-si) / (1 - si),
//   norm_total = norm / flux_FS_E0()  (:205), flux_FS_E0 = \int_0^zmax dz (1+z)^-si SFR/H  (:666-692).
// out[i * N + j] for source redshift zsrc[i] (i < S; the z-loop evaluates the source at z[i+1], :283) and bin j:
// the layout of the device source table.
void powerlaw_source(int N, int S, const double *zsrc, const double *Emin, const double *Emax, double zmax_eff, double si,
                     double norm, double *out) {
  const double E0 = 1e14;
  const double fs0 = gl3_panels(0, zmax_eff, [&](double zz) { return std::pow(1 + zz, -si) * cosmo_SFR(zz) / cosmo_H(zz); });
  const double norm_total = norm / fs0;
  for (int i = 0; i < S; ++i)
    for (int j = 0; j < N; ++j) {
      const double Ep = Emax[j], Em = Emin[j], zz = zsrc[i];
      out[(size_t)i * N + j] = norm_total / 3.0 * cosmo_SFR(zz) *
                               (Ep * std::pow(Ep / E0 * (1 + zz), -si) - Em * std::pow(Em / E0 * (1 + zz), -si)) / (1 - si);
    }
}

double freestream_energy(double zmax_eff, double lEmin, double lEmax, double si, double norm) {
  const double E0 = 1e14;  // pivot energy, nuSIprop.hpp:549
  const double fs0 = gl3_panels(0, zmax_eff, [&](double z) { return std::pow(1 + z, -si) * cosmo_SFR(z) / cosmo_H(z); });
  const double norm_total = norm / fs0;
  auto lum_times_E = [&](double z) {
    const double Em = std::pow(10, lEmin), Ep = std::pow(10, lEmax);
    if (std::fabs(si - 2) < 1e-5)
      return norm_total * cosmo_SFR(z) * std::pow(E0 / (1 + z), si) *
             (NUSI_HOST_LOG(Ep / Em) + (2 - si) / 2.0 * (sqr(NUSI_HOST_LOG(Ep)) - sqr(NUSI_HOST_LOG(Em))));
    return norm_total * cosmo_SFR(z) * std::pow(E0 / (1 + z), si) * (std::pow(Ep, 2 - si) - std::pow(Em, 2 - si)) / (2 - si);
  };
  return gl3_panels(0, zmax_eff, [&](double z) { return lum_times_E(z) / cosmo_H(z); });
}

// nuSIprop.hpp:638-646, expression for expression (log(exp(-x) + 1), not log1p; pow-free prefactor handed in)
static double dsnb_lum_int(double pref, double z, double E) {
  const double Tnue = 6e6;
  return pref * (-E * E * (1 + z) * NUSI_HOST_LOG(std::exp(-E * (1 + z) / Tnue) + 1) / Tnue + 2 * E * li2_neg(-std::exp(-E * (1 + z) / Tnue)) +
                 2 * Tnue * li3_neg(-std::exp(-E * (1 + z) / Tnue)) / (1 + z));
}
void dsnb_source(int N, int S, const double *zsrc, const double *rsn, double pref, const double *Emin, const double *Emax, double *out) {
  for (int i = 0; i < S; ++i)
    for (int j = 0; j < N; ++j)
      out[(size_t)i * N + j] = (dsnb_lum_int(pref, zsrc[i], Emax[j]) - dsnb_lum_int(pref, zsrc[i], Emin[j])) * rsn[i];
}

// E_FS of the supernova source: sum_i L^i(z, E(1+z)) = 3 R_SN(z) dN/dE(E(1+z)) (nuSIprop.hpp:607-626: the derivative of
// Lum_int with respect to E is dN/dE(E(1+z)) R_SN-free), so
//   E_FS = 3 \int_0^zmax dz R_SN(z)/H(z) \int_Emin^Emax dE E dN/dE(E (1+z)).
// z: the reference's 100 x 3 Gauss-Legendre panels (:694-729).  E: 3-point Gauss-Legendre on 400 panels uniform in ln E
// (the integrand E^2 dN/dE d lnE is smooth on that scale: T = 6 MeV against panels 2.9 % wide at 5 decades).
double freestream_energy_dsnb(double zmax_eff, double lEmin, double lEmax) {
  const double Etot = 3 * 6.24, Tnue = 6e6;
  auto dNdE = [&](double E) { return Etot * 120 * std::pow(E, 2) / (6 * 7 * std::pow(M_PI, 4) * std::pow(Tnue, 4) * (std::exp(E / Tnue) + 1)); };
  const double la = lEmin * NUSI_HOST_LOG(10.0), lb = lEmax * NUSI_HOST_LOG(10.0);
  auto e_integral = [&](double z) {
    const int n_panels = 400;
    double res = 0;
    for (int f = 0; f < n_panels; ++f) {
      const double a = la + f * (lb - la) / n_panels, b = la + (f + 1.0) * (lb - la) / n_panels;
      for (int q = 0; q < 3; ++q) {
        const double E = std::exp((b - a) / 2. * gl_x(q) + (b + a) / 2.);
        res += (b - a) / 2. * gl_w(q) * E * E * dNdE(E * (1 + z));  // dE E f = dlnE E^2 f
      }
    }
    return res;
  };
  return gl3_panels(0, zmax_eff, [&](double z) { return 3.0 * cosmo_RSN(z) * e_integral(z) / cosmo_H(z); });
}

double scalar_width(double mphi, double g, bool majorana) {
  if (majorana) return sqr(g) * mphi / (16.0 * M_PI);
  return sqr(g) * mphi / (8.0 * M_PI);
}

// Roots of a quartic by simultaneous (Durand-Kerner) iteration in long double.  The reference
// uses gsl_poly_complex_solve (aux.hpp:34-36); any solver works because getmL accepts at most
// one root for physical inputs -- what matters is the acceptance test below.
static void quartic_roots(const double a[5], double out_re[4], double out_im[4]) {
  typedef std::complex<long double> C;
  const int deg = 4;
  C r[4];
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
      const C step = p / d;
      r[i] -= step;
      const long double m = std::abs(step);
      if (m > mv) mv = m;
    }
    if (mv < 1e-30L * scale) break;
  }
  for (int i = 0; i < deg; ++i) {
    out_re[i] = (double)r[i].real();
    out_im[i] = (double)r[i].imag();
  }
}

bool solve_masses(double mSum, bool normal_ordering, double mn[3]) {
  const double dmqSL = 7.42e-5;
  const double dmqAT = normal_ordering ? 2.514e-3 : -2.497e-3;
  double coeff[5];
  if (dmqAT > 0) {
    coeff[0] = -(sqr(sqr(mSum)) - 2 * sqr(mSum) * (dmqSL + dmqAT) + sqr(dmqSL - dmqAT));
    coeff[1] = 4 * mSum * (sqr(mSum) - dmqSL - dmqAT);
    coeff[2] = -2 * (sqr(mSum) - dmqSL - dmqAT);
    coeff[3] = -4 * mSum;
    coeff[4] = 3;
  } else {
    coeff[0] = -(sqr(sqr(mSum)) - 2 * sqr(mSum) * (-dmqSL - 2 * dmqAT) + sqr(dmqSL));
    coeff[1] = 4 * mSum * (sqr(mSum) + dmqSL + 2 * dmqAT);
    coeff[2] = -2 * (sqr(mSum) + dmqSL + 2 * dmqAT);
    coeff[3] = -4 * mSum;
    coeff[4] = 3;
  }
  double re[4], im[4];
  quartic_roots(coeff, re, im);
  double mL = 0;
  bool found = false;
  for (int i = 0; i < 4 && !found; ++i) {
    const double ml = re[i];
    if ((im[i] < 1.0e-7) && (ml >= 0) && (mSum - ml > 1.0e-7) &&
        ((dmqAT > 0 && (sqr(mSum) - dmqAT - dmqSL - sqr(ml) - 2 * ml * mSum > 1.0e-7)) ||
         (dmqAT < 0 && (sqr(mSum) + 2 * dmqAT + dmqSL - sqr(ml) - 2 * ml * mSum > 1.0e-7)))) {
      mL = ml;
      found = true;
    }
  }
  if (!found) return false;
  if (normal_ordering) {
    mn[0] = mL;
    mn[1] = std::sqrt(dmqSL + sqr(mL));
    mn[2] = std::sqrt(dmqAT + sqr(mL));
  } else {
    mn[2] = mL;
    mn[1] = std::sqrt(sqr(mL) - dmqAT);
    mn[0] = std::sqrt(sqr(mn[1]) - dmqSL);
  }
  return true;
}

// ---------------------------------------------------------------------------------------
// phi-phi tables
// ---------------------------------------------------------------------------------------
static bool read_floats(const char *path, std::vector<float> &buf, std::string &err) {
  FILE *fp = std::fopen(path, "rb");
  if (!fp) {
    err = std::string("Error at interp: the input file ") + path + " does not exist";
    return false;
  }
  std::fseek(fp, 0, SEEK_END);
  const long sz = std::ftell(fp);
  std::fseek(fp, 0, SEEK_SET);
  buf.resize((size_t)sz / sizeof(float));
  const size_t got = std::fread(buf.data(), sizeof(float), buf.size(), fp);
  std::fclose(fp);
  if (got != buf.size()) {
    err = std::string("short read on ") + path;
    return false;
  }
  return true;
}

// Records are (x_0, ..., x_{D-1}, f) float32 with the LAST coordinate running fastest
// (interp.hpp:259-290).  The reference hard-codes the node counts (5000x100, 1000x1000x100);
// here they are recovered from the coordinate columns so that smaller tables of the same
// format also load.
static bool infer_dims(const std::vector<float> &buf, int D, int dims[3], std::string &err) {
  const size_t rec = (size_t)D + 1, nrec = buf.size() / rec;
  if (nrec == 0 || buf.size() % rec) {
    err = "phi-phi table: file size is not a whole number of records";
    return false;
  }
  size_t stride = 1;
  for (int ax = D - 1; ax >= 0; --ax) {
    size_t cnt = 1;
    while (cnt * stride < nrec && buf[(cnt * stride) * rec + ax] != buf[ax]) ++cnt;
    if (ax == 0) cnt = nrec / stride;
    dims[ax] = (int)cnt;
    stride *= cnt;
  }
  if (stride != nrec) {
    err = "phi-phi table: could not recover the node grid";
    return false;
  }
  return true;
}

std::string load_phiphi(const char *at_path, const char *a_path, double l10d, HostPhiPhi &T) {
  std::string err;
  std::vector<float> buf;
  // ---- alphaTilde: (|tbar+|, log10 delta, f)
  if (!read_floats(at_path, buf, err)) return err;
  int d2[3] = {0, 0, 0};
  if (!infer_dims(buf, 2, d2, err)) return err;
  {
    const int n0 = d2[0], n1 = d2[1];
    if (n0 < 4 || n1 < 4) return "phi-phi alphaTilde table: too few nodes";
    std::vector<double> x1(n1);
    for (int c = 0; c < n1; ++c) x1[c] = (double)buf[(size_t)c * 3 + 1];
    T.delta_in_range = interp_inside(x1.data(), n1, l10d) ? 1 : 0;
    T.at_x0.resize(n0);
    T.at_f.assign(n0, 0.0);
    for (int a = 0; a < n0; ++a) T.at_x0[a] = NUSI_HOST_LOG((double)buf[((size_t)a * n1) * 3 + 0]);
    if (T.delta_in_range) {
      AxisStencil S;
      interp_axis(x1.data(), n1, l10d, S);
      for (int a = 0; a < n0; ++a) {
        double acc = 0;
        for (int c = 0; c < S.cnt; ++c) acc += (double)buf[((size_t)a * n1 + S.i0 + c) * 3 + 2] * S.c[c];
        T.at_f[a] = acc;
      }
    }
  }
  // ---- alpha: (sbar, n, log10 delta, f)
  if (!read_floats(a_path, buf, err)) return err;
  int d3[3] = {0, 0, 0};
  if (!infer_dims(buf, 3, d3, err)) return err;
  {
    const int n0 = d3[0], n1 = d3[1], n2 = d3[2];
    if (n0 < 4 || n1 < 4 || n2 < 4) return "phi-phi alpha table: too few nodes";
    std::vector<double> x2(n2);
    for (int c = 0; c < n2; ++c) x2[c] = (double)buf[(size_t)c * 4 + 2];
    const int in2 = interp_inside(x2.data(), n2, l10d) ? 1 : 0;
    T.delta_in_range = T.delta_in_range && in2;
    T.a_x0.resize(n0);
    T.a_x1.resize(n1);
    T.a_f.assign((size_t)n0 * n1, 0.0);
    for (int a = 0; a < n0; ++a) T.a_x0[a] = NUSI_HOST_LOG((double)buf[((size_t)a * n1 * n2) * 4 + 0]);
    for (int b = 0; b < n1; ++b) T.a_x1[b] = (double)buf[((size_t)b * n2) * 4 + 1];
    if (in2) {
      AxisStencil S;
      interp_axis(x2.data(), n2, l10d, S);
      for (int a = 0; a < n0; ++a)
        for (int b = 0; b < n1; ++b) {
          double acc = 0;
          for (int c = 0; c < S.cnt; ++c)
            acc += (double)buf[(((size_t)a * n1 + b) * n2 + S.i0 + c) * 4 + 3] * S.c[c];
          T.a_f[(size_t)a * n1 + b] = acc;
        }
    }
  }
  return "";
}

}  // namespace nusi
