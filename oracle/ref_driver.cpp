// TEST INFRASTRUCTURE -- the oracle.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may build, load or run this.
//
// Compiles the UNMODIFIED reference headers (/root/reference/nuSIprop.hpp, aux.hpp,
// interp.hpp, located with -I at build time; never copied into this repo) against
// the stand-ins in oracle/refshim/ for the two third-party dependencies that are
// absent (GSL 2.7, polylogarithm f21dea8).  Outputs go to oracle/_ref/ only.
//
// Built twice from this one file (oracle/Makefile):
//   oracle/_ref/libnusi_ref.so  C entry points for ctypes (tests)
//   oracle/_ref/nusi_ref        CLI, one process per parameter point (timing,
//                               %.17g dumps)
//
// `#define private public` exposes the reference's private members so that the
// cross-section tables (nuSIprop.hpp:218-253) can be dumped entry by entry;
// libstdc++ headers are pre-included first because they break under that macro.
#include <sstream>
#include <complex>
#include <iostream>
#include <cmath>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <string>
#include <chrono>
#include <unistd.h>
#ifdef NUSI_REF_ULP
// Noise-floor ensemble member: the same unmodified reference, but every
// result of log/log1p/exp/atan/atanh/carg/clog/cabs that the reference's own code obtains from libm is
// moved by +-1 ulp for a deterministic quarter of the argument values.  How far the fluxes move under
// this perturbation is how far two conforming libms can pull the reference apart.  The stand-in headers
// are included BEFORE the macros so that only the reference's own call sites are affected.
#include <gsl/gsl_linalg.h>
#include <gsl/gsl_sf_dilog.h>
#include <gsl/gsl_poly.h>
#include "polylogarithm/src/cpp/Li2.hpp"
#include "polylogarithm/src/cpp/Li3.hpp"
#include <complex.h>
#include <stdint.h>
#ifndef NUSI_REF_ULP_SEED
#define NUSI_REF_ULP_SEED 0  // other seeds = other members of the ensemble (which quarter of the values moves, and which way)
#endif
static inline double nusi_ulp(double v) {
  uint64_t b;
  std::memcpy(&b, &v, 8);
  uint64_t h = (b ^ ((uint64_t)NUSI_REF_ULP_SEED * 0xD6E8FEB86659FD93ull)) * 0x9E3779B97F4A7C15ull;
  h ^= h >> 29;
  const unsigned r = (unsigned)(h >> 40) & 7u;
  if (r == 0) return nextafter(v, INFINITY);
  if (r == 1) return nextafter(v, -INFINITY);
  return v;
}
static inline double _Complex nusi_ulpc(double _Complex v) { return nusi_ulp(creal(v)) + I * nusi_ulp(cimag(v)); }
#define log(x) nusi_ulp(log(x))
#define log1p(x) nusi_ulp(log1p(x))
#define exp(x) nusi_ulp(exp(x))
#define atan(x) nusi_ulp(atan(x))
#define atanh(x) nusi_ulp(atanh(x))
#define carg(x) nusi_ulp(carg(x))
#define cabs(x) nusi_ulp(cabs(x))
#define clog(x) nusi_ulpc(clog(x))
#endif
#ifdef NUSI_REF_LOCKED
// The libm-locked oracle O2: the same unmodified reference, with every log / log1p / atan / atanh /
// carg / cabs / clog of its own code routed to nusiprop_b200/csrc/nusi_locked_math.h (IEEE basic operations only: the
// same bits on the host and on the GPU) and the two GSL dilogarithms served by the product's Li2 (REFSHIM_SHARED_LI2).
// Compared with the product's own locked build, what remains is the product's arithmetic, not libm noise.
#include <gsl/gsl_linalg.h>
#include <gsl/gsl_sf_dilog.h>
#include <gsl/gsl_poly.h>
#include "polylogarithm/src/cpp/Li2.hpp"
#include "polylogarithm/src/cpp/Li3.hpp"
#include <complex.h>
#include "../nusiprop_b200/csrc/nusi_locked_math.h"
static inline double _Complex nusi_lk_clog(double _Complex v) {
  return nusi_locked::log(nusi_locked::hypot(creal(v), cimag(v))) + I * nusi_locked::atan2(cimag(v), creal(v));
}
#define log(x) nusi_locked::log(x)
#define log1p(x) nusi_locked::log1p(x)
#define atan(x) nusi_locked::atan(x)
#define atanh(x) nusi_locked::atanh(x)
#define carg(x) nusi_locked::atan2(cimag(x), creal(x))
#define cabs(x) nusi_locked::hypot(creal(x), cimag(x))
#define clog(x) nusi_lk_clog(x)
#endif
#define private public
#include "nuSIprop.hpp"
#undef private
#ifdef NUSI_REF_LOCKED
#undef log
#undef log1p
#undef atan
#undef atanh
#undef carg
#undef cabs
#undef clog
#endif
#ifdef NUSI_REF_ULP
#undef log
#undef log1p
#undef exp
#undef atan
#undef atanh
#undef carg
#undef cabs
#undef clog
#endif

using nuSIprop::calculate_flux;

extern "C" {

// evolve() of the reference: nuSIprop.hpp:176-337.  flux_mass/flux_fla: [3][N] row-major.
int nusi_ref_evolve(double mphi, double g, double mntot, double si, double norm, int majorana,
                    int non_resonant, int normal_ordering, int N_bins_E, double lEmin, double lEmax,
                    double zmax, int flav, int phiphi, double *flux_mass, double *flux_fla,
                    double *energies) {
  calculate_flux ev(mphi, g, mntot, si, norm, majorana != 0, non_resonant != 0, normal_ordering != 0,
                    N_bins_E, lEmin, lEmax, zmax, flav, phiphi != 0);
  ev.evolve();
  for (int k = 0; k < 3; ++k)
    for (int j = 0; j < N_bins_E; ++j) {
      if (flux_mass) flux_mass[k * N_bins_E + j] = ev.get_flux(k, j);
      if (flux_fla) flux_fla[k * N_bins_E + j] = ev.get_flux_fla(k, j);
    }
  if (energies)
    for (int j = 0; j < N_bins_E; ++j) energies[j] = ev.get_energy(j);
  return 0;
}

// Sizes and grids of a configuration: nuSIprop.hpp:113-128.
int nusi_ref_grid(int N_bins_E, double lEmin, double lEmax, double zmax, int *N_steps_z, double *Emin,
                  double *Emax, double *Ecen, double *z /* >= N_steps_z, pass NULL to query */) {
  calculate_flux ev(1e2, 1e-2, 0.1, 2, 1, true, false, true, N_bins_E, lEmin, lEmax, zmax, 2, false);
  *N_steps_z = ev.N_steps_z;
  for (int j = 0; j < N_bins_E; ++j) {
    if (Emin) Emin[j] = ev.Emin[j];
    if (Emax) Emax[j] = ev.Emax[j];
    if (Ecen) Ecen[j] = ev.E_nu[j];
  }
  if (z)
    for (int i = 0; i < ev.N_steps_z; ++i) z[i] = ev.z[i];
  return 0;
}

static void set_masses(calculate_flux &ev) {  // nuSIprop.hpp:184-203
  double dmq21 = 7.42e-5;
  double dmqAT = ev.normal_ordering ? 2.514e-3 : -2.497e-3;
  double mL = nuSIaux::getmL(ev.mntot, dmq21, dmqAT);
  if (ev.normal_ordering) {
    ev.mn[0] = mL;
    ev.mn[1] = sqrt(dmq21 + SQR(mL));
    ev.mn[2] = sqrt(dmqAT + SQR(mL));
  } else {
    ev.mn[2] = mL;
    ev.mn[1] = sqrt(SQR(mL) - dmqAT);
    ev.mn[0] = sqrt(SQR(ev.mn[1]) - dmq21);
  }
}

int nusi_ref_masses(double mntot, int normal_ordering, double *mn) {
  calculate_flux ev(1e2, 1e-2, mntot, 2, 1, true, false, normal_ordering != 0, 10, 4, 9, 5, 2, false);
  set_masses(ev);
  for (int k = 0; k < 3; ++k) mn[k] = ev.mn[k];
  return 0;
}

int nusi_ref_pmns(int normal_ordering, double *U2 /*[3][3] |U_ab|^2*/) {
  calculate_flux ev(1e2, 1e-2, 0.1, 2, 1, true, false, normal_ordering != 0, 10, 4, 9, 5, 2, false);
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) U2[a * 3 + b] = std::norm(ev.U[a][b]);
  return 0;
}

// The three tables of nuSIprop.hpp:218-253.  tbl_alpha is returned dense NT x NT
// row-major with only m > i filled (others 0).  alpha==NULL skips the O(NT^2) part.
int nusi_ref_tables(double mphi, double g, double mntot, int majorana, int non_resonant,
                    int normal_ordering, int N_bins_E, double lEmin, double lEmax, double zmax, int flav,
                    int phiphi, double *Gamma, double *alphaTilde, double *alpha) {
  calculate_flux ev(mphi, g, mntot, 2, 1, majorana != 0, non_resonant != 0, normal_ordering != 0, N_bins_E,
                    lEmin, lEmax, zmax, flav, phiphi != 0);
  set_masses(ev);
  const int N = N_bins_E, NT = N + ev.N_steps_z - 2;
  std::vector<double> lo(NT), hi(NT);
  for (int i = 0; i < NT; ++i) {
    if (i < N) { lo[i] = ev.Emin[i]; hi[i] = ev.Emax[i]; }
    else { lo[i] = ev.Emin[N - 1] * (1 + ev.z[i - N + 1]); hi[i] = ev.Emax[N - 1] * (1 + ev.z[i - N + 1]); }
  }
  for (int i = 0; i < NT; ++i) {
    if (Gamma) Gamma[i] = ev.Gamma(lo[i], hi[i]);
    if (alphaTilde) alphaTilde[i] = ev.alphaTilde(lo[i], hi[i]);
    if (alpha)
      for (int m = 0; m < NT; ++m)
        alpha[(size_t)i * NT + m] = (m > i) ? ev.alpha(lo[i], hi[i], lo[m], hi[m]) : 0.0;
  }
  return 0;
}

// Source term Lum(z[i], Emin[j], Emax[j]) of nuSIprop.hpp:659-662 for i=1..N_z-1, j=0..N-1
// -> out[(i-1)*N + j]; and H(z), nd(z) helpers (nuSIprop.hpp:573-589).
int nusi_ref_source(int N_bins_E, double lEmin, double lEmax, double zmax, double *out) {
  calculate_flux ev(1e2, 1e-2, 0.1, 2, 1, true, false, true, N_bins_E, lEmin, lEmax, zmax, 2, false);
  for (int i = 1; i < ev.N_steps_z; ++i)
    for (int j = 0; j < N_bins_E; ++j) out[(size_t)(i - 1) * N_bins_E + j] = ev.Lum(ev.z[i], ev.Emin[j], ev.Emax[j], 0);
  return 0;
}

// Special functions exactly as the reference reaches them (through the shim).
double nusi_ref_dilog(double x) { return gsl_sf_dilog(x); }
void nusi_ref_cdilog(double x, double y, double *re, double *im) {
  gsl_sf_result a, b;
  gsl_sf_complex_dilog_xy_e(x, y, &a, &b);
  *re = a.val; *im = b.val;
}
double nusi_ref_Li2(double x) { return polylogarithm::Li2(x); }
double nusi_ref_Li3(double x) { return polylogarithm::Li3(x); }
double nusi_ref_atandiff(double x, double y) { return nuSIaux::atandiff(x, y); }
double nusi_ref_dilogdiff(double x, double y) { return nuSIaux::dilogdiff(x, y); }
double nusi_ref_dilog1mdiff(double x, double y) { return nuSIaux::dilog1mdiff(x, y); }
double nusi_ref_dilog1pdiff(double x, double y) { return nuSIaux::dilog1pdiff(x, y); }
double nusi_ref_dilog1over1mdiff(double x, double y) { return nuSIaux::dilog1over1mdiff(x, y); }
void nusi_ref_dilogdiff_complex(double xr, double xi, double yr, double yi, double *re, double *im) {
  double _Complex x = xr + I * xi, y = yr + I * yi;
  double _Complex r = nuSIaux::dilogdiff_complex(x, y);
  *re = creal(r); *im = cimag(r);
}
double nusi_ref_check_energy_conservation(double mphi, double g, double mntot, double si, double norm,
                                          int majorana, int non_resonant, int normal_ordering, int N_bins_E,
                                          double lEmin, double lEmax, double zmax, int flav) {
  calculate_flux ev(mphi, g, mntot, si, norm, majorana != 0, non_resonant != 0, normal_ordering != 0,
                    N_bins_E, lEmin, lEmax, zmax, flav, false);
  // energy_FS() reads norm_total, which only evolve() sets (nuSIprop.hpp:205): call evolve() first,
  // as every caller of the reference does, so the integral is not taken with an uninitialised member
  ev.evolve();
  return ev.check_energy_conservation();
}

}  // extern "C"

#ifdef NUSI_REF_MAIN
// CLI: nusi_ref evolve mphi g mntot si norm maj nonres NO N lEmin lEmax zmax flav phiphi [repeat]
//   prints N rows: E  m1 m2 m3  e mu tau  (%.17g).  With repeat>1 it re-runs evolve() and prints
//   the mean wall seconds per evolve() on stderr-free stdout line "#time <s>".
int main(int argc, char **argv) {
  if (argc < 16 || std::strcmp(argv[1], "evolve") != 0) {
    std::fprintf(stderr, "usage: %s evolve mphi g mntot si norm majorana non_resonant normal_ordering N lEmin lEmax zmax flav phiphi [repeat] [quiet]\n", argv[0]);
    return 2;
  }
  double mphi = atof(argv[2]), g = atof(argv[3]), mntot = atof(argv[4]), si = atof(argv[5]), norm = atof(argv[6]);
  int maj = atoi(argv[7]), nr = atoi(argv[8]), no = atoi(argv[9]), N = atoi(argv[10]);
  double lEmin = atof(argv[11]), lEmax = atof(argv[12]), zmax = atof(argv[13]);
  int flav = atoi(argv[14]), pp = atoi(argv[15]);
  int repeat = argc > 16 ? atoi(argv[16]) : 1;
  int quiet = argc > 17 ? atoi(argv[17]) : 0;
  calculate_flux ev(mphi, g, mntot, si, norm, maj != 0, nr != 0, no != 0, N, lEmin, lEmax, zmax, flav, pp != 0);
  auto t0 = std::chrono::steady_clock::now();
  for (int r = 0; r < repeat; ++r) ev.evolve();
  auto t1 = std::chrono::steady_clock::now();
  if (!quiet)
    for (int j = 0; j < N; ++j)
      std::printf("%.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", ev.get_energy(j), ev.get_flux(0, j),
                  ev.get_flux(1, j), ev.get_flux(2, j), ev.get_flux_fla(0, j), ev.get_flux_fla(1, j),
                  ev.get_flux_fla(2, j));
  std::printf("#time %.6f\n", std::chrono::duration<double>(t1 - t0).count() / repeat);
  return 0;
}
#endif
