// nusi_host.hpp -- parameter-independent host setup of a context and per-point host
// preparation.  These are the pieces of the reference that run once per object / once per
// evolve() and cost nothing; they stay on the host so that their inputs to the kernels are the
// same doubles the reference's own glibc calls produce (pow, cos, sin, sqrt).
#ifndef NUSI_HOST_HPP
#define NUSI_HOST_HPP
#include <vector>
#include <string>
#include "nusi_aux.h"

namespace nusi {

struct HostGrid {
  int N = 0;    // N_bins_E
  int Nz = 0;   // N_steps_z
  int NT = 0;   // N + Nz - 2 table rows
  double zmax_eff = 0;
  std::vector<double> Emin, Ecen, Emax, dE;  // [N]
  std::vector<double> z;                     // [Nz]
  std::vector<double> row_lo, row_hi;        // [NT] bin edges of the table rows (nuSIprop.hpp:223-233)
  // per z-step i = 1..Nz-1 (index i-1): factors of the implicit scheme (nuSIprop.hpp:256-294)
  std::vector<double> cz;    // (1+z[i-1]) * dlogz / H(z[i-1])
  std::vector<double> ndz;   // nd(z[i-1]) / (1+z[i-1])^2
  std::vector<double> rsn;   // R_SN(z[i])   (source is evaluated at z[i], rates at z[i-1])
  std::vector<double> zsrc;  // z[i]
  double lum_pref = 0;       // Etot*120/(6*7*pi^4*T^2) of Lum_int (nuSIprop.hpp:645)
  double U2[9];              // |U_ab|^2
};

// nuSIprop.hpp:113-128 (grids), :130-163 (PMNS), :573-616 (cosmology)
void build_grid(int N_bins_E, double lEmin, double lEmax, double zmax, bool normal_ordering, HostGrid &G);

// nuSIaux::getmL (aux.hpp:12-50) + nuSIprop.hpp:184-203.  false = no admissible root.
bool solve_masses(double mntot, bool normal_ordering, double mn[3]);

// scalar_width (nuSIprop.hpp:748-757)
double scalar_width(double mphi, double g, bool majorana);

// free-streaming diagnostics: energy_FS (nuSIprop.hpp:694-744) for norm_total = norm / flux_FS_E0 (:205, :666-692)
void powerlaw_source(int N, int S, const double *zsrc, const double *Emin, const double *Emax, double zmax_eff, double si,
                     double norm, double *out);
double freestream_energy(double zmax_eff, double lEmin, double lEmax, double si, double norm);
// This fork's source, Lum(z_i, bin j) = (Lum_int(z, Ep) - Lum_int(z, Em)) R_SN(z)  (nuSIprop.hpp:638-662), with the same
// glibc calls and the same Li2 / Li3 evaluation as the reference build: out[i * N + j], i < S.
void dsnb_source(int N, int S, const double *zsrc, const double *rsn, double pref, const double *Emin, const double *Emax, double *out);
// counterpart of energy_FS() (nuSIprop.hpp:694-729) for that source: \int dz/H \int dE E sum_i L^i(z, E(1+z))
double freestream_energy_dsnb(double zmax_eff, double lEmin, double lEmax);

// phi-phi tables: binary loader (interp.hpp:249-292) + contraction of the log10(delta) axis.
struct HostPhiPhi {
  std::vector<double> at_x0, at_f;        // ln|tbar+| nodes, contracted values
  std::vector<double> a_x0, a_x1, a_f;    // ln sbar nodes, n nodes, contracted values [n0][n1]
  int delta_in_range = 0;
};
// returns "" on success, else an error message
std::string load_phiphi(const char *alphatilde_path, const char *alpha_path, double log10_delta, HostPhiPhi &T);

}  // namespace nusi
#endif
