// nuSIprop.hpp -- drop-in for the reference's header of the same name: the class
// nuSIprop::calculate_flux (reference nuSIprop.hpp:22-540) with the same constructor signature and
// defaults (:61-65), public members mphi/g/mntot/si/norm (:174), evolve() (:176),
// check_energy_conservation() (:339), get_flux/get_flux_fla/get_N_bins_E/get_energy (:359-429)
// and value (deep-copy) semantics (:434-540).
//
// It contains NO physics: evolve() forwards to the C ABI of include/nusiprop_b200.h
// (libnusiprop_b200.so), where the cross-section builder and the z-stepping integrator run as
// CUDA kernels on a B200.  Compile a consumer exactly like the reference's test.cpp, plus the library:
//     g++ -std=c++11 -O2 -I nusiprop_b200/cxx -I include test.cpp -L nusiprop_b200/lib -lnusiprop_b200
//
// Differences a user can observe (all deliberate):
//   * objects with the same fixed options share one GPU context (grids, source term, phi-phi tables); the library
//     serialises calls on one context and the registry of contexts is locked, so one object per thread (an OpenMP
//     loop over a scan, as with the reference) is safe -- though evolve_scan() below is the fast way to do that;
//     it is created at the first evolve() -- or in the constructor when non_resonant && phiphi, where
//     the reference loads its tables and exit(1)s if a file is missing (interp.hpp:251-254);
//   * fatal conditions print the reference's message and exit(1) like the reference, unless
//     NUSIPROP_B200_EXCEPTIONS is defined, in which case they throw std::runtime_error (the Cython
//     wrapper is built that way: its pxd declares `except +`);
//   * the getters reject bin == N_bins_E (the reference tests `j > N_bins_E` and reads past the end);
//   * nuSIprop::evolve_scan() is new: many (mphi, g, mntot) points in one call;
//   * one extra, trailing constructor argument source_model (default 0 = the reference's DSNB source, for which si
//     and norm are inert as in the reference; 1 = the upstream power-law source the reference keeps commented out,
//     nuSIprop.hpp:648-657, with which check_energy_conservation() is meaningful).
#ifndef NUSIPROP_B200_DROPIN_HPP
#define NUSIPROP_B200_DROPIN_HPP

#include <cmath>
#include <cstdlib>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "nusiprop_b200.h"

namespace nuSIprop {

namespace detail {

inline void fatal(const std::string &msg) {
#ifdef NUSIPROP_B200_EXCEPTIONS
  throw std::runtime_error(msg);
#else
  std::cerr << msg << std::endl;
  std::exit(1);
#endif
}

// one GPU context per distinct set of fixed options, shared by all objects that use it
struct ctx_handle {
  nusi_ctx *c = nullptr;
  std::mutex mu;  // held across "set the power-law source, then evolve" (source_model 1: si and norm are per object)
  explicit ctx_handle(nusi_ctx *c_) : c(c_) {}
  ~ctx_handle() { nusi_ctx_destroy(c); }
  ctx_handle(const ctx_handle &) = delete;
  ctx_handle &operator=(const ctx_handle &) = delete;
};
// fixed options ..., device, source_model.  si and norm are not part of the key: with source_model 1 evolve() re-uploads
// the power-law table of the shared context when they differ (nusi_ctx_set_powerlaw), so sampling si creates no contexts.
typedef std::tuple<bool, bool, bool, int, double, double, double, int, bool, int, int> ctx_key;

inline std::shared_ptr<ctx_handle> acquire_ctx(const ctx_key &key, double si = 2.0, double norm = 1.0) {
  static std::map<ctx_key, std::weak_ptr<ctx_handle>> registry;
  static std::mutex registry_mu;
  std::lock_guard<std::mutex> lk(registry_mu);
  std::shared_ptr<ctx_handle> h = registry[key].lock();
  if (h) return h;
  nusi_opts o;
  nusi_opts_default(&o);
  o.majorana = std::get<0>(key);
  o.non_resonant = std::get<1>(key);
  o.normal_ordering = std::get<2>(key);
  o.N_bins_E = std::get<3>(key);
  o.lEmin = std::get<4>(key);
  o.lEmax = std::get<5>(key);
  o.zmax = std::get<6>(key);
  o.flav = std::get<7>(key);
  o.phiphi = std::get<8>(key);
  o.device = std::get<9>(key);
  o.source_model = std::get<10>(key);
  if (o.source_model == 1) {
    o.si = si;
    o.norm = norm;
  }
  nusi_ctx *c = nullptr;
  const int rc = nusi_ctx_create(&c, &o);
  if (rc != NUSI_OK) {
    fatal(std::string("nuSIprop (B200): cannot create the GPU context: ") + nusi_strerror(rc) + " (" + nusi_last_error() + ")");
    return h;
  }
  h = std::make_shared<ctx_handle>(c);
  registry[key] = h;
  return h;
}

inline int default_device() {
  const char *e = std::getenv("NUSIPROP_DEVICE");
  return e ? std::atoi(e) : 0;
}

}  // namespace detail

class calculate_flux {
 public:
  calculate_flux() : calculate_flux(1e7, 0.1, 0.1, 2, 1, true, false) {}  // as the reference (:59): needed by Cython

  calculate_flux(double mphi_, double g_, double mntot_, double si_, double norm_ = 1, bool majorana_ = true,
                 bool non_resonant_ = true, bool normal_ordering_ = true, int N_bins_E_ = 300, double lEmin_ = 12.0,
                 double lEmax_ = 17.0, double zmax_ = 5.0, int flav_ = 2, bool phiphi_ = false, int source_model_ = 0)
      : mphi(mphi_), g(g_), mntot(mntot_), si(si_), norm(norm_), majorana(majorana_), non_resonant(non_resonant_),
        normal_ordering(normal_ordering_), N_bins_E(N_bins_E_), lEmin(lEmin_), lEmax(lEmax_), zmax_arg(zmax_), flav(flav_),
        phiphi(phiphi_), source_model(source_model_), device(detail::default_device()) {
    if (N_bins_E < 2 || !(lEmax > lEmin) || !(zmax_arg > 0) || flav < 0 || flav > 2)
      detail::fatal("nuSIprop (B200): invalid binning / flavour options");
    flux.assign(3, std::vector<double>(N_bins_E, 0.0));
    flux_fla.assign(3, std::vector<double>(N_bins_E, 0.0));
    E_nu.assign(N_bins_E, 0.0);
    Emin.assign(N_bins_E, 0.0);
    Emax.assign(N_bins_E, 0.0);
    // energy and redshift grids (reference :113-128): host arithmetic inside the library, no GPU needed
    if (nusi_grid(N_bins_E, lEmin, lEmax, zmax_arg, &N_steps_z, &zmax, Emin.data(), E_nu.data(), Emax.data()) != NUSI_OK)
      detail::fatal("nuSIprop (B200): invalid binning options");
    // the reference loads the phi-phi tables here and exits if they are missing (:166-170)
    if (non_resonant && phiphi) {
      ctx = detail::acquire_ctx(key(), si, norm);
    }
  }

  /* Parameters that can be modified after an object has been created */
  double mphi, g, mntot, si, norm;

  void evolve() {
    if (!ctx) ctx = detail::acquire_ctx(key(), si, norm);
    if (!ctx) return;
    std::vector<double> fm(3 * (size_t)N_bins_E), ff(3 * (size_t)N_bins_E);
    status = 0;
    int rc;
    if (source_model == 1) {  // si and norm are public, modifiable members: the shared context gets this object's source first
      std::lock_guard<std::mutex> lk(ctx->mu);
      rc = nusi_ctx_set_powerlaw(ctx->c, si, norm);
      if (rc == NUSI_OK) rc = nusi_evolve_batch(ctx->c, 1, &mphi, &g, &mntot, fm.data(), ff.data(), &status);
    } else {
      rc = nusi_evolve_batch(ctx->c, 1, &mphi, &g, &mntot, fm.data(), ff.data(), &status);
    }
    if (rc != NUSI_OK) {
      detail::fatal(std::string("nuSIprop (B200): evolve failed: ") + nusi_strerror(rc) + " (" + nusi_last_error() + ")");
      return;
    }
    if (status & NUSI_ST_NO_MASS) {  // aux.hpp:48-49
      detail::fatal("No neutrino mass spectrum was found corresponding to \\sum m = " + std::to_string(mntot) + ". Exiting...");
      return;
    }
    if (status & NUSI_ST_PP_RANGE) {  // interp.hpp:355-361
      detail::fatal("Error at interp: a coordinate of the phi-phi cross section lookup is outside the bounds of the table");
      return;
    }
    // the reference's stderr diagnostics (:909-918, :1215-1231, :1505-1516), once per evolve()
    if (status & NUSI_ST_WARN_GAMMA) std::cerr << "Negative cross section when computing Gamma" << std::endl;
    if (status & NUSI_ST_WARN_ALPHATILDE) std::cerr << "Negative cross section when computing alphaTilde" << std::endl;
    if (status & NUSI_ST_WARN_ALPHA) std::cerr << "Negative cross section when computing alpha" << std::endl;
    for (int k = 0; k < 3; ++k)
      for (int i = 0; i < N_bins_E; ++i) {
        flux[k][i] = fm[(size_t)k * N_bins_E + i];
        flux_fla[k][i] = ff[(size_t)k * N_bins_E + i];
      }
  }

  double check_energy_conservation(void) {
    // (E_int - E_FS)/E_FS, reference :339-357; E_FS from the library's host code (:666-744)
    double E_FS = 0;
    nusi_freestream_energy(zmax, lEmin, lEmax, si, norm, &E_FS);
    evolve();
    double E_int = 0;
    for (int i = 0; i < N_bins_E; ++i)
      for (int k = 0; k < 3; ++k) E_int += (std::log(Emax[i]) - std::log(Emin[i])) * (E_nu[i] * E_nu[i]) * flux[k][i];
    return (E_int - E_FS) / E_FS;
  }

  /* extra (not in the reference): the same ratio with E_FS the free-streaming energy of the source that is really
   * injected (this fork's supernova background for source_model 0), nusi_ctx_freestream_energy: what
   * check_energy_conservation() is meant to measure.  The reference's own number above is -1 in the fork's default mode. */
  double check_energy_conservation_consistent(void) {
    evolve();
    if (!ctx) return 0;
    double E_FS = 0;
    nusi_ctx_freestream_energy(ctx->c, &E_FS);
    double E_int = 0;
    for (int i = 0; i < N_bins_E; ++i)
      for (int k = 0; k < 3; ++k) E_int += (std::log(Emax[i]) - std::log(Emin[i])) * (E_nu[i] * E_nu[i]) * flux[k][i];
    return (E_int - E_FS) / E_FS;
  }

  inline double get_flux(int i, int j) {
    if (!index_ok("mass", i, j)) return 0;
    return flux[i][j];
  }
  inline double get_flux_fla(int i, int j) {
    if (!index_ok("flavor", i, j)) return 0;
    return flux_fla[i][j];
  }
  inline int get_N_bins_E() { return N_bins_E; }
  inline double get_energy(int i) {
    if (i < 0) {
      std::cerr << "You asked for the energy at the bin " << i << "<0! Zero will be returned." << std::endl;
      return 0;
    } else if (i >= N_bins_E) {
      std::cerr << "You asked for the energy at the bin " << i << ", but there are only " << N_bins_E
                << " bins! Zero will be returned." << std::endl;
      return 0;
    }
    return E_nu[i];
  }

  /* extras (not in the reference) */
  int get_status() const { return status; }          // NUSI_ST_* bits of the last evolve()
  int get_N_steps_z() const { return N_steps_z; }
  double get_zmax() const { return zmax; }            // effective zmax = z[N_steps_z-1] (:128)

 private:
  bool majorana, non_resonant, normal_ordering;
  int N_bins_E;
  double lEmin, lEmax, zmax_arg;
  int flav;
  bool phiphi;
  int source_model;  // extra: 0 = the reference's DSNB source, 1 = upstream power law norm (E/1e14 eV)^-si (hpp:648-657)
  int device;
  int N_steps_z = 0;
  double zmax = 0;
  int status = 0;
  std::vector<std::vector<double>> flux, flux_fla;  // [3][N_bins_E]; copied with the object
  std::vector<double> E_nu, Emin, Emax;
  std::shared_ptr<detail::ctx_handle> ctx;  // shared, never copied

  detail::ctx_key key() const {
    return detail::ctx_key(majorana, non_resonant, normal_ordering, N_bins_E, lEmin, lEmax, zmax_arg, flav, phiphi, device,
                           source_model);
  }
  bool index_ok(const char *what, int i, int j) const {
    if ((i < 0) || (i >= 3)) {
      std::cerr << "You asked for the flux of the " << what << " eigenstate " << i << ", not in [0,1,2]. Zero will be returned."
                << std::endl;
      return false;
    } else if (j < 0) {
      std::cerr << "You asked for the flux at the energy bin " << j << "<0! Zero will be returned." << std::endl;
      return false;
    } else if (j >= N_bins_E) {
      std::cerr << "You asked for the flux at the energy bin " << j << ", but there are only " << N_bins_E
                << " bins! Zero will be returned." << std::endl;
      return false;
    }
    return true;
  }
};

// Many parameter points in one call (the GPU's natural unit of work).  flux_mass / flux_fla are
// [n][3][N_bins_E] row-major (either may be NULL), status [n] NUSI_ST_* bits (may be NULL).
// Returns NUSI_OK or a NUSI_E_* code; never exits.
inline int evolve_scan(int n, const double *mphi, const double *g, const double *mntot, double *flux_mass, double *flux_fla,
                       int *status, bool majorana = true, bool non_resonant = true, bool normal_ordering = true,
                       int N_bins_E = 300, double lEmin = 12.0, double lEmax = 17.0, double zmax = 5.0, int flav = 2,
                       bool phiphi = false) {
  std::shared_ptr<detail::ctx_handle> h = detail::acquire_ctx(detail::ctx_key(
      majorana, non_resonant, normal_ordering, N_bins_E, lEmin, lEmax, zmax, flav, phiphi, detail::default_device(), 0));
  if (!h) return NUSI_E_NO_DEVICE;
  return nusi_evolve_batch(h->c, n, mphi, g, mntot, flux_mass, flux_fla, status);
}

}  // namespace nuSIprop
#endif
