// Host-side compile of the product's __host__ __device__ math headers, for unit checks
// that must run without a GPU (pytest -m "not gpu").  Development/test aid only: this
// library is built under tests/ and is never loaded by the nusiprop_b200 package.
#include "nusi_math.h"
#ifdef NUSI_HOSTCHECK_XSEC
#include "nusi_gamma.h"
#include "nusi_alphatilde.h"
#include "nusi_alpha.h"
#endif

extern "C" {
double hc_li2_real(double x) { return nusi::li2_real(x); }
void hc_cli2(double x, double y, double *re, double *im) {
  nusi::cplx r = nusi::cli2(x, y);
  *re = r.re;
  *im = r.im;
}
double hc_li2_neg(double x) { return nusi::li2_neg(x); }
double hc_li3_neg(double x) { return nusi::li3_neg(x); }
}

#include "nusi_phiphi_gen.h"
extern "C" {
double hc_pp_integrand(double s, double t) { return nusi::pp_integrand(s, t); }
double hc_pp_alpha_node(double splus, double n, double l10d) { return nusi::pp_alpha_node(splus, n, l10d); }
double hc_pp_alphatilde_node(double abs_tplus, double l10d) { return nusi::pp_alphatilde_node(abs_tplus, l10d); }
}

#ifdef NUSI_HOSTCHECK_XSEC
static int hc_shared_logs = 0;
extern "C" {
void hc_set_shared_logs(int on) { hc_shared_logs = on; }
// lo/hi: the NT bin edges of the table rows (nuSIprop.hpp:223-233), built by the caller.
void hc_tables(double mphi, double g, double Ga, const double *mn, const double *U2f, int majorana,
               int non_resonant, int phiphi, int NT, const double *lo, const double *hi, double *Gamma,
               double *alphaTilde, double *alpha /* dense NT x NT, m > i */, int *warn) {
  nusi::PointParams P;
  P.mphi = mphi; P.g = g; P.Ga = Ga;
  for (int k = 0; k < 3; ++k) { P.mn[k] = mn[k]; P.U2f[k] = U2f[k]; }
  P.majorana = majorana; P.non_resonant = non_resonant; P.phiphi = phiphi;
  *warn = 0;
  for (int i = 0; i < NT; ++i) {
    const bool split = majorana && non_resonant && !phiphi && hc_shared_logs == 4;
    if (Gamma) Gamma[i] = split ? nusi::gamma_row_split(P, lo[i], hi[i], warn) : nusi::gamma_row(P, lo[i], hi[i], warn);
    if (alphaTilde)
      alphaTilde[i] = split ? nusi::alphatilde_row_split(P, lo[i], hi[i], warn) : nusi::alphatilde_row(P, nullptr, lo[i], hi[i], warn);
    if (alpha)
      for (int m = i + 1; m < NT; ++m)
        alpha[(size_t)i * NT + m] = (majorana && non_resonant && !phiphi && hc_shared_logs == 4)
                                        ? nusi::alpha_entry_grouped(P, lo[i], hi[i], lo[m], hi[m], warn)
                                    : (majorana && non_resonant && hc_shared_logs == 3)
                                        ? nusi::alpha_entry_compact<true>(P, nullptr, lo[i], hi[i], lo[m], hi[m], warn)
                                    : (majorana && non_resonant && hc_shared_logs == 2)
                                        ? nusi::alpha_entry_compact<false>(P, nullptr, lo[i], hi[i], lo[m], hi[m], warn)
                                    : (majorana && non_resonant && hc_shared_logs)
                                        ? nusi::alpha_entry_shared(P, nullptr, lo[i], hi[i], lo[m], hi[m], warn)
                                        : nusi::alpha_entry(P, nullptr, lo[i], hi[i], lo[m], hi[m], warn);
  }
}
}
#endif
