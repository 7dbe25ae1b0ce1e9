// TEST INFRASTRUCTURE (oracle side) -- never included by the product.
//
// Stand-in for the `polylogarithm` git submodule (Expander/polylogarithm @
// f21dea82009b064d5d130c1c6177298a1764c68a), whose sources are absent from
// /root/reference (directory polylogarithm/ is empty).  The reference calls
// polylogarithm::Li2 / Li3 only from Lum_int (nuSIprop.hpp:628-646) with
// argument -exp(-x) in [-1, 0].  We evaluate x * Chebyshev(Li_n(x)/x), table
// generated with mpmath by tools/gen_polylog_coeffs.py; accuracy <= 2e-16
// relative on [-1, 0] (tests/test_oracle_cpu.py and tests/test_host_math.py pin it against mpmath: tests/golden/specfun_mpmath.json).
#ifndef REFSHIM_POLYLOG_NEG_HPP
#define REFSHIM_POLYLOG_NEG_HPP
#include <cstdio>
#include <cstdlib>
#include "polylog_cheb.inc"

namespace polylogarithm {
namespace detail {
inline double cheb_neg(const double *c, double x) {
  // Clenshaw on t = 2x+1
  const double t = 2.0 * x + 1.0, t2 = 2.0 * t;
  double b1 = 0.0, b2 = 0.0;
  for (int j = NUSI_POLYLOG_NCHEB - 1; j >= 1; --j) {
    const double b0 = c[j] + (t2 * b1 - b2);
    b2 = b1;
    b1 = b0;
  }
  return x * (c[0] + (t * b1 - b2));
}
inline void range_check(double x, const char *nm) {
  if (!(x >= -1.0 && x <= 0.0)) {
    std::fprintf(stderr, "refshim polylogarithm::%s(%g): stand-in only covers [-1,0]\n", nm, x);
    std::abort();
  }
}
}  // namespace detail
inline double Li2(double x) {
  static const double c[NUSI_POLYLOG_NCHEB] = {NUSI_LI2_CHEB};
  detail::range_check(x, "Li2");
  return detail::cheb_neg(c, x);
}
inline double Li3(double x) {
  static const double c[NUSI_POLYLOG_NCHEB] = {NUSI_LI3_CHEB};
  detail::range_check(x, "Li3");
  return detail::cheb_neg(c, x);
}
}  // namespace polylogarithm
#endif
