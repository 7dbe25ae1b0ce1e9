// nusi_split.h -- device helpers shared by the role-split integrators (k_integrate_split.cu: one CTA per point;
// k_integrate_cluster.cu: one thread-block cluster per point): mbarrier / TMA bulk copy wrappers and the out-of-line
// general case of one implicit step.
#ifndef NUSI_SPLIT_H
#define NUSI_SPLIT_H
#include "nusi_implicit.h"

namespace nusi {

// ---- mbarrier / TMA bulk copy (PTX ISA 8.x, sm_90+) ------------------------------------------------------------
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
// all NTHR threads of the CTA, from any of the role loops (named barrier 1; __syncthreads() must not be used in divergent code)
template <int NTHR>
__device__ __forceinline__ void block_sync() {
  asm volatile("bar.sync 1, %0;" ::"n"(NTHR) : "memory");
}
// global -> shared bulk copy; dst, src 16-byte aligned, bytes a multiple of 16; completes on the mbarrier
__device__ __forceinline__ void tma_bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
               "r"(bar)
               : "memory");
}

__device__ __forceinline__ double2 lds_f64x2(unsigned addr) {  // 16-byte shared load that the compiler keeps where it is written
  double2 r;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "r"(addr));
  return r;
}

// the rare general case of one implicit step (M is not the identity to rounding, or a quotient overflowed): out of line,
// scalar arguments (arrays would force the caller's values into local memory on the hot path), results through shared memory
static __device__ __noinline__ void split_fallback(double U0, double U1, double U2, double cz, const double *slot, double s0, double s1,
                                                   double s2, double f0, double f1, double f2, double *out) {
  const double U2f[3] = {U0, U1, U2}, src[3] = {s0, s1, s2};
  double f[3] = {f0, f1, f2};
  const double x[3] = {f0 + s0, f1 + s1, f2 + s2};
  const double Z0 = slot[4 * 32], Z1 = slot[6 * 32], Z2 = slot[8 * 32], dEb = slot[10 * 32];
  double v[3];
  v[0] = x[0] / Z0;
  v[1] = x[1] / Z1;
  v[2] = x[2] / Z2;
  const double vmax = fmax(fabs(v[0]), fmax(fabs(v[1]), fabs(v[2]))), vmin = fmin(fabs(v[0]), fmin(fabs(v[1]), fabs(v[2])));
  // same sufficient condition for "the 3x3 solve returns v" as implicit_post(); both branches give the same bits
  if (slot[32] * vmax <= 6.9388939039072284e-18 * vmin) {
    f[0] = v[0];
    f[1] = v[1];
    f[2] = v[2];
  } else {
    implicit_bin(U2f, cz, slot[2 * 32], slot[3 * 32], dEb, src, f);
  }
  out[0] = f[0];
  out[32] = f[1];
  out[64] = f[2];
  out[96] = (U2f[0] * f[0] + U2f[1] * f[1] + U2f[2] * f[2]) / dEb;
}

}  // namespace nusi
#endif
