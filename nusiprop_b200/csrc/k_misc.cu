// k_misc.cu -- source term (reference nuSIprop.hpp:638-662), unit-test and calibration kernels
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
#include "nusi_launch.h"
#include "nusi_math.h"

namespace nusi {

// ------------------------------------------------------------------------------------------
// source term: Lum(z, Em, Ep) = (Lum_int(z,Ep) - Lum_int(z,Em)) * R_SN(z)   nuSIprop.hpp:638-662
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double lum_int(double pref, double z, double E) {
  const double Tnue = 6e6;
  const double u = -exp(-E * (1 + z) / Tnue);
  return pref * (-E * E * (1 + z) * log(exp(-E * (1 + z) / Tnue) + 1) / Tnue + 2 * E * li2_neg(u) +
                 2 * Tnue * li3_neg(u) / (1 + z));
}
__global__ void source_kernel(int N, int S, const double *__restrict__ zsrc, const double *__restrict__ rsn,
                              double pref, const double *__restrict__ Emin, const double *__restrict__ Emax,
                              double *__restrict__ out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * S) return;
  const int i = idx / N, j = idx - i * N;
  const double z = zsrc[i];
  out[idx] = (lum_int(pref, z, Emax[j]) - lum_int(pref, z, Emin[j])) * rsn[i];
}

// ------------------------------------------------------------------------------------------
// unit-test and calibration kernels
// ------------------------------------------------------------------------------------------
__global__ void specfun_kernel(int which, int n, const double *__restrict__ x, const double *__restrict__ y,
                               double *__restrict__ re, double *__restrict__ im) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (which == 0) {
    re[i] = li2_real(x[i]);
  } else if (which == 1) {
    const cplx r = cli2(x[i], y[i]);
    re[i] = r.re;
    im[i] = r.im;
  } else if (which == 2) {
    re[i] = li2_neg(x[i]);
  } else if (which == 3) {
    re[i] = li3_neg(x[i]);
  } else if (which == 4) {  // the elementary functions as the builders call them
    re[i] = nlog(x[i]);
  } else if (which == 5) {
    re[i] = nlog1p(x[i]);
  } else if (which == 6) {
    re[i] = natan2(x[i], y[i]);
  } else {
    re[i] = exp(x[i]);
  }
}

// 8 independent FMA chains per thread: measures the DFMA issue rate of the FP64 pipe
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *out, int iters, double a, double b) {
  double v0 = threadIdx.x * 1e-3, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3, v4 = v0 + 4, v5 = v0 + 5, v6 = v0 + 6, v7 = v0 + 7;
  for (int i = 0; i < iters; ++i) {
    v0 = fma(v0, a, b); v1 = fma(v1, a, b); v2 = fma(v2, a, b); v3 = fma(v3, a, b);
    v4 = fma(v4, a, b); v5 = fma(v5, a, b); v6 = fma(v6, a, b); v7 = fma(v7, a, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((v0 + v1) + (v2 + v3)) + ((v4 + v5) + (v6 + v7));
}

// Energy integral of check_energy_conservation (nuSIprop.hpp:351-355): E_int = sum_i sum_k (ln Emax_i - ln Emin_i) E_i^2 flux_k[i],
// one warp per point, fluxes read where the integrator left them in HBM.  wE2[i] = (ln Emax_i - ln Emin_i) E_i^2 (host, glibc).
// Summation order differs from the reference's serial loop (tree over lanes): agreement to ~N ulp, not bitwise.
__global__ void __launch_bounds__(128) energy_integral_kernel(int n, int N, const double *__restrict__ flux_mass,
                                                              const double *__restrict__ wE2, double *__restrict__ out) {
  const int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (p >= n) return;
  const double *f = flux_mass + (size_t)p * 3 * N;
  double acc = 0.0;
  for (int i = lane; i < N; i += 32) acc += wE2[i] * ((f[i] + f[N + i]) + f[2 * N + i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) out[p] = acc;
}

cudaError_t launch_source(int N, int S, const double *zsrc, const double *rsn, double pref, const double *Emin, const double *Emax,
                          double *out, cudaStream_t s) {
  source_kernel<<<(S * N + 127) / 128, 128, 0, s>>>(N, S, zsrc, rsn, pref, Emin, Emax, out);
  return cudaGetLastError();
}
cudaError_t launch_specfun(int which, int n, const double *x, const double *y, double *re, double *im) {
  specfun_kernel<<<(n + 127) / 128, 128>>>(which, n, x, y, re, im);
  return cudaGetLastError();
}
cudaError_t launch_fp64_peak(int blocks, int threads, double *out, int iters) {
  fp64_peak_kernel<<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
  return cudaGetLastError();
}
cudaError_t launch_energy_integral(int n, int N, const double *flux_mass, const double *wE2, double *out, cudaStream_t s) {
  energy_integral_kernel<<<(n + 3) / 4, 128, 0, s>>>(n, N, flux_mass, wE2, out);
  return cudaGetLastError();
}

}  // namespace nusi
