// k_integrate_check.cu -- z-stepping integrators kept as cross-checks (reference-shaped serial kernel, bin-owned wavefront) and the resonant-only recurrence
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
#include "nusi_launch.h"
#include "nusi_implicit.h"
#include "nusi_cpasync.h"

namespace nusi {

#ifndef NUSI_WAVE_MINB
#define NUSI_WAVE_MINB 2
#endif

// ------------------------------------------------------------------------------------------
// z-stepping integrator, general (non-resonant) tables.  One CTA per parameter point; thread
// `tid` owns bins tid, tid+T, ... (BPT of them).  Inside a z-step the bins are a back-
// substitution from the highest energy down (nuSIprop.hpp:266): when bin m is final, its
// weighted flux W_m = sum_l |U_fl|^2 flux_l[m] / dE_m is broadcast and every lower bin b does
//     acc[b] += T[b+i-1][m+i-1] * W_m
// which is one coalesced read of table column m+i-1.  src_k(b) = cz (Lum + nd' |U_fk|^2 acc[b]).
// ------------------------------------------------------------------------------------------
template <int BPT>
__global__ void __launch_bounds__(256)
integrate_kernel(DeviceGrid G, const PointParams *__restrict__ params, const double *__restrict__ rows,
                 const double *__restrict__ tri, size_t tstride, double *__restrict__ flux_mass,
                 double *__restrict__ flux_fla, const int *__restrict__ out_index, const int *__restrict__ status_in,
                      int *__restrict__ status_out) {
  const int p = blockIdx.x, tid = threadIdx.x, nthr = blockDim.x;
  const int N = G.N, NT = G.NT;
  const PointParams P = params[p];
  __shared__ double sW[2];
  const bool dead = !(P.mphi > 0);
  const double *rowG = rows + ((size_t)p * 2 + 0) * NT;
  const double *rowA = rows + ((size_t)p * 2 + 1) * NT;
  const double *T = tri + (size_t)p * tstride;
  const double U2f[3] = {P.U2f[0], P.U2f[1], P.U2f[2]};

  double f[BPT][3], acc[BPT], dEb[BPT];
#pragma unroll
  for (int r = 0; r < BPT; ++r) {
    const int b = tid + r * nthr;
    f[r][0] = f[r][1] = f[r][2] = 0.0;
    dEb[r] = (b < N) ? G.dE[b] : 1.0;
  }

  for (int i = G.Nz - 1; i > 0 && !dead; --i) {
    const double cz = G.cz[i - 1], ndz = G.ndz[i - 1];
    double Gam[BPT], aT[BPT], lum[BPT];
#pragma unroll
    for (int r = 0; r < BPT; ++r) {
      const int b = tid + r * nthr;
      acc[r] = 0.0;
      if (b < N) {
        Gam[r] = ndz * rowG[b + i - 1];
        aT[r] = ndz * rowA[b + i - 1];
        lum[r] = cz * G.source[(size_t)(i - 1) * N + b];
      }
    }
    for (int m = N - 1; m >= 0; --m) {
      const int c = m + i - 1;  // table column of source bin m at this redshift
      const double *col = T + (size_t)c * (c - 1) / 2;
      double tv[BPT];
#pragma unroll
      for (int r = 0; r < BPT; ++r) {
        const int b = tid + r * nthr;
        tv[r] = (b < m) ? col[b + i - 1] : 0.0;
      }
      const int rm = m / nthr;
      if (tid == m - rm * nthr) {
#pragma unroll
        for (int r = 0; r < BPT; ++r)
          if (r == rm) {
            double src[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) src[k] = lum[r] + cz * (ndz * acc[r]) * U2f[k];
            implicit_bin(U2f, cz, Gam[r], aT[r], dEb[r], src, f[r]);
            sW[m & 1] = (U2f[0] * f[r][0] + U2f[1] * f[r][1] + U2f[2] * f[r][2]) / dEb[r];
          }
      }
      __syncthreads();
      const double W = sW[m & 1];
#pragma unroll
      for (int r = 0; r < BPT; ++r) acc[r] = fma(tv[r], W, acc[r]);
    }
  }

  // bin-width divide and flavour projection (nuSIprop.hpp:329-336) into this point's output slot
  const int q = out_index ? out_index[p] : p;
#pragma unroll
  for (int r = 0; r < BPT; ++r) {
    const int b = tid + r * nthr;
    if (b < N) store_bin(G, flux_mass, flux_fla, q, b, f[r], dEb[r], dead);
  }
  if (status_out && tid == 0) status_out[q] = status_in[p];
}

// ------------------------------------------------------------------------------------------
// z-stepping integrator, wavefront form (non-resonant tables).
//
// Cell (s, m) = bin m at z-step s (s = 0 is the highest redshift) needs cell (s-1, m) and the
// cells (s, m' > m).  With wave index  w = s + (N-1-m)  every cell of a wave is independent,
// and -- because the redshift step equals the energy step -- ALL z-steps of a wave read the SAME
// table column  c = NT-1-w  (nuSIprop.hpp:272: both table indices shift by i-1).  So the packed
// table is streamed through shared memory exactly once per band of SB z-steps, column by column
// (a ring of NSTAGE columns filled by cp.async, NSTAGE-1 columns ahead of the wave that reads them, so
// the HBM latency of a column is off the wave-to-wave critical path), instead of once per z-step; the critical path shrinks from N*(Nz-1) dependent bin updates to
// N+SB-1 waves per band.  One CTA per parameter point; thread t owns bins t, t+256, ... with their
// flux and the SB partial regeneration sums acc[s] in registers.  Accumulation order per cell is
// source bins from high to low, as in integrate_kernel, so both produce the same bits.
// ------------------------------------------------------------------------------------------
template <int BPT, int SB, int NSTAGE>
__global__ void __launch_bounds__(256, (BPT == 1 ? NUSI_WAVE_MINB : 1))
integrate_wave_kernel(DeviceGrid G, const PointParams *__restrict__ params, const double *__restrict__ rows,
                      const double *__restrict__ tri, size_t tstride, double *__restrict__ flux_mass,
                      double *__restrict__ flux_fla, const int *__restrict__ out_index, const int *__restrict__ status_in,
                      int *__restrict__ status_out) {
  extern __shared__ __align__(16) double wsm[];
  const int p = blockIdx.x, tid = threadIdx.x, nthr = blockDim.x;
  const int N = G.N, NT = G.NT, S = G.Nz - 1;
  const PointParams P = params[p];
  const bool dead = !(P.mphi > 0);
  double *Tcol = wsm;                 // [NSTAGE][NT] ring of table columns
  double *Wsh = Tcol + NSTAGE * NT;   // [2][SB]
  double *sG = Wsh + 2 * SB;          // [NT]  Gamma rows
  double *sA = sG + NT;               // [NT]  alphaTilde rows
  double *sCz = sA + NT;              // [S]
  double *sNd = sCz + S;              // [S]
  const double *T = tri + (size_t)p * tstride;
  const double U2f[3] = {P.U2f[0], P.U2f[1], P.U2f[2]};
  constexpr int PRE = BPT + 2;        // NT <= 256*PRE is checked on the host

  for (int i = tid; i < NT; i += nthr) {
    sG[i] = rows[((size_t)p * 2 + 0) * NT + i];
    sA[i] = rows[((size_t)p * 2 + 1) * NT + i];
  }
  for (int i = tid; i < S; i += nthr) {
    sCz[i] = G.cz[i];
    sNd[i] = G.ndz[i];
  }
  double f[BPT][3], dEb[BPT];
#pragma unroll
  for (int r = 0; r < BPT; ++r) {
    const int b = tid + r * nthr;
    f[r][0] = f[r][1] = f[r][2] = 0.0;
    dEb[r] = (b < N) ? G.dE[b] : 1.0;
  }
  __syncthreads();

  for (int s0 = 0; s0 < S && !dead; s0 += SB) {
    const int cnt = min(SB, S - s0);
    double acc[BPT][SB];
#pragma unroll
    for (int r = 0; r < BPT; ++r)
#pragma unroll
      for (int sl = 0; sl < SB; ++sl) acc[r][sl] = 0.0;
    const int w_begin = s0, w_end = s0 + cnt - 1 + N - 1;
    // column c = NT-1-w into ring slot w % NSTAGE; rows >= c (above the diagonal) are zero-filled: a z-step
    // whose front has already passed bin b then adds 0 * W to an accumulator that is no longer read
    auto issue_column = [&](int w) {
      const int c = NT - 1 - w;
      const double *col = T + (size_t)c * (c - 1) / 2;
      double *dst = Tcol + (w % NSTAGE) * NT;
#pragma unroll
      for (int q = 0; q < PRE; ++q) {
        const int a = tid + q * nthr;
        if (a < NT) cp_async8(dst + a, col + (a < c ? a : 0), a < c);
      }
    };
#pragma unroll
    for (int j = 0; j < NSTAGE - 1; ++j) {
      if (w_begin + j <= w_end) issue_column(w_begin + j);
      cp_async_commit();
    }
    // flux-independent half of this thread's first cell of the band
    BinPre Bn[BPT];
    double snext[BPT];  // source term of the z-step after the prepared one, loaded one wave early (off the critical path)
#pragma unroll
    for (int r = 0; r < BPT; ++r) {
      const int b = tid + r * nthr;
      const int sgl = max(w_begin - (N - 1 - b), s0);
      snext[r] = 0.0;
      if (b < N && sgl < s0 + cnt) {
        const int zi = S - 1 - sgl;
        implicit_pre(U2f, sCz[zi], sNd[zi], sG[b + zi], sA[b + zi], G.source[(size_t)zi * N + b], dEb[r], Bn[r]);
        if (sgl + 1 < s0 + cnt) snext[r] = G.source[(size_t)(zi - 1) * N + b];
      }
    }
    for (int w = w_begin; w <= w_end; ++w) {
      const int buf = w & 1;
      cp_async_wait<NSTAGE - 2>();  // this thread's part of column w has landed (the barrier below publishes it)
      // finalise this thread's cells of the wave
#pragma unroll
      for (int r = 0; r < BPT; ++r) {
        const int b = tid + r * nthr;
        const int sgl = w - (N - 1 - b);  // z-step whose front is at bin b
        if (b < N && sgl >= s0 && sgl < s0 + cnt) {
          const int sl = sgl - s0;
          // acc[r][sl] by a select tree (registers cannot be indexed)
          double t[SB];
#pragma unroll
          for (int q = 0; q < SB; ++q) t[q] = acc[r][q];
#pragma unroll
          for (int h = SB / 2; h >= 1; h /= 2)
#pragma unroll
            for (int q = 0; q < h; ++q) t[q] = (sl & h) ? t[q + h] : t[q];
          implicit_post(U2f, Bn[r], dEb[r], t[0], f[r]);
          Wsh[buf * SB + sl] = (U2f[0] * f[r][0] + U2f[1] * f[r][1] + U2f[2] * f[r][2]) / dEb[r];
          if (sgl + 1 < s0 + cnt) {  // flux-independent half of the next z-step at this bin
            const int zi = S - 2 - sgl;
            implicit_pre(U2f, sCz[zi], sNd[zi], sG[b + zi], sA[b + zi], snext[r], dEb[r], Bn[r]);
            if (sgl + 2 < s0 + cnt) snext[r] = G.source[(size_t)(zi - 1) * N + b];
          }
        }
      }
      __syncthreads();
      // refill the slot that update(w-1) was the last to read (every thread is past it now)
      if (w + NSTAGE - 1 <= w_end) issue_column(w + NSTAGE - 1);
      cp_async_commit();
      // rank-1 updates acc[s][b] += T[b+i-1][c] * W_s for the z-steps 0..hi of the band that have started
      // (hi is uniform over the CTA); steps whose front is already below bin b hit the zero padding.
      {
        const int hi = min(w - s0, cnt - 1);
        double Wr[SB];
#pragma unroll
        for (int sl = 0; sl < SB; sl += 2) {
          const double2 w2 = *reinterpret_cast<const double2 *>(Wsh + buf * SB + sl);
          Wr[sl] = w2.x;
          Wr[sl + 1] = w2.y;
        }
#pragma unroll
        for (int r = 0; r < BPT; ++r) {
          const int b = tid + r * nthr;
          if (b < N) {
            const double *tc = Tcol + (w % NSTAGE) * NT + b + (S - 1 - s0);
#pragma unroll
            for (int sl = 0; sl < SB; ++sl) {
              if (sl > hi) break;
              acc[r][sl] = fma(tc[-sl], Wr[sl], acc[r][sl]);
            }
          }
        }
      }
    }
    __syncthreads();
  }

  // bin-width divide and flavour projection (nuSIprop.hpp:329-336) into this point's output slot
  const int q = out_index ? out_index[p] : p;
#pragma unroll
  for (int r = 0; r < BPT; ++r) {
    const int b = tid + r * nthr;
    if (b < N) store_bin(G, flux_mass, flux_fla, q, b, f[r], dEb[r], dead);
  }
  if (status_out && tid == 0) status_out[q] = status_in[p];
}

// ------------------------------------------------------------------------------------------
// resonant-only integrator (non_resonant = false): the regeneration sum collapses to the running
// quantity alpha_cum of nuSIprop.hpp:261-287, a strictly serial recurrence in the bin index.
// One thread per point; scratch = the point's own flux_mass rows in HBM.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64)
integrate_resonant_kernel(DeviceGrid G, int npts, const PointParams *__restrict__ params,
                          const double *__restrict__ rows, const double *__restrict__ sup, size_t tstride,
                          double *__restrict__ work /*[npts][3][N]*/, double *__restrict__ flux_mass,
                          double *__restrict__ flux_fla, const int *__restrict__ out_index,
                          const int *__restrict__ status_in, int *__restrict__ status_out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npts) return;
  const int N = G.N, NT = G.NT;
  const PointParams P = params[p];
  const bool dead = !(P.mphi > 0);
  const double *rowG = rows + ((size_t)p * 2 + 0) * NT;
  const double *rowA = rows + ((size_t)p * 2 + 1) * NT;
  const double *sd = sup + (size_t)p * tstride;
  double *fx = work + (size_t)p * 3 * N;
  const double U2f[3] = {P.U2f[0], P.U2f[1], P.U2f[2]};
  for (int j = 0; j < 3 * N; ++j) fx[j] = 0.0;
  for (int i = G.Nz - 1; i > 0 && !dead; --i) {
    const double cz = G.cz[i - 1], ndz = G.ndz[i - 1];
    double cum[3] = {0, 0, 0};
    for (int j = N; j > 0; --j) {
      const double Gam = ndz * rowG[j + i - 2];
      const double aT = ndz * rowA[j + i - 2];
      if (j != N) {
        const double aw = ndz * sd[j + i - 2];
#pragma unroll
        for (int l = 0; l < 3; ++l) cum[l] += fx[l * N + j] * aw / (G.dE[j]) / (G.dE[j - 1]);
      }
      double src[3], f[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        double s = cz * G.source[(size_t)(i - 1) * N + (j - 1)];
        if (j != N)
#pragma unroll
          for (int l = 0; l < 3; ++l) s += cz * cum[l] * U2f[k] * U2f[l] * (G.dE[j - 1]);
        src[k] = s;
        f[k] = fx[k * N + j - 1];
      }
      implicit_bin(U2f, cz, Gam, aT, G.dE[j - 1], src, f);
#pragma unroll
      for (int k = 0; k < 3; ++k) fx[k * N + j - 1] = f[k];
    }
  }
  const int q = out_index ? out_index[p] : p;
  for (int b = 0; b < N; ++b) {
    const double f[3] = {fx[b], fx[N + b], fx[2 * N + b]};
    store_bin(G, flux_mass, flux_fla, q, b, f, G.dE[b], dead);
  }
  if (status_out) status_out[q] = status_in[p];
}

cudaError_t launch_integrate_serial(const IntegArgs &A, cudaStream_t s) {
  const int N = A.G.N;
  if (N > 256) return cudaErrorInvalidValue;
  integrate_kernel<1><<<A.cnt, ((N + 31) / 32) * 32, 0, s>>>(A.G, A.params, A.rows, A.tri, A.tstride, A.flux_mass, A.flux_fla, A.out_index, A.status_in, A.status_out);
  return cudaGetLastError();
}
cudaError_t launch_integrate_wave(const IntegArgs &A, cudaStream_t s) {
  const DeviceGrid &G = A.G;
  const int N = G.N, NT = G.NT, S = G.Nz - 1;
  cudaError_t e = cudaSuccess;
#define NUSI_WAVE(BPT, SB, NST)                                                                             \
  do {                                                                                                      \
    const size_t sh = (size_t)((NST + 2) * NT + 2 * 32 + 2 * S) * sizeof(double);                           \
    e = cudaFuncSetAttribute(integrate_wave_kernel<BPT, SB, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh); \
    if (e == cudaSuccess)                                                                                   \
      integrate_wave_kernel<BPT, SB, NST><<<A.cnt, 256, sh, s>>>(G, A.params, A.rows, A.tri, A.tstride, A.flux_mass, A.flux_fla, A.out_index, A.status_in, A.status_out); \
  } while (0)
  if (N <= 256 && NT <= 256 * 3)
    NUSI_WAVE(1, 32, 8);
  else if (N <= 512 && NT <= 256 * 4)
    NUSI_WAVE(2, 16, 8);
  else if (N <= 1024 && NT <= 256 * 6)
    NUSI_WAVE(4, 8, 8);
  else if (N <= 2048 && NT <= 256 * 10)
    NUSI_WAVE(8, 4, 6);
  else
    return cudaErrorInvalidValue;
#undef NUSI_WAVE
  return e != cudaSuccess ? e : cudaGetLastError();
}
cudaError_t launch_integrate_resonant(const IntegArgs &A, cudaStream_t s) {
  integrate_resonant_kernel<<<(A.cnt + 63) / 64, 64, 0, s>>>(A.G, A.cnt, A.params, A.rows, A.tri, A.tstride, A.work, A.flux_mass, A.flux_fla, A.out_index, A.status_in, A.status_out);
  return cudaGetLastError();
}

}  // namespace nusi
