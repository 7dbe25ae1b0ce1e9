// k_integrate.cu -- the implicit z-stepping integrator (reference nuSIprop.hpp:255-336): product kernel
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
#include "nusi_launch.h"
#include "nusi_implicit.h"
#include "nusi_cpasync.h"

namespace nusi {

#ifndef NUSI_WAVE_MINB
#define NUSI_WAVE_MINB 2
#endif
#ifndef NUSI_DIAG_SB_SMALL
#define NUSI_DIAG_SB_SMALL 32  // z-steps per band for N_bins_E <= 225
#endif

// ------------------------------------------------------------------------------------------
// z-stepping integrator, wavefront form with DIAGONAL-owned accumulators (default).
//
// Same wavefront as integrate_wave_kernel, different ownership of the partial regeneration sums.
// Cell (z-step sl of the band, bin b) reads table row  b - sl + const  of the wave's column: the row depends
// only on the diagonal d = b - sl.  So the thread that owns diagonal d needs ONE table value per wave and
// multiplies it with the SB broadcast W's of the wave (SB FMAs per shared-memory load instead of one), and --
// because all front cells of wave w lie on the single diagonal d_w = N-1-(w-s0) -- the owner of d_w hands its SB
// finished sums to the bin owners through shared memory; no register array is ever indexed dynamically.
// Per wave: publish (1 thread) | barrier | finalise <= SB cells (bin owners) | barrier | update (live diagonals).
// The accumulation order of every cell is unchanged (source bins from high to low), so the fluxes equal
// those of integrate_kernel / integrate_wave_kernel bit for bit.
// ------------------------------------------------------------------------------------------
template <int BPT, int SB, int DPT, int NSTAGE, int NTHR = 256, int PREQ = 0>
__global__ void __launch_bounds__(NTHR, ((BPT == 1 && DPT == 1 && NTHR == 256) ? NUSI_WAVE_MINB : 1))
integrate_diag_kernel(DeviceGrid G, const PointParams *__restrict__ params, const double *__restrict__ rows,
                      const double *__restrict__ tri, size_t tstride, double *__restrict__ flux_mass,
                      double *__restrict__ flux_fla, const int *__restrict__ out_index, const int *__restrict__ status_in,
                      int *__restrict__ status_out) {
  extern __shared__ __align__(16) double wsm[];
#ifndef NUSI_TID_PLAIN  // (5.72 -> 5.63 ms per 8192 points)
  // read once through an opaque move: otherwise the compiler re-reads the special register (S2R, ~20 cycles) in every
  // wave instead of keeping the value in a register
  int tid;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
  const int p = blockIdx.x;
#else
  const int p = blockIdx.x, tid = threadIdx.x;
#endif
  constexpr int nthr = NTHR;  // == blockDim.x (power of two)
  const int N = G.N, NT = G.NT, S = G.Nz - 1;
  const PointParams P = params[p];
  const bool dead = !(P.mphi > 0);
  double *Tcol = wsm;                           // [NSTAGE][NT] ring of table columns
  double *Ash = Tcol + NSTAGE * NT;             // [SB] finished sums of the wave's diagonal
  double *Wsh = Ash + SB;                       // [SB] weighted fluxes of the wave's front cells
  double *Pre = Wsh + SB;                       // [2][SB][NUSI_PRE_DOUBLES] prepared half-steps, by wave parity
  double *sG = Pre + 2 * SB * NUSI_PRE_DOUBLES;  // [NT]  Gamma rows
  double *sA = sG + NT;                         // [NT]  alphaTilde rows
  double *sCz = sA + NT;                        // [S]
  double *sNd = sCz + S;                        // [S]
  double *sdE = sNd + S;                        // [N]
  double *sRdE = sdE + N;                       // [N]  1/dE (IEEE)
  const double U2f[3] = {P.U2f[0], P.U2f[1], P.U2f[2]};
  constexpr int PRE = PREQ ? PREQ : BPT + 2;  // table rows per thread and column; NT <= NTHR*PRE is checked on the host
                                              // (PREQ = 1: NT <= NTHR, one cp.async per thread and wave)
  // Both are derived from special registers (blockIdx, the shared window); made opaque so that the compiler keeps
  // them in registers instead of re-reading the special registers (S2UR, ~30 cycles each) in every wave.
  const double *T;
  unsigned tcol_s;
  asm volatile("mov.u64 %0, %1;" : "=l"(T) : "l"(tri + (size_t)p * tstride));
  asm volatile("mov.u32 %0, %1;" : "=r"(tcol_s) : "r"((unsigned)__cvta_generic_to_shared(Tcol)));

  for (int i = tid; i < NT; i += nthr) {
    sG[i] = rows[((size_t)p * 2 + 0) * NT + i];
    sA[i] = rows[((size_t)p * 2 + 1) * NT + i];
  }
  for (int i = tid; i < S; i += nthr) {
    sCz[i] = G.cz[i];
    sNd[i] = G.ndz[i];
  }
  for (int i = tid; i < N; i += nthr) {
    sdE[i] = G.dE[i];
    sRdE[i] = 1.0 / G.dE[i];
  }
  double f[BPT][3], dEb[BPT], rdE[BPT];
#pragma unroll
  for (int r = 0; r < BPT; ++r) {
    const int b = tid + r * nthr;
    f[r][0] = f[r][1] = f[r][2] = 0.0;
    dEb[r] = (b < N) ? G.dE[b] : 1.0;
    rdE[r] = 1.0 / dEb[r];
  }
  __syncthreads();
  const int htid = (tid + NTHR / 2) & (NTHR - 1);  // this thread prepares the half-steps of bins htid + r*nthr (never its own front cells)

  for (int s0 = 0; s0 < S && !dead; s0 += SB) {
    const int cnt = min(SB, S - s0);
    double acc[DPT][SB];  // sums of diagonal d = tid + r*nthr - (SB-1): cell (sl, b = d + sl)
#pragma unroll
    for (int r = 0; r < DPT; ++r)
#pragma unroll
      for (int sl = 0; sl < SB; ++sl) acc[r][sl] = 0.0;
    const int w_begin = s0, w_end = s0 + cnt - 1 + N - 1;
    auto issue_column = [&](int w) {  // column c = NT-1-w into ring slot w % NSTAGE, rows >= c zero-filled
      const int c = NT - 1 - w;
      const double *col = T + (size_t)c * (c - 1) / 2;
      const unsigned dst = tcol_s + (unsigned)((w % NSTAGE) * NT + tid) * 8u;
#pragma unroll
      for (int q = 0; q < PRE; ++q) {
        const int a = tid + q * nthr;
        if (a < NT) cp_async8s(dst + (unsigned)(q * nthr) * 8u, col + (a < c ? a : 0), a < c);
      }
    };
#pragma unroll
    for (int j = 0; j < NSTAGE - 1; ++j) {
      if (w_begin + j <= w_end) issue_column(w_begin + j);
      cp_async_commit();
    }
    // Helper role: prepare the flux-independent half of every cell that is finalised in wave wn, into Pre[wn & 1].
    // The source term of a helped bin's next z-step is loaded one wave ahead (hsrc), the first one two waves ahead.
    double hsrc[BPT];
#pragma unroll
    for (int r = 0; r < BPT; ++r) hsrc[r] = 0.0;
    auto help = [&](int wn) {
#pragma unroll
      for (int r = 0; r < BPT; ++r) {
        const int bh = htid + r * nthr;
        const int sgn = wn - (N - 1 - bh);  // z-step of bin bh that is finalised in wave wn
        if (bh < N && sgn >= s0 - 1 && sgn < s0 + cnt) {
          const int zi = S - 1 - sgn;
          if (sgn >= s0)
            pre_to_slot(U2f, sCz[zi], sNd[zi], sG[bh + zi], sA[bh + zi], hsrc[r], sdE[bh], sRdE[bh],
                        Pre + ((wn & 1) * SB + (sgn - s0)) * NUSI_PRE_DOUBLES);
          if (sgn + 1 < s0 + cnt) hsrc[r] = G.source[(size_t)(zi - 1) * N + bh];
        }
      }
    };
    // prologue: the bin whose first cell is in wave w_begin needs its source term now
#pragma unroll
    for (int r = 0; r < BPT; ++r) {
      const int bh = htid + r * nthr;
      if (bh < N && w_begin - (N - 1 - bh) == s0) hsrc[r] = G.source[(size_t)(S - 1 - s0) * N + bh];
    }
    help(w_begin);
    for (int w = w_begin; w <= w_end; ++w) {
      const int dw = N - 1 - (w - s0);  // the diagonal whose cells are finalised in this wave
      // publish: the owner of diagonal dw hands over its sums
#pragma unroll
      for (int r = 0; r < DPT; ++r)
        if (tid + r * nthr - (SB - 1) == dw) {
#pragma unroll
          for (int sl = 0; sl < SB; sl += 2) *reinterpret_cast<double2 *>(Ash + sl) = make_double2(acc[r][sl], acc[r][sl + 1]);
        }
      cp_async_wait<NSTAGE - 2>();  // this thread's part of column w has landed
      __syncthreads();
      // finalise this thread's cells of the wave from the prepared slots
#pragma unroll
      for (int r = 0; r < BPT; ++r) {
        const int b = tid + r * nthr;
        const int sgl = w - (N - 1 - b);  // z-step whose front is at bin b
        if (b < N && sgl >= s0 && sgl < s0 + cnt) {
          const int sl = sgl - s0;
          Wsh[sl] = implicit_post_smem(U2f, Pre + ((w & 1) * SB + sl) * NUSI_PRE_DOUBLES, dEb[r], rdE[r], Ash[sl], f[r]);
        }
      }
      // meanwhile the helpers prepare the next wave's cells
      if (w < w_end) help(w + 1);
      __syncthreads();
      // refill the ring slot that update(w-1) was the last to read
      if (w + NSTAGE - 1 <= w_end) issue_column(w + NSTAGE - 1);
      cp_async_commit();
      // update the live diagonals d < dw with the z-steps 0..hi that have started:
      //   acc[d][sl] += T[d + S-1-s0][c] * W_sl       (cell (sl, b = d+sl); source bin dw + sl)
      {
        const int hi = min(w - s0, cnt - 1);
        const double *tc = Tcol + (w % NSTAGE) * NT + (S - 1 - s0);
#pragma unroll
        for (int r = 0; r < DPT; ++r) {
          const int d = tid + r * nthr - (SB - 1);
          if (d < dw) {
            const double tv = tc[d];
            if (hi == SB - 1) {
#pragma unroll
              for (int sl = 0; sl < SB; sl += 2) {
                const double2 w2 = *reinterpret_cast<const double2 *>(Wsh + sl);
                acc[r][sl] = fma(tv, w2.x, acc[r][sl]);
                acc[r][sl + 1] = fma(tv, w2.y, acc[r][sl + 1]);
              }
            } else {
#pragma unroll
              for (int sl = 0; sl < SB; ++sl)
                if (sl <= hi) acc[r][sl] = fma(tv, Wsh[sl], acc[r][sl]);
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

// shapes the kernel family covers: N <= 256 BPT bins, N + SB - 1 <= NTHR DPT diagonals, NT <= NTHR PRE table rows
bool integrate_diag_supports(int N, int NT) { return (N <= 2048 && NT <= 256 * 10) || (N + 8 - 1 <= 512 * 4 && NT <= 512 * 6); }

cudaError_t launch_integrate_diag(const IntegArgs &A, int fine_variant, cudaStream_t s) {
  const DeviceGrid &G = A.G;
  const int N = G.N, NT = G.NT, S = G.Nz - 1, cnt = A.cnt;
  cudaError_t e = cudaSuccess;
#define NUSI_DIAG_T(BPT, SB, DPT, NST, NTHR, PREQ)                                                          \
  do {                                                                                                      \
    const size_t sh = (size_t)((NST + 2) * NT + 2 * SB + 2 * SB * NUSI_PRE_DOUBLES + 2 * S + 2 * N + 2) * sizeof(double); \
    e = cudaFuncSetAttribute(integrate_diag_kernel<BPT, SB, DPT, NST, NTHR, PREQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh); \
    if (e == cudaSuccess)                                                                                   \
      integrate_diag_kernel<BPT, SB, DPT, NST, NTHR, PREQ><<<cnt, NTHR, sh, s>>>(G, A.params, A.rows, A.tri, A.tstride, A.flux_mass, A.flux_fla, A.out_index, A.status_in, A.status_out); \
  } while (0)
  // More threads per point for finer binning = fewer bins and diagonals per thread, so wide bands still fit the
  // registers and the table is streamed fewer times (bands x waves is what a latency-bound point pays for).
  // fine_variant 1 selects the 256-thread configurations throughout (kept for comparison; same bits).
  const bool wide = fine_variant == 0;
  if (N <= 225 && NT <= 256)  // a table column fits one row per thread (N_E = 200: NT = 231)
    NUSI_DIAG_T(1, NUSI_DIAG_SB_SMALL, 1, 8, 256, 1);
  else if (N <= 225 && NT <= 256 * 3)
    NUSI_DIAG_T(1, NUSI_DIAG_SB_SMALL, 1, 8, 256, 0);
  else if (N <= 256 && NT <= 256 * 3)
    NUSI_DIAG_T(1, 32, 2, 8, 256, 0);
  else if (wide && N + 32 - 1 <= 512 && NT <= 512 * 3)
    NUSI_DIAG_T(1, 32, 1, 8, 512, 0);
  else if (N <= 512 && NT <= 256 * 4)
    NUSI_DIAG_T(2, 16, 3, 8, 256, 0);
  else if (wide && N <= 1024 && NT <= 512 * 4)
    NUSI_DIAG_T(2, 16, 3, 8, 512, 0);
  else if (N <= 1024 && NT <= 256 * 6)
    NUSI_DIAG_T(4, 8, 5, 8, 256, 0);
  else if (wide && N + 8 - 1 <= 512 * 4 && NT <= 512 * 6)
    NUSI_DIAG_T(4, 8, 4, 6, 512, 0);
  else if (N <= 2048 && NT <= 256 * 10)
    NUSI_DIAG_T(8, 4, 9, 6, 256, 0);
  else
    return cudaErrorInvalidValue;
#undef NUSI_DIAG_T
  return e != cudaSuccess ? e : cudaGetLastError();
}

}  // namespace nusi
