// k_rows.cu -- Gamma / alphaTilde rows of the cross-section tables (reference nuSIprop.hpp:759-1235)
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
#include "nusi_launch.h"
#include "nusi_gamma.h"
#include "nusi_alphatilde.h"

namespace nusi {

#ifndef NUSI_ROWS_MINB
#define NUSI_ROWS_MINB 4  // resident CTAs of 192 threads per SM the row kernels are compiled for (80 registers)
#endif

// ------------------------------------------------------------------------------------------
// cross-section builder, straightforward form: one thread per table row / per table entry
// ------------------------------------------------------------------------------------------
// Gamma and alphaTilde rows.  Thread (row r, mass k, which) evaluates one mass term of one of
// the two functions; the channel values go through shared memory and are summed by one thread
// per (row, function) in the reference's order (mass 0 channels, mass 1 channels, ...).
// One launch per function (WHICH 0 Gamma, 1 alphaTilde): every warp of the grid then runs the same
// (large) code body, which keeps it in the instruction cache -- with both functions in one CTA the kernel
// spent most of its time in instruction-fetch stalls.
template <int WHICH>
__global__ void __launch_bounds__(NUSI_ROWS_PER_CTA * 3, NUSI_ROWS_MINB)
rows_kernel(DeviceGrid G, const PointParams *__restrict__ params, double *__restrict__ rows, int *__restrict__ status) {
  __shared__ double ch[NUSI_ROWS_PER_CTA][3][NUSI_ATILDE_NCH + 1];
  __shared__ int nch;
  const int p = blockIdx.y;
  const PointParams P = params[p];
  if (P.mphi <= 0) return;  // point rejected on the host (no mass spectrum)
  const int tid = threadIdx.x;
  const int rl = tid % NUSI_ROWS_PER_CTA, k = tid / NUSI_ROWS_PER_CTA;
  const int i = blockIdx.x * NUSI_ROWS_PER_CTA + rl;
  int warn = 0;
  if (i < G.NT) {
    const PhiPhiTables *pp = G.phiphi ? &G.pp : nullptr;
    const double lo = G.row_lo[i], hi = G.row_hi[i];
    const int n = (WHICH == 0) ? gamma_mass(P, k, lo, hi, ch[rl][k], &warn) : alphatilde_mass(P, pp, k, lo, hi, ch[rl][k], &warn);
    if (rl == 0 && k == 0) nch = n;
  }
  __syncthreads();
  if (i < G.NT && k == 0) {
    double tot = 0;
    const int n = nch;
    for (int kk = 0; kk < 3; ++kk)
      for (int c = 0; c < n; ++c) tot += ch[rl][kk][c];
    rows[((size_t)p * 2 + WHICH) * G.NT + i] = tot;
  }
  if (warn) atomicOr(&status[p], warn);
}

// Gamma / alphaTilde rows of a parameter scan (Majorana, non-resonant, phiphi off): one CTA per (64 rows, GROUP of
// points with the same mphi and neutrino masses).  A thread = (row, target mass) evaluates the part of its closed
// forms that does not contain g once (gamma_mass_shared / alphatilde_mass_shared: every real dilogarithm, atanh
// and most logarithms), then loops over the points of the group with the per-point rest (Breit-Wigner terms,
// complex dilogarithms, args).  Same bits as rows_kernel (split functions, asserted in the tests).
template <int WHICH>
__global__ void __launch_bounds__(NUSI_ROWS_PER_CTA * 3, NUSI_ROWS_MINB)
rows_group_kernel(DeviceGrid G, const PointParams *__restrict__ params, const int2 *__restrict__ groups,
                  double *__restrict__ rows, int *__restrict__ status) {
  constexpr int NCH = (WHICH == 0) ? NUSI_GAMMA_NCH : NUSI_ATILDE_NCH;
  __shared__ double ch[NUSI_ROWS_PER_CTA][3][NUSI_ATILDE_NCH + 1];
  const int2 grp = __ldg(groups + blockIdx.y);
  const int p0 = grp.x, glen = grp.y;
  const double mphi0 = params[p0].mphi;
  if (mphi0 <= 0) return;  // rejected points form groups of their own
  const int tid = threadIdx.x;
  const int rl = tid % NUSI_ROWS_PER_CTA, k = tid / NUSI_ROWS_PER_CTA;
  const int i = blockIdx.x * NUSI_ROWS_PER_CTA + rl;
  const bool act = i < G.NT;
  // blockIdx.z of gridDim.z CTAs share a group (more CTAs than groups x row blocks: the per-point part dominates,
  // so the g-independent part is cheap to repeat a few times)
  const int per = (glen + gridDim.z - 1) / gridDim.z;
  const int j0 = blockIdx.z * per, j1 = min(glen, j0 + per);
  if (j0 >= j1) return;
  GammaShared SG;
  ATildeShared SA;
  if (act) {
    const double lo = G.row_lo[i], hi = G.row_hi[i], mnk = params[p0].mn[k];
    if (WHICH == 0)
      gamma_mass_shared(mphi0, mnk, lo, hi, SG);
    else
      alphatilde_mass_shared(mphi0, mnk, lo, hi, SA);
  }
#pragma unroll 1
  for (int j = j0; j < j1; ++j) {
    const int p = p0 + j;
    int warn = 0;
    if (act) {
      const PointParams P = params[p];
      if (WHICH == 0)
        gamma_mass_point(P, k, SG, ch[rl][k], &warn);
      else
        alphatilde_mass_point(P, k, SA, ch[rl][k], &warn);
    }
    __syncthreads();
    if (act && k == 0) {
      double tot = 0;
      for (int kk = 0; kk < 3; ++kk)
        for (int c = 0; c < NCH; ++c) tot += ch[rl][kk][c];
      rows[((size_t)p * 2 + WHICH) * G.NT + i] = tot;
    }
    if (warn) atomicOr(&status[p], warn);
    __syncthreads();
  }
}

// The same in two kernels (product path of a scan): the g-independent part of a (row, target mass) ONCE per group into a small
// global array, then the per-point part with as many CTAs per group as fill the GPU -- rows_group_kernel repeats the shared
// part in every one of its `split` CTAs (about half of its instructions at 8 points per CTA) and carries both code bodies
// through the instruction cache.  Same functions, same operands, same bits.
static_assert(sizeof(GammaShared) <= NUSI_ROWSH * sizeof(double) && sizeof(ATildeShared) <= NUSI_ROWSH * sizeof(double), "NUSI_ROWSH too small");
template <int WHICH>
__global__ void __launch_bounds__(NUSI_ROWS_PER_CTA * 3, NUSI_ROWS_MINB)
rows_shared_kernel(DeviceGrid G, const PointParams *__restrict__ params, const int2 *__restrict__ groups, double *__restrict__ rowsh) {
  const int2 grp = __ldg(groups + blockIdx.y);
  const int p0 = grp.x;
  const double mphi0 = params[p0].mphi;
  if (mphi0 <= 0) return;
  const int tid = threadIdx.x, rl = tid % NUSI_ROWS_PER_CTA, k = tid / NUSI_ROWS_PER_CTA;
  const int i = blockIdx.x * NUSI_ROWS_PER_CTA + rl;
  if (i >= G.NT) return;
  const double lo = G.row_lo[i], hi = G.row_hi[i], mnk = params[p0].mn[k];
  double *o = rowsh + (((size_t)blockIdx.y * G.NT + i) * 3 + k) * NUSI_ROWSH;
  if (WHICH == 0) {
    GammaShared SG;
    gamma_mass_shared(mphi0, mnk, lo, hi, SG);
    *reinterpret_cast<GammaShared *>(o) = SG;
  } else {
    ATildeShared SA;
    alphatilde_mass_shared(mphi0, mnk, lo, hi, SA);
    *reinterpret_cast<ATildeShared *>(o) = SA;
  }
}
template <int WHICH>
__global__ void __launch_bounds__(NUSI_ROWS_PER_CTA * 3, NUSI_ROWS_MINB)
rows_point_kernel(DeviceGrid G, const PointParams *__restrict__ params, const int2 *__restrict__ groups, const double *__restrict__ rowsh,
                  double *__restrict__ rows, int *__restrict__ status) {
  constexpr int NCH = (WHICH == 0) ? NUSI_GAMMA_NCH : NUSI_ATILDE_NCH;
  __shared__ double ch[NUSI_ROWS_PER_CTA][3][NUSI_ATILDE_NCH + 1];
  const int2 grp = __ldg(groups + blockIdx.y);
  const int p0 = grp.x, glen = grp.y;
  if (params[p0].mphi <= 0) return;
  const int tid = threadIdx.x, rl = tid % NUSI_ROWS_PER_CTA, k = tid / NUSI_ROWS_PER_CTA;
  const int i = blockIdx.x * NUSI_ROWS_PER_CTA + rl;
  const bool act = i < G.NT;
  const int per = (glen + gridDim.z - 1) / gridDim.z;
  const int j0 = blockIdx.z * per, j1 = min(glen, j0 + per);
  if (j0 >= j1) return;
  GammaShared SG;
  ATildeShared SA;
  if (act) {
    const double *o = rowsh + (((size_t)blockIdx.y * G.NT + i) * 3 + k) * NUSI_ROWSH;
    if (WHICH == 0)
      SG = *reinterpret_cast<const GammaShared *>(o);
    else
      SA = *reinterpret_cast<const ATildeShared *>(o);
  }
#pragma unroll 1
  for (int j = j0; j < j1; ++j) {
    const int p = p0 + j;
    int warn = 0;
    if (act) {
      const PointParams P = params[p];
      if (WHICH == 0)
        gamma_mass_point(P, k, SG, ch[rl][k], &warn);
      else
        alphatilde_mass_point(P, k, SA, ch[rl][k], &warn);
    }
    __syncthreads();
    if (act && k == 0) {
      double tot = 0;
      for (int kk = 0; kk < 3; ++kk)
        for (int c = 0; c < NCH; ++c) tot += ch[rl][kk][c];
      rows[((size_t)p * 2 + WHICH) * G.NT + i] = tot;
    }
    if (warn) atomicOr(&status[p], warn);
    __syncthreads();
  }
}

cudaError_t launch_rows(const BuildArgs &A, bool grouped, int split, cudaStream_t s) {
  const DeviceGrid &G = A.G;
  const unsigned nblk = (G.NT + NUSI_ROWS_PER_CTA - 1) / NUSI_ROWS_PER_CTA;
  if (grouped && A.rowsh && A.ngroups <= A.rowsh_groups) {
    dim3 gs(nblk, A.ngroups), gp(nblk, A.ngroups, split);
    rows_shared_kernel<0><<<gs, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.groups, A.rowsh);
    rows_point_kernel<0><<<gp, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.groups, A.rowsh, A.rows, A.status);
    rows_shared_kernel<1><<<gs, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.groups, A.rowsh);
    rows_point_kernel<1><<<gp, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.groups, A.rowsh, A.rows, A.status);
  } else if (grouped) {
    dim3 gg(nblk, A.ngroups, split);
    rows_group_kernel<0><<<gg, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.groups, A.rows, A.status);
    rows_group_kernel<1><<<gg, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.groups, A.rows, A.status);
  } else {
    dim3 gr(nblk, A.cnt);
    rows_kernel<0><<<gr, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.rows, A.status);
    rows_kernel<1><<<gr, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, A.params, A.rows, A.status);
  }
  return cudaGetLastError();
}

}  // namespace nusi
