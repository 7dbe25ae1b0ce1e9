// nusi_launch.h -- host-callable launchers of the CUDA kernels.  The kernels live in several translation units
// (k_*.cu: they compile in parallel and none of them needs the others' code); nusi_capi.cu holds the host pipeline
// and calls them through these functions only.
#ifndef NUSI_LAUNCH_H
#define NUSI_LAUNCH_H
#include <cuda_runtime.h>
#include "nusi_device.h"

namespace nusi {

// cross-section builder: one launch covers `cnt` points (slot i of rows / tri / edges belongs to params[i])
struct BuildArgs {
  DeviceGrid G;
  int nE;                     // distinct bin-edge energies of the table cells
  const PointParams *params;  // [cnt]
  int cnt;
  int *status;                // [cnt] NUSI_WARN_* bits, OR'ed
  double *rows;               // [cnt][2][NT]
  double *tri;                // [cnt][tstride]
  size_t tstride;
  double *edges;              // [cnt][edge_block_doubles]
  const double *edgeE;        // [nE]
  const int2 *tiles;          // alpha_tile_kernel: (row tile, column tile)
  int n_tiles;
  const int4 *ctiles;         // compact / scan builders: cells [a0, a1) x [b0, b1)
  int n_ctiles;
  const int2 *groups;         // scan builders: (first point, length) of runs with identical (mphi, m_nu)
  int ngroups;
  double *rowsh;              // scan rows: [rowsh_groups][NT][3][NUSI_ROWSH] g-independent part per (group, row, mass), or NULL
  int rowsh_groups;
};

// k_rows.cu -- Gamma / alphaTilde rows (nuSIprop.hpp:759-1235)
cudaError_t launch_rows(const BuildArgs &A, bool grouped, int split, cudaStream_t s);

// k_alpha.cu -- alpha table, product builders (nuSIprop.hpp:1237-1520)
cudaError_t alpha_setup_attributes();
cudaError_t launch_edges(const BuildArgs &A, cudaStream_t s);
cudaError_t launch_edges_group(const BuildArgs &A, cudaStream_t s);
cudaError_t launch_alpha_group(const BuildArgs &A, cudaStream_t s);
cudaError_t launch_alpha_compact(const BuildArgs &A, cudaStream_t s);
size_t alpha_group_smem_bytes();

// k_alpha_check.cu -- strict / per-entry builders (cross-checks, Dirac, resonant-only, phi-phi)
cudaError_t alpha_check_setup_attributes();
cudaError_t launch_alpha_tile(const BuildArgs &A, int mode, cudaStream_t s);
cudaError_t launch_alpha_entry(const BuildArgs &A, int compact, cudaStream_t s);

// z-stepping integrators (nuSIprop.hpp:255-336).  Point i of the launch writes its fluxes to output slot
// out_index[i] (NULL: slot i) and, when status_out is given, copies its status word status_in[i] there too.
struct IntegArgs {
  DeviceGrid G;
  const PointParams *params;
  int cnt;
  const double *rows, *tri;
  size_t tstride;
  double *work;          // resonant-only kernel: [cnt][3][N] scratch
  double *flux_mass, *flux_fla;  // output arrays, slot-indexed (either may be NULL)
  const int *out_index;  // [cnt] or NULL
  const int *status_in;  // [cnt] or NULL
  int *status_out;       // slot-indexed or NULL
};
// k_integrate.cu -- product kernel (diagonal-owned wavefront); returns cudaErrorInvalidValue when the binning is
// outside what the kernel family supports (N_bins_E <= 2041)
cudaError_t launch_integrate_diag(const IntegArgs &A, int fine_variant, cudaStream_t s);
bool integrate_diag_supports(int N, int NT);
// k_integrate_split.cu -- product kernel for 32 <= N_bins_E <= 576: role-split wavefront (critical warp / helper / bulk warps, TMA ring)
cudaError_t launch_integrate_split(const IntegArgs &A, int variant, cudaStream_t s);
bool integrate_split_supports(int N, int NT);
// k_integrate_cluster.cu -- product kernel for fine binnings (576 < N_bins_E <= 2017): the role-split wavefront with one point
// spread over a thread-block cluster of 4 CTAs
cudaError_t launch_integrate_cluster(const IntegArgs &A, int variant, cudaStream_t s);
bool integrate_cluster_supports(int N, int NT);
// k_integrate_check.cu -- bin-owned wavefront, serial reference-shaped, resonant-only
cudaError_t launch_integrate_wave(const IntegArgs &A, cudaStream_t s);
cudaError_t launch_integrate_serial(const IntegArgs &A, cudaStream_t s);
cudaError_t launch_integrate_resonant(const IntegArgs &A, cudaStream_t s);

// k_misc.cu
cudaError_t launch_source(int N, int S, const double *zsrc, const double *rsn, double pref, const double *Emin, const double *Emax,
                          double *out, cudaStream_t s);
cudaError_t launch_specfun(int which, int n, const double *x, const double *y, double *re, double *im);
cudaError_t launch_phiphi_table(int which, long long first, long long count, int n1, int n2, const double *x0, const double *x1,
                                const double *x2, float *out);
cudaError_t launch_fp64_peak(int blocks, int threads, double *out, int iters);
// sum_bins sum_k (ln Emax - ln Emin) E^2 flux_k  per point: the energy integral of check_energy_conservation (nuSIprop.hpp:351-355)
cudaError_t launch_energy_integral(int n, int N, const double *flux_mass, const double *wE2, double *out, cudaStream_t s);

}  // namespace nusi
#endif
