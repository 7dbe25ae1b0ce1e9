// k_alpha_check.cu -- alpha table: strict tiled builder (modes 0/1) and the per-entry builder (Dirac, resonant-only, phi-phi; cross-checks)
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
#include "nusi_launch.h"
#include "nusi_alpha.h"

namespace nusi {

// packed-by-column index -> (row a, column b), a < b
__device__ __forceinline__ void tri_unpack(long long idx, int &a, int &b) {
  int c = (int)((1.0 + sqrt(1.0 + 8.0 * (double)idx)) * 0.5);
  while ((long long)c * (c - 1) / 2 > idx) --c;
  while ((long long)(c + 1) * c / 2 <= idx) ++c;
  b = c;
  a = (int)(idx - (long long)c * (c - 1) / 2);
}

__global__ void __launch_bounds__(128)
alpha_kernel(DeviceGrid G, const PointParams *__restrict__ params, double *__restrict__ tri, size_t tstride,
             int *__restrict__ status, int compact) {
  const int p = blockIdx.y;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const PointParams P = params[p];
  if (P.mphi <= 0) return;
  int a, b;
  if (G.non_resonant) {
    if (idx >= (long long)G.NT * (G.NT - 1) / 2) return;
    tri_unpack(idx, a, b);
  } else {  // resonant-only: the integrator reads the super-diagonal alone (nuSIprop.hpp:275)
    if (idx >= G.NT - 1) return;
    a = (int)idx;
    b = a + 1;
  }
  int warn = 0;
  const PhiPhiTables *pp = G.phiphi ? &G.pp : nullptr;
  // compact: the per-entry evaluation of what alpha_compact_kernel computes (cross-check of its tiling)
  tri[(size_t)p * tstride + idx] = (compact && P.majorana && P.non_resonant)
                                       ? alpha_entry_compact<true>(P, pp, G.row_lo[a], G.row_hi[a], G.row_lo[b], G.row_hi[b], &warn)
                                       : alpha_entry(P, pp, G.row_lo[a], G.row_hi[a], G.row_lo[b], G.row_hi[b], &warn);
  if (warn) atomicOr(&status[p], warn);
}

// ------------------------------------------------------------------------------------------
// cross-section builder, tiled form (Majorana + non-resonant: the expensive case).
//
// All bins are cells of one edge list.  Main-grid cells c < N share edges with their neighbours
// bit for bit (Emax[c] and Emin[c+1] are the same pow() call); the blue-shifted cells c >= N
// carry their own two edges, as the reference computes them (nuSIprop.hpp:229-233):
//     edge id of cell c:  lo = c, hi = c+1            (c <  N)
//                         lo = N+1+2(c-N), hi = lo+1  (c >= N)
// One CTA builds a TS x TS tile of table entries for one parameter point.  Per target mass k it
//   (1) evaluates the Vertex data (all dilogarithms, ln(1+s+t), arg) once per (row edge, column
//       edge) pair of the tile and the RowEdge / ColEdge data once per edge, into shared memory;
//   (2) lets each thread combine 4 vertices + 4 edges into its table entry (alpha_cell_mass),
//       accumulating over k in a register in the reference's order.
// A vertex is shared by up to 4 entries, so the transcendental work drops ~4x against the
// one-thread-per-entry kernel, and each big code block exists once (instruction cache).
// ------------------------------------------------------------------------------------------
#ifndef NUSI_TILE_MINB
#define NUSI_TILE_MINB 2
#endif
#define NUSI_VSTRIDE (NUSI_MAXE + 1)  // odd row stride: the 15-row cell pattern stays bank-conflict free
struct TileSmem {
  double V[NUSI_VF][NUSI_MAXE][NUSI_VSTRIDE];  // [field][col edge][row edge]     (one mass at a time)
  double LR[NUSI_MAXE][NUSI_TS + 1];           // [col edge][row cell]  t-channel log on a horizontal cell edge
  double LY[NUSI_TS][NUSI_VSTRIDE];            // [col cell][row edge]  t-channel log on a vertical cell edge
  double R[3][NUSI_RF][NUSI_MAXE];             // edge data of all three masses, loaded once
  double C[3][NUSI_CF][NUSI_MAXE];
  double fs[3][NUSI_FF][NUSI_TS + 1];
};

// MODE 0: strict (make_vertex: every logarithm / argument evaluated where the reference does)
// MODE 1: vertices through shared logarithms (make_vertex_shared), cell ratio logs evaluated
// (a MODE 2 that also shared the t-channel ratio logs of each entry was measured: 13 % faster, but flux
//  parity fell from 89 % to 83 % of bins within 1e-9 on the S-scan sample, so it is not built)
template <int MODE>
__global__ void __launch_bounds__(256, NUSI_TILE_MINB)
alpha_tile_kernel(DeviceGrid G, int nE, const PointParams *__restrict__ params, const double *__restrict__ edgeE,
                  const double *__restrict__ edges, const int2 *__restrict__ tiles, double *__restrict__ tri,
                  size_t tstride, int *__restrict__ status) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TileSmem &S = *reinterpret_cast<TileSmem *>(smem_raw);
  const int p = blockIdx.y, tid = threadIdx.x;
  const PointParams P = params[p];
  if (P.mphi <= 0) return;
  const int N = G.N, NT = G.NT;
  const int2 tl = tiles[blockIdx.x];
  const int a0 = tl.x * NUSI_TS, b0 = tl.y * NUSI_TS;
  const int a1 = min(a0 + NUSI_TS, NT), b1 = min(b0 + NUSI_TS, NT);
  const int ncr = a1 - a0, ncc = b1 - b0;
  const int r0 = edge_lo_id(a0, N), c0 = edge_lo_id(b0, N);
  const int nre = edge_hi_id(a1 - 1, N) - r0 + 1, nce = edge_hi_id(b1 - 1, N) - c0 + 1;
  const bool diag = (tl.x == tl.y);
  const double gr = P.Ga / P.mphi;
  const PhiPhiTables *pp = G.phiphi ? &G.pp : nullptr;
  const double *ebase = edges + (size_t)p * edge_block_doubles(nE, NT);

  // edge data of this tile, all three masses, from the per-point edge block (one global pass)
  {
    const int e = tid & 31, f = tid >> 5;  // 8 fields x 32 edge slots
    for (int k = 0; k < 3; ++k) {
      const double *gR = ebase + (size_t)k * NUSI_RF * nE;
      const double *gC = ebase + (size_t)3 * NUSI_RF * nE + (size_t)k * NUSI_CF * nE;
      if (e < nre) S.R[k][f][e] = gR[(size_t)f * nE + r0 + e];
      if (e < nce) S.C[k][f][e] = gC[(size_t)f * nE + c0 + e];
    }
    if (tid < 3 * NUSI_FF * 16) {
      const int cc = tid & 15, kf = tid >> 4;  // kf = k*NUSI_FF + field
      if (cc < ncc) S.fs[kf / NUSI_FF][kf % NUSI_FF][cc] = ebase[(size_t)3 * (NUSI_RF + NUSI_CF) * nE + (size_t)kf * NT + b0 + cc];
    }
  }
  __syncthreads();

  // this thread's table entry
  const int ca = a0 + tid % ncr, cb = b0 + tid / ncr;
  const bool have_cell = (tid < ncr * ncc) && (ca < cb);
  const int rm = edge_lo_id(ca, N) - r0, rp = edge_hi_id(ca, N) - r0;
  const int cm = edge_lo_id(cb, N) - c0, cp = edge_hi_id(cb, N) - c0;
  double tot = 0;
  int warn = 0;

  for (int k = 0; k < 3; ++k) {
    const double mn = P.mn[k];
    // (1) vertices, and the t-channel logarithms that live on cell edges
    for (int v = tid; v < nre * nce; v += blockDim.x) {
      const int rl = v % nre, cl = v / nre;
      bool need = true;
      if (diag) need = max(edge_cell_min(r0 + rl, N), a0) < min(edge_cell_max(c0 + cl, N), b1 - 1);
      if (need) {
        Vertex X;
        if (MODE == 0) {
          make_vertex(s_of(mn, edgeE[c0 + cl], P.mphi), t_of(mn, edgeE[r0 + rl], P.mphi), gr, X);
        } else {
          RowEdge Re;
          ColEdge Ce;
          Re.t = S.R[k][0][rl]; Re.l1m = S.R[k][1][rl]; Re.labs = S.R[k][2][rl]; Re.lnD = S.R[k][6][rl]; Re.phi = S.R[k][7][rl];
          Ce.s = S.C[k][0][cl]; Ce.ls = S.C[k][1][cl]; Ce.args = S.C[k][3][cl]; Ce.l2 = S.C[k][6][cl]; Ce.lnN = S.C[k][7][cl];
          make_vertex_shared(Re, Ce, gr, X);
        }
        S.V[0][cl][rl] = X.l1;
        S.V[1][cl][rl] = X.lf;
        S.V[2][cl][rl] = X.dF;
        S.V[3][cl][rl] = X.lq;
        S.V[4][cl][rl] = X.d2;
        S.V[5][cl][rl] = X.zr_re;
        S.V[6][cl][rl] = X.zr_im;
        S.V[7][cl][rl] = X.zc_re;
        S.V[8][cl][rl] = X.zc_im;
        S.V[9][cl][rl] = X.ca;
      }
    }
    for (int q = tid; q < ncr * nce; q += blockDim.x) {  // LR(row cell, column edge)
      const int al = q % ncr, cl = q / ncr, a = a0 + al;
      if (!diag || a < min(edge_cell_max(c0 + cl, N), b1 - 1))
        S.LR[cl][al] = alpha_log_LR(S.C[k][0][cl], S.R[k][0][edge_lo_id(a, N) - r0], S.R[k][0][edge_hi_id(a, N) - r0]);
    }
    for (int q = tid; q < nre * ncc; q += blockDim.x) {  // LY(row edge, column cell)
      const int rl = q % nre, cc = q / nre, b = b0 + cc;
      if (!diag || max(edge_cell_min(r0 + rl, N), a0) < b)
        S.LY[cc][rl] = alpha_log_LY(S.C[k][0][edge_lo_id(b, N) - c0], S.C[k][0][edge_hi_id(b, N) - c0], S.R[k][0][rl]);
    }
    __syncthreads();
    // (2) combine
    if (have_cell) {
      RowEdge Rm, Rp;
      ColEdge Cm, Cp;
      Vertex Vmm, Vpm, Vmp, Vpp;
      CellLogs L;
#define NUSI_LD_R(X, i) X.t = S.R[k][0][i]; X.l1m = S.R[k][1][i]; X.labs = S.R[k][2][i]; X.lfabs = S.R[k][3][i]; X.lam2 = S.R[k][4][i]; X.argn = S.R[k][5][i];
#define NUSI_LD_C(X, i) X.s = S.C[k][0][i]; X.ls = S.C[k][1][i]; X.lam = S.C[k][2][i]; X.args = S.C[k][3][i]; X.l2s = S.C[k][4][i]; X.ls2 = S.C[k][5][i];
#define NUSI_LD_V(X, c, r) X.l1 = S.V[0][c][r]; X.lf = S.V[1][c][r]; X.dF = S.V[2][c][r]; X.lq = S.V[3][c][r]; X.d2 = S.V[4][c][r]; \
      X.zr_re = S.V[5][c][r]; X.zr_im = S.V[6][c][r]; X.zc_re = S.V[7][c][r]; X.zc_im = S.V[8][c][r]; X.ca = S.V[9][c][r];
      NUSI_LD_R(Rm, rm)
      NUSI_LD_R(Rp, rp)
      NUSI_LD_C(Cm, cm)
      NUSI_LD_C(Cp, cp)
      NUSI_LD_V(Vmm, cm, rm)
      NUSI_LD_V(Vpm, cp, rm)
      NUSI_LD_V(Vmp, cm, rp)
      NUSI_LD_V(Vpp, cp, rp)
#undef NUSI_LD_R
#undef NUSI_LD_C
#undef NUSI_LD_V
      L.lmp = S.fs[k][1][cb - b0];
      L.lpm = S.fs[k][2][cb - b0];
      L.LR1 = S.LR[cm][ca - a0];
      L.LR2 = S.LR[cp][ca - a0];
      L.LYm = S.LY[cb - b0][rm];
      L.LYp = S.LY[cb - b0][rp];
      alpha_cell_mass_t<true>(P, pp, k, S.fs[k][0][cb - b0], L, Rm, Rp, Cm, Cp, Vmm, Vpm, Vmp, Vpp, tot, &warn);
    }
    __syncthreads();
  }
  if (have_cell) tri[(size_t)p * tstride + (size_t)cb * (cb - 1) / 2 + ca] = tot;
  if (warn) atomicOr(&status[p], warn);
}

cudaError_t alpha_check_setup_attributes() {
  cudaError_t e = cudaFuncSetAttribute(alpha_tile_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(alpha_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem));
  return e;
}
cudaError_t launch_alpha_tile(const BuildArgs &A, int mode, cudaStream_t s) {
  dim3 gt(A.n_tiles, A.cnt);
  if (mode == 0)
    alpha_tile_kernel<0><<<gt, 256, sizeof(TileSmem), s>>>(A.G, A.nE, A.params, A.edgeE, A.edges, A.tiles, A.tri, A.tstride, A.status);
  else
    alpha_tile_kernel<1><<<gt, 256, sizeof(TileSmem), s>>>(A.G, A.nE, A.params, A.edgeE, A.edges, A.tiles, A.tri, A.tstride, A.status);
  return cudaGetLastError();
}
cudaError_t launch_alpha_entry(const BuildArgs &A, int compact, cudaStream_t s) {
  const DeviceGrid &G = A.G;
  const long long nent = G.non_resonant ? (long long)G.NT * (G.NT - 1) / 2 : (long long)(G.NT - 1);
  dim3 ga((unsigned)((nent + 127) / 128), A.cnt);
  alpha_kernel<<<ga, 128, 0, s>>>(G, A.params, A.tri, A.tstride, A.status, compact);
  return cudaGetLastError();
}

}  // namespace nusi
