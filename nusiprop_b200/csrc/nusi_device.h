// nusi_device.h -- types shared by the kernel translation units and the host pipeline (nusi_capi.cu):
// the per-context DeviceGrid, the tiling constants of the cross-section builders, and the data layout in HBM.
//
// Data layout in HBM (per launch chunk of P points, all FP64):
//   params   [P]                PointParams (mphi, g, Ga, m_k, |U_fk|^2, switches)
//   rows     [P][2][NT]         Gamma and alphaTilde per table row            (nuSIprop.hpp:234-235)
//   tri      [P][NT(NT-1)/2]    alpha table, strictly-upper triangle packed BY COLUMN:
//                               entry (row a, column b), a < b, at b(b-1)/2 + a.
//                               The integrator reads one column per finalised bin with
//                               consecutive threads on consecutive rows (coalesced).
//            [P][NT]            resonant-only contexts: only the super-diagonal (a, a+1) at [a]
//   source   [Nz-1][N]          Lum(z_i, bin j), context constant                (nuSIprop.hpp:659-662)
//   flux     [P][3][N] x2       outputs: mass basis, flavour basis              (nuSIprop.hpp:329-336)
#ifndef NUSI_DEVICE_H
#define NUSI_DEVICE_H
#include "nusi_aux.h"
#include "nusi_interp.h"

namespace nusi {

struct DeviceGrid {
  int N, Nz, NT;
  int majorana, non_resonant, phiphi, flav;
  const double *row_lo, *row_hi;  // [NT]
  const double *dE;               // [N]
  const double *cz, *ndz;         // [Nz-1]
  const double *source;           // [Nz-1][N]
  double U2[9];
  PhiPhiTables pp;
};

#define NUSI_TS 15             // tile side in cells: (TS+1)^2 = 256 vertices = one pass of 256 threads
#define NUSI_MAXE (2 * NUSI_TS) // most edges a tile side can have (all cells blue-shifted)
#define NUSI_VF 10             // doubles per Vertex
#define NUSI_RF 10             // doubles per RowEdge
#define NUSI_CF 9              // doubles per ColEdge
#define NUSI_FF 3              // doubles per column cell: fs, ln(s-/s+), ln(s+/s-)

NUSI_HD int edge_lo_id(int c, int N) { return c < N ? c : N + 1 + 2 * (c - N); }
NUSI_HD int edge_hi_id(int c, int N) { return c < N ? c + 1 : N + 2 + 2 * (c - N); }
NUSI_HD int edge_cell_min(int e, int N) { return e <= N ? (e == 0 ? 0 : e - 1) : N + (e - N - 1) / 2; }
NUSI_HD int edge_cell_max(int e, int N) { return e < N ? e : (e == N ? N - 1 : N + (e - N - 1) / 2); }

// Per-point edge data for the tiled builder: RowEdge / ColEdge for every distinct edge and mass,
// and the s-channel column factor of every cell.  Layout (doubles) per point:
//   [k][NUSI_RF][nE]  row-edge fields | [k][NUSI_CF][nE]  column-edge fields | [k][NUSI_FF][NT]  per column cell
NUSI_HD size_t edge_block_doubles(int nE, int NT) {
  return (size_t)3 * (NUSI_RF + NUSI_CF) * nE + (size_t)3 * NUSI_FF * NT;
}
#define NUSI_CE 16  // edges per tile side of the compact / scan builders
#define NUSI_ROWS_PER_CTA 64
#define NUSI_ROWSH 17  // doubles per (group, row, target mass) of the row kernels' g-independent part (>= sizeof(GammaShared), sizeof(ATildeShared))

}  // namespace nusi
#endif
