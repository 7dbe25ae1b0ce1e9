// k_alpha.cu -- the regeneration table alpha (reference nuSIprop.hpp:1237-1520): product builders
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
#include "nusi_launch.h"
#include "nusi_alpha.h"
#include "nusi_cpasync.h"

namespace nusi {

__global__ void __launch_bounds__(128)
edges_kernel(DeviceGrid G, int nE, const PointParams *__restrict__ params, const double *__restrict__ edgeE,
             double *__restrict__ edges) {
  const int p = blockIdx.y;
  const PointParams P = params[p];
  if (P.mphi <= 0) return;
  const int N = G.N, NT = G.NT;
  const int per_k = 2 * nE + NT;
  const int task = blockIdx.x * blockDim.x + threadIdx.x;
  if (task >= 3 * per_k) return;
  const int k = task / per_k, q = task - k * per_k;
  const double mn = P.mn[k], gr = P.Ga / P.mphi;
  double *base = edges + (size_t)p * edge_block_doubles(nE, NT);
  double *gR = base + (size_t)k * NUSI_RF * nE;
  double *gC = base + (size_t)3 * NUSI_RF * nE + (size_t)k * NUSI_CF * nE;
  double *gF = base + (size_t)3 * (NUSI_RF + NUSI_CF) * nE + (size_t)k * NUSI_FF * NT;
  if (q < nE) {
    RowEdge Re;
    make_row_edge(t_of(mn, edgeE[q], P.mphi), gr, true, Re);
    gR[0 * nE + q] = Re.t;
    gR[1 * nE + q] = Re.l1m;
    gR[2 * nE + q] = Re.labs;
    gR[3 * nE + q] = Re.lfabs;
    gR[4 * nE + q] = Re.lam2;
    gR[5 * nE + q] = Re.argn;
    gR[6 * nE + q] = Re.lnD;
    gR[7 * nE + q] = Re.phi;
    gR[8 * nE + q] = Re.i1p;
    gR[9 * nE + q] = Re.i1m;
  } else if (q < 2 * nE) {
    const int e = q - nE;
    ColEdge Ce;
    make_col_edge(s_of(mn, edgeE[e], P.mphi), gr, true, Ce);
    gC[0 * nE + e] = Ce.s;
    gC[1 * nE + e] = Ce.ls;
    gC[2 * nE + e] = Ce.lam;
    gC[3 * nE + e] = Ce.args;
    gC[4 * nE + e] = Ce.l2s;
    gC[5 * nE + e] = Ce.ls2;
    gC[6 * nE + e] = Ce.l2;
    gC[7 * nE + e] = Ce.lnN;
    gC[8 * nE + e] = Ce.is;
  } else {
    const int b = q - 2 * nE;
    const double sm = s_of(mn, edgeE[edge_lo_id(b, N)], P.mphi), sp = s_of(mn, edgeE[edge_hi_id(b, N)], P.mphi);
    gF[0 * NT + b] = alpha_s_colfactor(P, sm, sp);
    gF[1 * NT + b] = nlog(sm / sp);
    gF[2 * NT + b] = nlog(sp / sm);
  }
}

// Edge data for the scan builder: alpha_group_kernel reads the FULL edge block only of the first point of a group (the
// g-independent fields are the group's) and, of every other point, just the seven fields that contain g -- row edge lam2,
// lnD, phi; column edge lam, args, lnN; the s-channel column factor.  One CTA = (128 edge tasks, an eighth of a group): the first point
// gets everything, the others those seven (same expressions as make_row_edge / make_col_edge: same bits), which saves
// two thirds of the logarithms and of the 117 KB per point that edges_kernel writes.
__global__ void __launch_bounds__(128)
edges_group_kernel(DeviceGrid G, int nE, const PointParams *__restrict__ params, const double *__restrict__ edgeE,
                   const int2 *__restrict__ groups, double *__restrict__ edges) {
  const int2 grp = __ldg(groups + blockIdx.y);
  const int p0 = grp.x, glen = grp.y;
  if (params[p0].mphi <= 0) return;
  const int N = G.N, NT = G.NT;
  const int per_k = 2 * nE + NT;
  const int task = blockIdx.x * blockDim.x + threadIdx.x;
  if (task >= 3 * per_k) return;
  const int k = task / per_k, q = task - k * per_k;
  const size_t eblock = edge_block_doubles(nE, NT);
  for (int j = blockIdx.z; j < glen; j += gridDim.z) {  // gridDim.z CTAs share a group
    const PointParams P = params[p0 + j];
    const double mn = P.mn[k], gr = P.Ga / P.mphi;
    double *base = edges + (size_t)(p0 + j) * eblock;
    double *gR = base + (size_t)k * NUSI_RF * nE;
    double *gC = base + (size_t)3 * NUSI_RF * nE + (size_t)k * NUSI_CF * nE;
    double *gF = base + (size_t)3 * (NUSI_RF + NUSI_CF) * nE + (size_t)k * NUSI_FF * NT;
    if (q < nE) {
      const double t = t_of(mn, edgeE[q], P.mphi);
      if (j == 0) {
        RowEdge Re;
        make_row_edge(t, gr, true, Re);
        gR[0 * nE + q] = Re.t; gR[1 * nE + q] = Re.l1m; gR[2 * nE + q] = Re.labs; gR[3 * nE + q] = Re.lfabs; gR[4 * nE + q] = Re.lam2;
        gR[5 * nE + q] = Re.argn; gR[6 * nE + q] = Re.lnD; gR[7 * nE + q] = Re.phi; gR[8 * nE + q] = Re.i1p; gR[9 * nE + q] = Re.i1m;
      } else {
        const double lam2 = nlog1p(sqr(2 + t) / sqr(gr));
        gR[4 * nE + q] = lam2;
        gR[6 * nE + q] = lam2 + 2 * nlog(gr);
        gR[7 * nE + q] = natan2(gr, 2 + t);
      }
    } else if (q < 2 * nE) {
      const int e = q - nE;
      const double sv = s_of(mn, edgeE[e], P.mphi);
      if (j == 0) {
        ColEdge Ce;
        make_col_edge(sv, gr, true, Ce);
        gC[0 * nE + e] = Ce.s; gC[1 * nE + e] = Ce.ls; gC[2 * nE + e] = Ce.lam; gC[3 * nE + e] = Ce.args; gC[4 * nE + e] = Ce.l2s;
        gC[5 * nE + e] = Ce.ls2; gC[6 * nE + e] = Ce.l2; gC[7 * nE + e] = Ce.lnN; gC[8 * nE + e] = Ce.is;
      } else {
        const double lam = nlog1p(sqr(-1 + sv) / sqr(gr));
        gC[2 * nE + e] = lam;
        gC[3 * nE + e] = natan2(gr, -1 + sv);
        gC[7 * nE + e] = lam + 2 * nlog(gr);
      }
    } else {
      const int b = q - 2 * nE;
      const double sm = s_of(mn, edgeE[edge_lo_id(b, N)], P.mphi), sp = s_of(mn, edgeE[edge_hi_id(b, N)], P.mphi);
      gF[0 * NT + b] = alpha_s_colfactor(P, sm, sp);
      if (j == 0) {
        gF[1 * NT + b] = nlog(sm / sp);
        gF[2 * NT + b] = nlog(sp / sm);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// cross-section builder, compact tiled form (MODE 2, default for Majorana + non-resonant).
//
// Same decomposition as alpha_tile_kernel, with three changes that remove most of its non-FP64 work:
//  * the cell axis is cut into segments of <= 15 cells AND <= 16 edges (15 main-grid cells, or 8 blue-shifted
//    ones, which own two edges each), so a tile has at most 16 x 16 vertices = one vertex per thread, and every
//    index is a shift/mask of threadIdx;
//  * a vertex is stored as 3 doubles (l1, G, H: see make_vertex_compact) instead of 10, so the data of all
//    three target masses fit in 38 KB of shared memory: ONE barrier between "vertices + edge logarithms" and
//    "combine" instead of two per mass, and a cell loads 12 vertex doubles per mass instead of 40;
//  * no division in a vertex.
// ------------------------------------------------------------------------------------------
#ifndef NUSI_COMPACT_MINB
#define NUSI_COMPACT_MINB 4
#endif
struct CompactSmem {
  double V[3][3][NUSI_CE][NUSI_CE];  // [mass][l1, G, H][col edge][row edge]
  double LR[3][NUSI_CE][NUSI_CE];    // [mass][col edge][row cell]
  double LY[3][NUSI_CE][NUSI_CE];    // [mass][col cell][row edge]
  double R[3][NUSI_RF][NUSI_CE];
  double C[3][NUSI_CF][NUSI_CE];
  double fs[3][NUSI_FF][NUSI_CE];
  double wgt[3], U2f[3];  // m4/(2 m_k), |U_fk|^2
  CellConst K;
};

__global__ void __launch_bounds__(256, NUSI_COMPACT_MINB)
alpha_compact_kernel(DeviceGrid G, int nE, const PointParams *__restrict__ params, const double *__restrict__ edges,
                     const int4 *__restrict__ tiles, double *__restrict__ tri, size_t tstride, int *__restrict__ status) {
  __shared__ CompactSmem S;
  const int p = blockIdx.y, tid = threadIdx.x;
  const int4 tl = __ldg(tiles + blockIdx.x);  // cells [a0, a1) x [b0, b1): one load, nothing depends on another load
  if (params[p].mphi <= 0) return;
  const int N = G.N, NT = G.NT;
  const int a0 = tl.x, a1 = tl.y, b0 = tl.z, b1 = tl.w;
  const int ncr = a1 - a0, ncc = b1 - b0;
  const int r0 = edge_lo_id(a0, N), c0 = edge_lo_id(b0, N);
  const int nre = edge_hi_id(a1 - 1, N) - r0 + 1, nce = edge_hi_id(b1 - 1, N) - c0 + 1;
  const bool diag = (a0 == b0);
  const PhiPhiTables *pp = G.phiphi ? &G.pp : nullptr;
  const double *ebase = edges + (size_t)p * edge_block_doubles(nE, NT);
  const int lo = tid & 15, hi = tid >> 4;

  // phase 0: edge data of this tile, all three masses; per-point constants
  if (tid == 255) {
    const PointParams P = params[p];
    make_cell_const(P, S.K);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      S.wgt[k] = S.K.m4 / (2 * P.mn[k]);
      S.U2f[k] = P.U2f[k];
    }
  }
  {
    // [mass][field][edge slot]: row-edge fields first, then column-edge fields
    for (int idx = tid; idx < 3 * (NUSI_RF + NUSI_CF) * NUSI_CE; idx += 256) {
      const int e = idx & (NUSI_CE - 1), kf = idx / NUSI_CE, k = kf / (NUSI_RF + NUSI_CF), f = kf - k * (NUSI_RF + NUSI_CF);
      if (f < NUSI_RF) {
        if (e < nre) S.R[k][f][e] = ebase[(size_t)k * NUSI_RF * nE + (size_t)f * nE + r0 + e];
      } else {
        const int fc = f - NUSI_RF;
        if (e < nce) S.C[k][fc][e] = ebase[(size_t)3 * NUSI_RF * nE + (size_t)k * NUSI_CF * nE + (size_t)fc * nE + c0 + e];
      }
    }
    if (hi < 3 * NUSI_FF && lo < ncc)
      S.fs[hi / NUSI_FF][hi % NUSI_FF][lo] = ebase[(size_t)3 * (NUSI_RF + NUSI_CF) * nE + (size_t)hi * NT + b0 + lo];
  }
  __syncthreads();

  // this thread's cell (row cell lo, column cell hi) and its edges
  const int ca = a0 + lo, cb = b0 + hi;
  const int rm = edge_lo_id(min(ca, NT - 1), N) - r0, rp = edge_hi_id(min(ca, NT - 1), N) - r0;
  const int cm = edge_lo_id(min(cb, NT - 1), N) - c0, cp = edge_hi_id(min(cb, NT - 1), N) - c0;

  // phase 1: per mass one vertex (row edge lo, col edge hi), one LR (row cell lo, col edge hi), one LY (row edge lo, col cell hi)
  bool need_v = (lo < nre) && (hi < nce);
  bool need_lr = (lo < ncr) && (hi < nce);
  bool need_ly = (lo < nre) && (hi < ncc);
  if (diag) {
    need_v = need_v && (max(edge_cell_min(r0 + lo, N), a0) < min(edge_cell_max(c0 + hi, N), b1 - 1));
    need_lr = need_lr && (ca < min(edge_cell_max(c0 + hi, N), b1 - 1));
    need_ly = need_ly && (max(edge_cell_min(r0 + lo, N), a0) < cb);
  }
#pragma unroll 1
  for (int k = 0; k < 3; ++k) {
    if (need_v) {
      RowEdge Re;
      ColEdge Ce;
      Re.t = S.R[k][0][lo]; Re.l1m = S.R[k][1][lo]; Re.labs = S.R[k][2][lo]; Re.lfabs = S.R[k][3][lo];
      Re.lam2 = S.R[k][4][lo]; Re.argn = S.R[k][5][lo]; Re.lnD = S.R[k][6][lo]; Re.phi = S.R[k][7][lo];
      Ce.s = S.C[k][0][hi]; Ce.ls = S.C[k][1][hi]; Ce.lam = S.C[k][2][hi]; Ce.args = S.C[k][3][hi];
      Ce.l2 = S.C[k][6][hi]; Ce.lnN = S.C[k][7][hi];
      VertexC X;
      make_vertex_compact(Re, Ce, S.K.gr, X);
      S.V[k][0][hi][lo] = X.l1;
      S.V[k][1][hi][lo] = X.G;
      S.V[k][2][hi][lo] = X.H;
    }
    if (need_lr) S.LR[k][hi][lo] = alpha_log_LR_c(S.C[k][0][hi], S.R[k][0][rm], S.R[k][0][rp]);
    if (need_ly) S.LY[k][hi][lo] = alpha_log_LY_c(S.C[k][0][cm], S.C[k][0][cp], S.R[k][0][lo]);
  }
  __syncthreads();

  // phase 2: combine
  if (!((lo < ncr) && (hi < ncc) && (ca < cb))) return;
  double tot = 0;
  int warn = 0;
#pragma unroll 1
  for (int k = 0; k < 3; ++k) {
    RowEdge Rm, Rp;
    ColEdge Cm, Cp;
    CellLogs L;
    Rm.t = S.R[k][0][rm]; Rm.l1m = S.R[k][1][rm]; Rm.labs = S.R[k][2][rm]; Rm.i1p = S.R[k][8][rm]; Rm.i1m = S.R[k][9][rm];
    Rp.t = S.R[k][0][rp]; Rp.l1m = S.R[k][1][rp]; Rp.labs = S.R[k][2][rp]; Rp.i1p = S.R[k][8][rp]; Rp.i1m = S.R[k][9][rp];
    Cm.s = S.C[k][0][cm]; Cm.ls = S.C[k][1][cm]; Cm.lam = S.C[k][2][cm]; Cm.args = S.C[k][3][cm]; Cm.l2s = S.C[k][4][cm]; Cm.ls2 = S.C[k][5][cm]; Cm.is = S.C[k][8][cm];
    Cp.s = S.C[k][0][cp]; Cp.ls = S.C[k][1][cp]; Cp.lam = S.C[k][2][cp]; Cp.args = S.C[k][3][cp]; Cp.l2s = S.C[k][4][cp]; Cp.ls2 = S.C[k][5][cp]; Cp.is = S.C[k][8][cp];
    L.lmp = S.fs[k][1][hi];
    L.lpm = S.fs[k][2][hi];
    L.LR1 = S.LR[k][cm][lo];
    L.LR2 = S.LR[k][cp][lo];
    L.LYm = S.LY[k][hi][rm];
    L.LYp = S.LY[k][hi][rp];
    const double Gs = S.V[k][1][cm][rm] - S.V[k][1][cp][rm] - S.V[k][1][cm][rp] + S.V[k][1][cp][rp];
    const double Hs = S.V[k][2][cm][rm] - S.V[k][2][cp][rm] - S.V[k][2][cm][rp] + S.V[k][2][cp][rp];
    alpha_cell_mass_compact<true>(S.K, pp, S.wgt[k], S.U2f[k], S.fs[k][0][hi], L, Rm, Rp, Cm, Cp, S.V[k][0][cm][rm], S.V[k][0][cp][rm], S.V[k][0][cm][rp],
                            S.V[k][0][cp][rp], Gs, Hs, tot, &warn);
  }
  tri[(size_t)p * tstride + (size_t)cb * (cb - 1) / 2 + ca] = tot;
  if (warn) atomicOr(&status[p], warn);
}

// ------------------------------------------------------------------------------------------
// cross-section builder for parameter SCANS (MODE 3): alpha_compact_kernel with the g-independent work
// hoisted out of a loop over the points of a group.
//
// A group is a run of points with identical (mphi, m_nu) -- a row of a (mphi, g) scan.  Within a group the
// edge grid (s', t), every real dilogarithm, ln(1+s+t), the bin-ratio logarithms and the brackets of the t
// channel and of the t-u interference are the same numbers for every point; g enters only through
// ga_red = Gamma/mphi (complex dilogarithm, arg and Breit-Wigner logarithms of the s-t interference, the s
// channel) and through prefactors.  One CTA owns one tile of one GROUP:
//   phase 0   edge data of the group's first point                                   (once)
//   phase 1a  per vertex and mass: l1, H, Re Li2(zr); the LR / LY logarithms         (once)
//   phase 2a  per cell and mass: X (t channel bracket), Y (t-u bracket) -> shared    (once)
//   per point of the group:  1b  G per vertex and mass (complex Li2 core)  |barrier|
//                            2b  combine X, Y, G and the per-point prefactors, store the entry
// The per-point loop is software-pipelined (G of point j+1 and the combine of point j between the same two
// barriers; G double buffered; the per-point edge fields -- 3 row, 3 column, 1 per column cell and mass -- and
// CellConst triple buffered, fetched with cp.async two points ahead), so a point costs ONE barrier.
// The split is the one of make_vertex_compact_shared/_G and alpha_cell_shared/_combine: the same
// operations on the same operands in the same order as alpha_compact_kernel, hence the same bits
// (asserted in tests/test_gpu_parity.py for whole tables, tests/test_host_math.py for the functions).
// ------------------------------------------------------------------------------------------
#define NUSI_RGF 3  // per-point row-edge fields: lam2, lnD, phi
#define NUSI_CGF 3  // per-point column-edge fields: lam, args, lnN
struct GroupSmem {
  double V[3][3][NUSI_CE][NUSI_CE];  // [mass][l1, Re Li2(zr), H -> G][col edge][row edge]
  double LR[3][NUSI_CE][NUSI_CE];    // phase 1a: LR [mass][col edge][row cell]; afterwards X [mass][col cell][row cell]
  double LY[3][NUSI_CE][NUSI_CE];    // phase 1a: LY [mass][col cell][row edge]; afterwards Y [mass][col cell][row cell]
  double R[3][NUSI_RF][NUSI_CE];     // edge data of the group's first point (the g-independent fields are the group's)
  double C[3][NUSI_CF][NUSI_CE];
  double fs[3][NUSI_FF][NUSI_CE];
  double G1[3][NUSI_CE][NUSI_CE];      // second G buffer (the first one is the H slot of V)
  double GL[3][NUSI_CE][NUSI_CE];      // t-channel fallback integral [mass][col cell][row cell] (0 where X >= 0)
  struct PerPoint {                    // per-point fields, triple buffered
    double Rg[3][NUSI_RGF][NUSI_CE];
    double Cg[3][NUSI_CGF][NUSI_CE];
    double fsg[3][NUSI_CE];
    CellConst K;
  } PP[3];
  double wgt[3], U2f[3];
};
#ifndef NUSI_GROUP_UNROLL_K
#define NUSI_GROUP_UNROLL_K 1  // masses unrolled in the per-point loop: 14.98 -> 14.48 ms per 8192 points
#endif
#if NUSI_GROUP_UNROLL_K
#define NUSI_GK_UNROLL _Pragma("unroll")
#else
#define NUSI_GK_UNROLL _Pragma("unroll 1")
#endif
#ifndef NUSI_GROUP_MINB
#define NUSI_GROUP_MINB 3  // 80 registers: the per-point loop keeps its indices and addresses in registers (4 CTAs/SM at 64 registers spills them: 20.3 vs 16.3 ms per 8192 points)
#endif

__global__ void __launch_bounds__(256, NUSI_GROUP_MINB)
alpha_group_kernel(DeviceGrid G, int nE, const PointParams *__restrict__ params, const double *__restrict__ edges,
                   const int4 *__restrict__ tiles, const int2 *__restrict__ groups, double *__restrict__ tri, size_t tstride,
                   int *__restrict__ status) {
  extern __shared__ __align__(16) unsigned char nusi_group_smem[];
  GroupSmem &S = *reinterpret_cast<GroupSmem *>(nusi_group_smem);
  const int tid = threadIdx.x;
  const int2 grp = __ldg(groups + blockIdx.y);  // points [p0, p0 + glen) of this launch
  const int4 tl = __ldg(tiles + blockIdx.x);
  const int p0 = grp.x, glen = grp.y;
  if (params[p0].mphi <= 0) return;  // rejected points form groups of their own (same mphi <= 0 marker)
  const int N = G.N, NT = G.NT;
  const int a0 = tl.x, a1 = tl.y, b0 = tl.z, b1 = tl.w;
  const int ncr = a1 - a0, ncc = b1 - b0;
  const int r0 = edge_lo_id(a0, N), c0 = edge_lo_id(b0, N);
  const int nre = edge_hi_id(a1 - 1, N) - r0 + 1, nce = edge_hi_id(b1 - 1, N) - c0 + 1;
  const bool diag = (a0 == b0);
  const size_t eblock = edge_block_doubles(nE, NT);
  const int lo = tid & 15, hi = tid >> 4;

  // per-point fields of point p into buffer b (cp.async; the caller commits / waits).  Each thread copies up to
  // two doubles per point; their source offset inside a point's edge block and destination offset inside a
  // PerPoint buffer are fixed for the whole group, so they are worked out once (negative source = nothing to read:
  // the slot is zero-filled).
  int fsrc[2];
  unsigned fdst[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int idx = tid + q * 256;
    fsrc[q] = -1;
    fdst[q] = 0xffffffffu;
    if (idx < 3 * (NUSI_RGF + NUSI_CGF + 1) * NUSI_CE) {
      const int e = idx & (NUSI_CE - 1), kf = idx / NUSI_CE, k = kf / (NUSI_RGF + NUSI_CGF + 1), f = kf - k * (NUSI_RGF + NUSI_CGF + 1);
      if (f < NUSI_RGF) {
        const int fr = (f == 0) ? 4 : (f == 1 ? 6 : 7);
        fdst[q] = (unsigned)(offsetof(GroupSmem::PerPoint, Rg) + ((k * NUSI_RGF + f) * NUSI_CE + e) * sizeof(double));
        if (e < nre) fsrc[q] = (k * NUSI_RF + fr) * nE + r0 + e;
      } else if (f < NUSI_RGF + NUSI_CGF) {
        const int fc = f - NUSI_RGF, fr = (fc == 0) ? 2 : (fc == 1 ? 3 : 7);
        fdst[q] = (unsigned)(offsetof(GroupSmem::PerPoint, Cg) + ((k * NUSI_CGF + fc) * NUSI_CE + e) * sizeof(double));
        if (e < nce) fsrc[q] = (3 * NUSI_RF + k * NUSI_CF + fr) * nE + c0 + e;
      } else {
        fdst[q] = (unsigned)(offsetof(GroupSmem::PerPoint, fsg) + (k * NUSI_CE + e) * sizeof(double));
        if (e < ncc) fsrc[q] = 3 * (NUSI_RF + NUSI_CF) * nE + (k * NUSI_FF) * NT + b0 + e;
      }
    }
  }
  const unsigned pp_base = (unsigned)__cvta_generic_to_shared(&S.PP[0]);
  auto fetch_point = [&](int p, int b) {
    const double *eb = edges + (size_t)p * eblock;
    const unsigned base = pp_base + (unsigned)(b * sizeof(GroupSmem::PerPoint));
    cp_async8s(base + fdst[0], eb + (fsrc[0] < 0 ? 0 : fsrc[0]), fsrc[0] >= 0);
    if (fdst[1] != 0xffffffffu) cp_async8s(base + fdst[1], eb + (fsrc[1] < 0 ? 0 : fsrc[1]), fsrc[1] >= 0);
    if (tid == 255) make_cell_const(params[p], S.PP[b].K);
  };

  // phase 0: edge data of this tile from the group's first point, all three masses
  const double *ebase = edges + (size_t)p0 * eblock;
  if (tid == 255) {
    const PointParams P = params[p0];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      S.wgt[k] = sqr(sqr(P.mphi)) / (2 * P.mn[k]);
      S.U2f[k] = P.U2f[k];
    }
  }
  fetch_point(p0, 0);
  cp_async_commit();
  for (int idx = tid; idx < 3 * (NUSI_RF + NUSI_CF) * NUSI_CE; idx += 256) {
    const int e = idx & (NUSI_CE - 1), kf = idx / NUSI_CE, k = kf / (NUSI_RF + NUSI_CF), f = kf - k * (NUSI_RF + NUSI_CF);
    if (f < NUSI_RF) {
      if (e < nre) S.R[k][f][e] = ebase[(size_t)k * NUSI_RF * nE + (size_t)f * nE + r0 + e];
    } else {
      const int fc = f - NUSI_RF;
      if (e < nce) S.C[k][fc][e] = ebase[(size_t)3 * NUSI_RF * nE + (size_t)k * NUSI_CF * nE + (size_t)fc * nE + c0 + e];
    }
  }
  if (hi < 3 * NUSI_FF && lo < ncc)
    S.fs[hi / NUSI_FF][hi % NUSI_FF][lo] = ebase[(size_t)3 * (NUSI_RF + NUSI_CF) * nE + (size_t)hi * NT + b0 + lo];
  __syncthreads();

  const int ca = a0 + lo, cb = b0 + hi;
  const int rm = edge_lo_id(min(ca, NT - 1), N) - r0, rp = edge_hi_id(min(ca, NT - 1), N) - r0;
  const int cm = edge_lo_id(min(cb, NT - 1), N) - c0, cp = edge_hi_id(min(cb, NT - 1), N) - c0;
  bool need_v = (lo < nre) && (hi < nce);
  bool need_lr = (lo < ncr) && (hi < nce);
  bool need_ly = (lo < nre) && (hi < ncc);
  if (diag) {
    need_v = need_v && (max(edge_cell_min(r0 + lo, N), a0) < min(edge_cell_max(c0 + hi, N), b1 - 1));
    need_lr = need_lr && (ca < min(edge_cell_max(c0 + hi, N), b1 - 1));
    need_ly = need_ly && (max(edge_cell_min(r0 + lo, N), a0) < cb);
  }
  const bool is_cell = (lo < ncr) && (hi < ncc) && (ca < cb);

  // phase 1a: the g-independent part of the vertices, LR, LY
#pragma unroll 1
  for (int k = 0; k < 3; ++k) {
    if (need_v) {
      RowEdge Re;
      ColEdge Ce;
      Re.t = S.R[k][0][lo]; Re.l1m = S.R[k][1][lo]; Re.labs = S.R[k][2][lo];
      Ce.s = S.C[k][0][hi]; Ce.ls = S.C[k][1][hi]; Ce.l2 = S.C[k][6][hi];
      VertexS X;
      make_vertex_compact_shared(Re, Ce, X);
      S.V[k][0][hi][lo] = X.l1;
      S.V[k][1][hi][lo] = X.zr_re;
      S.V[k][2][hi][lo] = X.H;
    }
    if (need_lr) S.LR[k][hi][lo] = alpha_log_LR_c(S.C[k][0][hi], S.R[k][0][rm], S.R[k][0][rp]);
    if (need_ly) S.LY[k][hi][lo] = alpha_log_LY_c(S.C[k][0][cm], S.C[k][0][cp], S.R[k][0][lo]);
  }
  __syncthreads();

  // phase 2a: the g-independent brackets of this thread's cell; they replace LR / LY of the same mass in
  // shared memory once every reader of those is done (slot [col cell][row cell], read back by its writer only)
#pragma unroll 1
  for (int k = 0; k < 3; ++k) {
    CellShared O;
    if (is_cell) {
      RowEdge Rm, Rp;
      ColEdge Cm, Cp;
      CellLogs L;
      Rm.t = S.R[k][0][rm]; Rm.l1m = S.R[k][1][rm]; Rm.labs = S.R[k][2][rm]; Rm.i1p = S.R[k][8][rm]; Rm.i1m = S.R[k][9][rm];
      Rp.t = S.R[k][0][rp]; Rp.l1m = S.R[k][1][rp]; Rp.labs = S.R[k][2][rp]; Rp.i1p = S.R[k][8][rp]; Rp.i1m = S.R[k][9][rp];
      Cm.s = S.C[k][0][cm]; Cm.ls = S.C[k][1][cm]; Cm.l2s = S.C[k][4][cm]; Cm.ls2 = S.C[k][5][cm];
      Cp.s = S.C[k][0][cp]; Cp.ls = S.C[k][1][cp]; Cp.l2s = S.C[k][4][cp]; Cp.ls2 = S.C[k][5][cp];
      L.lmp = S.fs[k][1][hi];
      L.lpm = S.fs[k][2][hi];
      L.LR1 = S.LR[k][cm][lo];
      L.LR2 = S.LR[k][cp][lo];
      L.LYm = S.LY[k][hi][rm];
      L.LYp = S.LY[k][hi][rp];
      const double Hs = S.V[k][2][cm][rm] - S.V[k][2][cp][rm] - S.V[k][2][cm][rp] + S.V[k][2][cp][rp];
      alpha_cell_shared(L, Rm, Rp, Cm, Cp, S.V[k][0][cm][rm], S.V[k][0][cp][rm], S.V[k][0][cm][rp], S.V[k][0][cp][rp], Hs, O);
    }
    __syncthreads();
    if (is_cell) {
      S.LR[k][hi][lo] = O.X;
      S.LY[k][hi][lo] = O.Y;
      S.GL[k][hi][lo] = O.GL;
    }
  }

  // the points of the group, software-pipelined: between two barriers a thread computes G of its vertex for point
  // j+1 AND combines its cell for point j, so vertex-heavy and cell-heavy threads even out and a point costs one
  // barrier.  G is double buffered (G2[0] is the H slot, dead by now), the per-point fields triple buffered.
  auto vertex_G = [&](int b3, int g2) {
    if (need_v) {
      const double gr = S.PP[b3].K.gr;
NUSI_GK_UNROLL
      for (int k = 0; k < 3; ++k) {
        RowEdge Re;
        ColEdge Ce;
        Re.t = S.R[k][0][lo]; Re.labs = S.R[k][2][lo]; Re.lfabs = S.R[k][3][lo]; Re.argn = S.R[k][5][lo];
        Re.lam2 = S.PP[b3].Rg[k][0][lo]; Re.lnD = S.PP[b3].Rg[k][1][lo]; Re.phi = S.PP[b3].Rg[k][2][lo];
        Ce.s = S.C[k][0][hi]; Ce.ls = S.C[k][1][hi];
        Ce.lam = S.PP[b3].Cg[k][0][hi]; Ce.args = S.PP[b3].Cg[k][1][hi]; Ce.lnN = S.PP[b3].Cg[k][2][hi];
        const double Gv = make_vertex_compact_G(Re, Ce, gr, S.V[k][0][hi][lo], S.V[k][1][hi][lo]);
        if (g2) S.G1[k][hi][lo] = Gv; else S.V[k][2][hi][lo] = Gv;
      }
    }
  };
  if (glen > 1) fetch_point(p0 + 1, 1);
  cp_async_commit();
  cp_async_wait<1>();  // the fields of the first point (fetched in phase 0) have landed
  __syncthreads();
  vertex_G(0, 0);
  int b3 = 0;  // j % 3
  double *tdst = tri + (size_t)p0 * tstride + (size_t)(cb * (cb - 1) / 2 + ca);  // this thread's entry of point j
#pragma unroll 1
  for (int j = 0; j < glen; ++j) {
    const int p = p0 + j, g2 = j & 1;
    const int b3n = (b3 == 2) ? 0 : b3 + 1, b3nn = (b3n == 2) ? 0 : b3n + 1;
    cp_async_wait<0>();  // this thread's copies for point j+1 have landed (issued one point ago)
    __syncthreads();     // G of point j and the fields of point j+1 are visible; everybody is done with point j-1
    if (j + 2 < glen) fetch_point(p + 2, b3nn);  // buffer (j+2)%3 == (j-1)%3
    cp_async_commit();
    if (j + 1 < glen) vertex_G(b3n, g2 ^ 1);
    if (is_cell) {
      double tot = 0;
      int warn = 0;
NUSI_GK_UNROLL
      for (int k = 0; k < 3; ++k) {
        RowEdge Rm, Rp;
        ColEdge Cm, Cp;
        Rm.t = S.R[k][0][rm]; Rm.l1m = S.R[k][1][rm]; Rm.i1p = S.R[k][8][rm];
        Rp.t = S.R[k][0][rp]; Rp.l1m = S.R[k][1][rp]; Rp.i1p = S.R[k][8][rp];
        Cm.s = S.C[k][0][cm]; Cm.ls = S.C[k][1][cm]; Cm.is = S.C[k][8][cm]; Cm.lam = S.PP[b3].Cg[k][0][cm]; Cm.args = S.PP[b3].Cg[k][1][cm];
        Cp.s = S.C[k][0][cp]; Cp.ls = S.C[k][1][cp]; Cp.is = S.C[k][8][cp]; Cp.lam = S.PP[b3].Cg[k][0][cp]; Cp.args = S.PP[b3].Cg[k][1][cp];
        const double(*Gk)[NUSI_CE] = g2 ? S.G1[k] : S.V[k][2];
        const double Gs = Gk[cm][rm] - Gk[cp][rm] - Gk[cm][rp] + Gk[cp][rp];
        alpha_cell_combine(S.PP[b3].K, S.wgt[k], S.U2f[k], S.PP[b3].fsg[k][hi], Rm, Rp, Cm, Cp, S.LR[k][hi][lo], S.LY[k][hi][lo], S.GL[k][hi][lo], Gs, tot, &warn);
      }
      *tdst = tot;
      if (warn) atomicOr(&status[p], warn);
    }
    tdst += tstride;
    b3 = b3n;
  }
}

#ifndef NUSI_GROUP_SMEM_PAD
#define NUSI_GROUP_SMEM_PAD 0  // tuning experiments: extra dynamic shared memory per CTA, to cap the resident CTAs per SM
#endif
cudaError_t alpha_setup_attributes() {
  return cudaFuncSetAttribute(alpha_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GroupSmem) + NUSI_GROUP_SMEM_PAD);
}
size_t alpha_group_smem_bytes() { return sizeof(GroupSmem); }
cudaError_t launch_edges(const BuildArgs &A, cudaStream_t s) {
  const int ntask = 3 * (2 * A.nE + A.G.NT);
  dim3 ge((ntask + 127) / 128, A.cnt);
  edges_kernel<<<ge, 128, 0, s>>>(A.G, A.nE, A.params, A.edgeE, A.edges);
  return cudaGetLastError();
}
cudaError_t launch_edges_group(const BuildArgs &A, cudaStream_t s) {  // only for a launch that alpha_group_kernel consumes
  const int ntask = 3 * (2 * A.nE + A.G.NT);
  dim3 ge((ntask + 127) / 128, A.ngroups, 8);
  edges_group_kernel<<<ge, 128, 0, s>>>(A.G, A.nE, A.params, A.edgeE, A.groups, A.edges);
  return cudaGetLastError();
}
cudaError_t launch_alpha_group(const BuildArgs &A, cudaStream_t s) {
  dim3 gg(A.n_ctiles, A.ngroups);
  alpha_group_kernel<<<gg, 256, sizeof(GroupSmem) + NUSI_GROUP_SMEM_PAD, s>>>(A.G, A.nE, A.params, A.edges, A.ctiles, A.groups, A.tri, A.tstride, A.status);
  return cudaGetLastError();
}
cudaError_t launch_alpha_compact(const BuildArgs &A, cudaStream_t s) {
  dim3 gc(A.n_ctiles, A.cnt);
  alpha_compact_kernel<<<gc, 256, 0, s>>>(A.G, A.nE, A.params, A.edges, A.ctiles, A.tri, A.tstride, A.status);
  return cudaGetLastError();
}

}  // namespace nusi
