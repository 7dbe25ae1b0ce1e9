// k_integrate_cluster.cu -- the implicit z-stepping integrator (reference nuSIprop.hpp:255-336) for FINE binnings
// (576 < N_bins_E <= 2017; BASELINE configs[3]: N_E = 2000): the role-split wavefront of k_integrate_split.cu with ONE POINT
// SPREAD OVER A THREAD-BLOCK CLUSTER of 4 CTAs (4 SMs; 2 CTAs up to N_bins_E = 993).
// (sm_100a, FP64 pipe; -fmad=false: contraction only where fma() is written)
//
// Why a cluster: the wavefront needs one running sum per (table diagonal, z-step of the band).  At N_E = 2000 that is
// 2031 diagonals x 32 z-steps x 8 B = 520 KB -- more than one SM's register file, which is why the one-CTA-per-point
// kernel (integrate_diag_kernel) has to cut the z-axis into 39 bands of 8 z-steps and walk the 2007 waves of a band 39
// times (78 k dependent waves, 1.9 us each: 0.15 s per point).  Four SMs hold the sums of a 32-step band in registers
// (508 diagonals per CTA), so a point takes 10 bands = 20 k waves, and the wave-to-wave chain is carried by one warp as
// in integrate_split_kernel instead of by two block-wide barriers per wave.
//
//   rank 0, warp 0         critical warp: lane = z-step.  Finalises one cell per lane and wave (flux of the previous z-step
//                          by warp shuffle, flux-independent half from a helper slot, regeneration sum = far part published
//                          by the diagonal's owners + the last D = 2B-1 sources kept in registers), and BROADCASTS the new
//                          weighted flux W of the wave into the shared memory of every CTA of the cluster (st.shared::cluster).
//   ranks 0-1, warps 1..2  helpers: slots (denominators, reciprocals, source term) one block of B waves ahead, one wave
//                          each; rank 1 writes its slots into rank 0's shared memory.
//   all ranks, 1 warp      producer (one lane per copy): the TMA copies of this CTA's table slices, three blocks ahead.
//   all ranks, 16 warps    bulk: a thread owns two neighbouring diagonals of its rank's 508 and half of the band's z-steps:
//                          acc[s] += T[row(d)][c] W_s for every source more than D waves behind its cells, one block behind
//                          the critical warp; the owners of the diagonal that is finalised in block kb+1 write its 32 sums
//                          into rank 0's shared memory (DSMEM).  Rank 0 owns the highest (short-lived) diagonals.
//   every CTA streams ITS slice of each table column (the rows of its diagonals) through a shared-memory ring with TMA
//   bulk copies (cp.async.bulk + mbarrier), three blocks ahead; rank 0 also fetches the D near-diagonal rows the critical
//   warp needs.  One cluster barrier (barrier.cluster.arrive.release / wait.acquire) per block of B = 4 waves.
//
// Summation order: as integrate_split_kernel -- a cell's regeneration sum is (far part) + (near part), each accumulated
// in the reference's order (source bins high -> low); it differs from the strictly sequential sum by rounding only.
#include <cfloat>
#include "nusi_launch.h"
#include "nusi_implicit.h"
#include "nusi_split.h"

namespace nusi {

#define NUSI_CL_C 4      // most CTAs per cluster (the kernel is instantiated for clusters of 2 and 4)
#ifndef NUSI_CL_HR
#define NUSI_CL_HR 2     // ranks whose helper warps prepare slots (ranks >= 1 write them into rank 0's shared memory)
#endif
#define NUSI_CL_PF 12    // slot fields: lum, offdiag, Gam, aT, Z0, rZ0, Z1, rZ1, Z2, rZ2, dE, 1/dE
#define NUSI_CL_RB 5     // ring depth in blocks of B columns (block kb-1 .. kb+3)
#define NUSI_CL_PAD 16   // doubles in front of the near ring: the critical warp reads up to D rows below row 0 of a column
#define NUSI_CL_NEAR 10  // doubles per near-ring slot: D <= 7 rows + 1 for alignment, rounded to an even count, + 2

__device__ __forceinline__ unsigned cl_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cl_mapa(unsigned saddr, unsigned rank) {  // the same shared-memory offset in CTA `rank` of the cluster
  unsigned r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cl_st_f64(unsigned addr, double v) {
  asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void cl_st_f64x2(unsigned addr, double a, double b) {
  asm volatile("st.shared::cluster.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(a), "d"(b) : "memory");
}
// every thread of every CTA of the cluster; orders the DSMEM stores before it against the loads after it
__device__ __forceinline__ void cl_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

struct ClCtx {  // at the start of dynamic shared memory of every CTA, filled by its thread 0
  int N, NT, S, NP, SLOT, DPC, rank, dbg, csize;  // csize: CTAs of the cluster
  int s0, cnt, n_waves, n_blocks, rowoff, a0;  // a0: table row of this CTA's first diagonal in this band (may be negative)
  int ring, near, wsh, ash, pre, sdE, sRdE, sCz, sNd, fband, tcoff, fb;  // offsets in doubles
  unsigned mbar0, ring0, near0, phase0, tpar, wsh_s, ash_s;              // shared::cta byte addresses
  double U2f[3];
  const double *T, *source, *rowG, *rowA;
};
#define NUSI_CL_CTX_DOUBLES ((int)((sizeof(ClCtx) + 15) / 16) * 2)

struct ClLayout {
  int SLOT, ring, near, wsh, ash, pre, sdE, sRdE, sCz, sNd, fband, mbar, tcoff, fb, total;
};
__host__ __device__ inline ClLayout cl_layout(int N, int S, int B, int DPC) {
  ClLayout L;
  L.SLOT = (DPC + 3) & ~1;  // a column slice, possibly shifted by one double for 16-byte alignment of its global source
  int o = NUSI_CL_CTX_DOUBLES;
  L.ring = o; o += NUSI_CL_RB * B * L.SLOT;
  o += NUSI_CL_PAD;
  L.near = o; o += NUSI_CL_RB * B * NUSI_CL_NEAR;
  L.wsh = o; o += 2 * B * 32;
  L.ash = o; o += 2 * B * 32;
  L.pre = o; o += 2 * B * NUSI_CL_PF * 32;
  L.sdE = o; o += (N + 1) & ~1;
  L.sRdE = o; o += (N + 1) & ~1;
  L.sCz = o; o += (S + 1) & ~1;
  L.sNd = o; o += (S + 1) & ~1;
  L.fband = o; o += 3 * ((N + 1) & ~1);
  L.mbar = o; o += NUSI_CL_RB + 1;
  L.tcoff = o; o += B;    // 2 x B ints: per wave of a block, where row c of its column sits in the near ring
  L.fb = o; o += 4 * 32;  // results of the out-of-line general step: f0, f1, f2, W per lane
  L.total = o;
  return L;
}

// packed offset (doubles) of row a of column c, and whether the 16-byte aligned copy that contains it starts one double early
__device__ __forceinline__ size_t cl_packed(int c, int a) { return (size_t)c * (size_t)(c - 1) / 2 + (size_t)a; }
__device__ __forceinline__ int cl_mis(unsigned tpar, int c, int a) { return (int)((tpar ^ (unsigned)(c >> 1) ^ (unsigned)a) & 1u); }

__device__ __forceinline__ unsigned cl_parity(unsigned phase0, int kb) {  // parity on which block kb of this band completes
  const int rb = kb % NUSI_CL_RB;
  return ((phase0 >> rb) & 1u) ^ (unsigned)((kb / NUSI_CL_RB) & 1);
}

// One wave of the critical warp (split_wave of k_integrate_split.cu; W goes to all CTAs of the cluster).
template <int D, bool ALL, int CC>
__device__ __forceinline__ void cl_wave(bool act, int sl, int cnt, int bb, int NP, double U0, double U1, double U2, double czl, double ndl,
                                        const double *__restrict__ slot, const double *__restrict__ ash, const unsigned (&wdst)[CC],
                                        unsigned woff, const double *__restrict__ tc, double *__restrict__ fband, double *__restrict__ fb,
                                        double (&ip)[D], double &o0, double &o1, double &o2) {
  const double u0 = __shfl_up_sync(0xffffffffu, o0, 1), u1 = __shfl_up_sync(0xffffffffu, o1, 1), u2 = __shfl_up_sync(0xffffffffu, o2, 1);
  const double f0 = (sl == 0) ? fband[bb] : u0, f1 = (sl == 0) ? fband[NP + bb] : u1, f2 = (sl == 0) ? fband[2 * NP + bb] : u2;
  const double lum = slot[0], offdiag = slot[32];
  const double Z0 = slot[4 * 32], r0 = slot[5 * 32], Z1 = slot[6 * 32], r1 = slot[7 * 32], Z2 = slot[8 * 32], r2 = slot[9 * 32];
  const double dEb = slot[10 * 32], rdE = slot[11 * 32];
  const double acc_c = ash[0] + ip[0];
  const double base = czl * (ndl * acc_c);
  const double s0 = lum + base * U0, s1 = lum + base * U1, s2 = lum + base * U2;
  double v0 = quot_fast(f0 + s0, Z0, r0), v1 = quot_fast(f1 + s1, Z1, r1), v2 = quot_fast(f2 + s2, Z2, r2);
  const double dot = U0 * v0 + U1 * v1 + U2 * v2;
  double W = quot_fast(dot, dEb, rdE);
  const int h0 = __double2hiint(v0) & 0x7fffffff, h1 = __double2hiint(v1) & 0x7fffffff, h2 = __double2hiint(v2) & 0x7fffffff;
  const int hmax = max(h0, max(h1, h2)), hmin = min(h0, min(h1, h2));
  const bool identity = (hmax < 0x7ff00000) && ((__double2hiint(W) & 0x7fffffff) < 0x7ff00000) &&
                        (offdiag * __hiloint2double(hmax + 1, 0) <= 6.9388939039072284e-18 /* 2^-57 */ * __hiloint2double(hmin, 0));
  if (!identity && (ALL || act)) {
    split_fallback(U0, U1, U2, czl, slot, s0, s1, s2, f0, f1, f2, fb);
    v0 = fb[0];
    v1 = fb[32];
    v2 = fb[64];
    W = fb[96];
  }
  const double Wp = (ALL || act) ? W : 0.0;
#pragma unroll
  for (int r = 0; r < CC; ++r) cl_st_f64(wdst[r] + woff, Wp);
  if (ALL) {
    o0 = v0;
    o1 = v1;
    o2 = v2;
#pragma unroll
    for (int j = 1; j < D; ++j) ip[j - 1] = fma(tc[-j], W, ip[j]);
    ip[D - 1] = tc[-D] * W;
  } else {
    if (act) {
      o0 = v0;
      o1 = v1;
      o2 = v2;
    }
#pragma unroll
    for (int j = 1; j < D; ++j) ip[j - 1] = act ? fma(tc[-j], W, ip[j]) : ip[j];
    ip[D - 1] = act ? tc[-D] * W : 0.0;
  }
  if (sl == cnt - 1 && (ALL || act)) {  // the band's last z-step: this bin's flux after the band
    fband[bb] = v0;
    fband[NP + bb] = v1;
    fband[2 * NP + bb] = v2;
  }
}

// ================= critical warp (rank 0) =================
template <int B, int CC>
static __device__ __noinline__ void cl_critical(double *ssm) {
  constexpr int D = 2 * B - 1, PF = NUSI_CL_PF;
  const ClCtx &C = *reinterpret_cast<const ClCtx *>(ssm);
  const int sl = threadIdx.x & 31, N = C.N, NP = C.NP, cnt = C.cnt, n_blocks = C.n_blocks;
  const double U0 = C.U2f[0], U1 = C.U2f[1], U2 = C.U2f[2];
  double *fband = ssm + C.fband, *fb = ssm + C.fb + sl;
  double ip[D];
#pragma unroll
  for (int j = 0; j < D; ++j) ip[j] = 0.0;
  double o0 = 0.0, o1 = 0.0, o2 = 0.0;
  const double czl = (sl < cnt) ? ssm[C.sCz + C.S - 1 - C.s0 - sl] : 0.0, ndl = (sl < cnt) ? ssm[C.sNd + C.S - 1 - C.s0 - sl] : 0.0;
  const unsigned mbar0 = C.mbar0, phase0 = C.phase0;
  unsigned wdst[CC];
#pragma unroll
  for (int r = 0; r < CC; ++r) wdst[r] = cl_mapa(C.wsh_s, (unsigned)r) + (unsigned)sl * 8u;
  for (int kb = 0; kb < n_blocks; ++kb) {
    if (!(C.dbg & 8)) mbar_wait(mbar0 + 8u * (kb % NUSI_CL_RB), cl_parity(phase0, kb));
    const double *slotB = ssm + C.pre + ((kb & 1) * B * PF) * 32 + sl;
    const double *ashB = ssm + C.ash + (kb & 1) * B * 32 + sl;
    const unsigned woffB = (unsigned)((kb & 1) * B * 32) * 8u;
    const int *tco = reinterpret_cast<const int *>(ssm + C.tcoff) + (kb & 1) * B;
    const int k0 = kb * B - sl;
    const bool all = (cnt == 32) && (kb * B - 31 >= 0) && (kb * B + B - 1 <= N - 1);
    if (C.dbg & 2) {  // timing experiment: no chain
#pragma unroll
      for (int r = 0; r < B; ++r)
#pragma unroll
        for (int q = 0; q < CC; ++q) cl_st_f64(wdst[q] + woffB + (unsigned)(r * 32) * 8u, 1e-3);
    } else if (all) {
#pragma unroll
      for (int r = 0; r < B; ++r)
        cl_wave<D, true, CC>(true, sl, cnt, N - 1 - (k0 + r), NP, U0, U1, U2, czl, ndl, slotB + r * PF * 32, ashB + r * 32, wdst, woffB + (unsigned)(r * 32) * 8u,
                         ssm + tco[r], fband, fb, ip, o0, o1, o2);
    } else {
#pragma unroll
      for (int r = 0; r < B; ++r) {
        const int k = k0 + r;
        const bool act = (sl < cnt) && (k >= 0) && (k <= N - 1);
        cl_wave<D, false, CC>(act, sl, cnt, act ? N - 1 - k : 0, NP, U0, U1, U2, czl, ndl, slotB + r * PF * 32, ashB + r * 32, wdst, woffB + (unsigned)(r * 32) * 8u,
                          ssm + tco[r], fband, fb, ip, o0, o1, o2);
      }
    }
    cl_sync();
  }
}

// ================= helper warps (rank 0): slots of block kb+1 =================
// what a helper fetches from global memory for one of its waves, two blocks before it is used: the source term of the
// lane's cell and the wave's Gamma / alphaTilde (the same for all lanes)
struct ClHelpIn {
  double src, G, A;
};
template <int B, int NHW>
__device__ __forceinline__ void cl_help_load(const ClCtx &C, int kb, int hw, int lane, ClHelpIn (&hs)[(B + NHW - 1) / NHW]) {
  const int zi = max(C.S - 1 - C.s0 - lane, 0);
#pragma unroll
  for (int q = 0; q < (B + NHW - 1) / NHW; ++q) {
    const int r = hw + q * NHW, w_rel = kb * B + r, b = min(max(C.N - 1 - (w_rel - lane), 0), C.N - 1);
    const int c = min(max(C.NT - 1 - (C.s0 + w_rel), 0), C.NT - 1);
    hs[q].src = __ldg(C.source + (size_t)zi * C.N + b);
    hs[q].G = __ldg(C.rowG + c);
    hs[q].A = __ldg(C.rowA + c);
  }
}
template <int B, int NHW>
__device__ __forceinline__ void cl_help_block(double *ssm, const ClCtx &C, int kb, int hw, int lane, unsigned pre0, const ClHelpIn (&hs)[(B + NHW - 1) / NHW]) {
  constexpr int RH = (B + NHW - 1) / NHW, PF = NUSI_CL_PF, D = 2 * B - 1;
  const int sl = lane, zi = max(C.S - 1 - C.s0 - sl, 0), N = C.N, NT = C.NT;
  if (hw == 0 && lane < B) {  // where row c of the wave's column sits in the near ring
    const int w_rel = kb * B + lane, c = NT - 1 - (C.s0 + w_rel), as = max(c - D, 0);
    reinterpret_cast<int *>(ssm + C.tcoff)[(kb & 1) * B + lane] =
        C.near + ((kb % NUSI_CL_RB) * B + lane) * NUSI_CL_NEAR + cl_mis(C.tpar, max(c, 0), as) + (c - as);
  }
  const double cz = ssm[C.sCz + zi], ndz = ssm[C.sNd + zi];
  const double U2f[3] = {C.U2f[0], C.U2f[1], C.U2f[2]};
  double Z[RH][3], rZ[RH][3], Gam[RH], aT[RH], dEb[RH], rdE[RH];
  bool ok = true;
#pragma unroll
  for (int q = 0; q < RH; ++q) {
    const int r = hw + q * NHW, w_rel = kb * B + r;
    const int b = min(max(N - 1 - (w_rel - sl), 0), N - 1);
    dEb[q] = ssm[C.sdE + b];
    rdE[q] = ssm[C.sRdE + b];
    Gam[q] = ndz * hs[q].G;
    aT[q] = ndz * hs[q].A;
#pragma unroll
    for (int kk = 0; kk < 3; ++kk) {
      const double x = cz * (Gam[q] * U2f[kk] - aT[q] * sqr(U2f[kk]));
      const double qq = quot_fast(x, dEb[q], rdE[q]);
      ok = ok && (fabs(qq) <= DBL_MAX);
      Z[q][kk] = 1.0 + qq;
    }
  }
  if (!ok) {
#pragma unroll
    for (int q = 0; q < RH; ++q)
#pragma unroll
      for (int kk = 0; kk < 3; ++kk) Z[q][kk] = 1.0 + cz * (Gam[q] * U2f[kk] - aT[q] * sqr(U2f[kk])) / dEb[q];
  }
  ok = true;
#pragma unroll
  for (int q = 0; q < RH; ++q)
#pragma unroll
    for (int kk = 0; kk < 3; ++kk) {
      rZ[q][kk] = recip_newton(Z[q][kk]);
      ok = ok && (fabs(Z[q][kk]) < 1e30) && (fabs(Z[q][kk]) > 1e-30);
    }
  if (!ok) {
#pragma unroll
    for (int q = 0; q < RH; ++q)
#pragma unroll
      for (int kk = 0; kk < 3; ++kk) rZ[q][kk] = 1.0 / Z[q][kk];
  }
#pragma unroll
  for (int q = 0; q < RH; ++q) {
    const int r = hw + q * NHW;
    if (r < B) {
      const double offdiag = fabs(aT[q]) * rdE[q] * fmax(fabs(rZ[q][0]), fmax(fabs(rZ[q][1]), fabs(rZ[q][2]))) * 1.00001;
      const unsigned o = pre0 + (unsigned)((((kb & 1) * B + r) * PF) * 32 + sl) * 8u;  // rank 0's slot (this CTA's own when rank == 0)
      const double fld[PF] = {cz * hs[q].src, offdiag, Gam[q], aT[q], Z[q][0], rZ[q][0], Z[q][1], rZ[q][1], Z[q][2], rZ[q][2], dEb[q], rdE[q]};
#pragma unroll
      for (int f = 0; f < PF; ++f) cl_st_f64(o + (unsigned)(f * 32) * 8u, fld[f]);
    }
  }
}

// Column of wave w_rel: c = NT-1-(s0+w_rel), rows 0..c-1 at packed offset c(c-1)/2.  One thread per CTA (lane 0 of the
// producer warp) issues the copies of block kb: this CTA's slice of each of the B columns (rows [a0, a0 + DPC) clipped
// to [0, c)), and on rank 0 the D rows below the diagonal as well.  Works on a register copy of the band's fields (the
// asm statements are memory barriers for the compiler: shared-memory fields would be re-read after every one).
struct ClIssue {
  int NT, s0, n_waves, a0, DPC, SLOT, rank;
  unsigned tpar, mbar0, ring0, near0;
  const double *T;
};
__device__ __forceinline__ ClIssue cl_issue_fields(const ClCtx &C) {
  ClIssue P;
  P.NT = C.NT; P.s0 = C.s0; P.n_waves = C.n_waves; P.a0 = C.a0; P.DPC = C.DPC; P.SLOT = C.SLOT; P.rank = C.rank;
  P.tpar = C.tpar; P.mbar0 = C.mbar0; P.ring0 = C.ring0; P.near0 = C.near0; P.T = C.T;
  return P;
}
// Called by the WHOLE producer warp: lane r (< B) copies this CTA's slice of column r of the block, lane B + r (rank 0 only)
// the D rows below the diagonal of that column; lane 0 announces the byte count first.
template <int B>
__device__ __forceinline__ void cl_issue_block(const ClIssue &P, int kb, int lane) {
  constexpr int D = 2 * B - 1;
  const int rb = kb % NUSI_CL_RB;
  const unsigned bar = P.mbar0 + 8u * rb;
  const int as = max(P.a0, 0);
  const int r = lane & (B - 1), kind = lane / B;  // B is a power of two
  const int w_rel = kb * B + r, c = P.NT - 1 - (P.s0 + w_rel);
  unsigned bytes = 0, dst = 0;
  const double *src = P.T;
  if (kind <= 1 && w_rel < P.n_waves && c > 0) {
    const size_t col = (size_t)c * (size_t)(c - 1) / 2;
    if (kind == 0) {
      const int len = min(P.a0 + P.DPC, c) - as;
      if (len > 0) {
        const int mis = cl_mis(P.tpar, c, as);
        bytes = (unsigned)((mis + len + 1) & ~1) * 8u;
        src = P.T + col + as - mis;
        dst = P.ring0 + (unsigned)((rb * B + r) * P.SLOT) * 8u;
      }
    } else if (P.rank == 0) {
      const int an = max(c - D, 0), mis = cl_mis(P.tpar, c, an);
      bytes = (unsigned)((mis + (c - an) + 1) & ~1) * 8u;
      src = P.T + col + an - mis;
      dst = P.near0 + (unsigned)((rb * B + r) * NUSI_CL_NEAR) * 8u;
    }
  }
  const unsigned total = __reduce_add_sync(0xffffffffu, bytes);
  if (lane == 0) mbar_expect_tx(bar, total);
  __syncwarp();
  if (bytes) tma_bulk_g2s(dst, src, bytes, bar);
}
// ================= producer warp (every rank): table copies of block kb+3 =================
template <int B>
static __device__ __noinline__ void cl_producer(double *ssm) {
  const ClCtx &C = *reinterpret_cast<const ClCtx *>(ssm);
  const int n_blocks = C.n_blocks;
  const int lane = threadIdx.x & 31;
  const bool on = !(C.dbg & 8);
  const ClIssue P = cl_issue_fields(C);
  const unsigned phase0 = C.phase0;
  for (int kb = 0; kb < n_blocks; ++kb) {
    if (on && kb + 3 < n_blocks) cl_issue_block<B>(P, kb + 3, lane);
    __syncwarp();
    cl_sync();
  }
  // nobody else waits for this CTA's copies of the band's last block: they must have landed before the ring is reused
  if (on && lane == 0 && n_blocks > 0) mbar_wait(P.mbar0 + 8u * ((n_blocks - 1) % NUSI_CL_RB), cl_parity(phase0, n_blocks - 1));
}

template <int B, int NHW>
static __device__ __noinline__ void cl_helper(double *ssm) {
  constexpr int NHT = NHW * NUSI_CL_HR, RH = (B + NHT - 1) / NHT;  // helper warps of the whole cluster; waves per helper and block
  const ClCtx &C = *reinterpret_cast<const ClCtx *>(ssm);
  const int lane = threadIdx.x & 31, hw = C.rank * NHW + (int)(threadIdx.x >> 5) - 1, n_blocks = C.n_blocks;
  const bool work = (C.rank < NUSI_CL_HR) && !(C.dbg & 4);
  const unsigned pre0 = cl_mapa((unsigned)__cvta_generic_to_shared(ssm + C.pre), 0u);
  ClHelpIn hs[RH], hsn[RH];
  if (work) {
    cl_help_load<B, NHT>(C, 1, hw, lane, hs);
    cl_help_load<B, NHT>(C, 2, hw, lane, hsn);
  }
  for (int kb = 0; kb < n_blocks; ++kb) {
    if (work) {
      if (kb + 1 < n_blocks) cl_help_block<B, NHT>(ssm, C, kb + 1, hw, lane, pre0, hs);
#pragma unroll
      for (int q = 0; q < RH; ++q) hs[q] = hsn[q];
      cl_help_load<B, NHT>(C, kb + 3, hw, lane, hsn);  // clamped indices: harmless beyond the last block
    }
    __syncwarp();
    cl_sync();
  }
}

// ================= bulk warps (every rank): sources of block kb-1, publish the diagonal of block kb+1 =================
// Thread u owns TWO neighbouring diagonals and HALF of the band's z-steps (the lower half of the bulk warps z-steps 0..15, the
// upper half 16..31): 2 x 16 running sums, as many registers as one diagonal x 32 z-steps, but every broadcast load of a pair
// of weighted fluxes feeds four FMAs instead of two (the update is bound by the shared-memory loads it has to wait for).
template <int B, int NBW, int NHW>
static __device__ __noinline__ void cl_bulk(double *ssm) {
  constexpr int D = 2 * B - 1, HALF = NBW * 16;  // HALF threads per z-half, each with two diagonals: 32 NBW diagonals per CTA
  const ClCtx &C = *reinterpret_cast<const ClCtx *>(ssm);
  double a0[16], a1[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) a0[j] = a1[j] = 0.0;
  const int N = C.N, u = (int)threadIdx.x - (2 + NHW) * 32;
  const int zh = (u >= HALF) ? 1 : 0, dp = u - zh * HALF;
  // Rank 0 -- the critical warp's CTA -- owns the HIGHEST diagonals: they are finalised in the first quarter of a band and
  // receive no update afterwards, so for three quarters of the band the chain shares its SM with nobody; the long-lived
  // low diagonals live on rank 3.
  const int d0 = -31 + (C.csize - 1 - C.rank) * C.DPC + 2 * dp;  // cells (sl, b = d + sl); the second diagonal is d0 + 1
  const bool own0 = (2 * dp < C.DPC) && (d0 <= N - 1), own1 = (2 * dp + 1 < C.DPC) && (d0 + 1 <= N - 1);
  const int n_blocks = C.n_blocks, n_waves = C.n_waves, cnt = C.cnt;
  const int wd0 = N - 1 - d0, wd1 = wd0 - 1;  // the waves in which the diagonals' cells are finalised
  const int row0 = C.a0 + 2 * dp;             // table row of d0 in this band (d0 + 1: the next row)
  const int as = max(C.a0, 0);
  const unsigned mbar0 = C.mbar0, phase0 = C.phase0, tpar = C.tpar;
  const unsigned ash0 = cl_mapa(C.ash_s, 0u) + (unsigned)(16 * zh) * 8u;  // rank 0's hand-over slots, this thread's z-half
  for (int kb = 0; kb < n_blocks; ++kb) {
    if (kb >= 1 && (own0 || own1) && !(C.dbg & 1)) {
      const int kbs = kb - 1;
      mbar_wait(mbar0 + 8u * (kbs % NUSI_CL_RB), cl_parity(phase0, kbs));
#pragma unroll
      for (int r = 0; r < B; ++r) {
        const int w_rel = kbs * B + r;
        const bool act0 = own0 && w_rel < n_waves && wd0 - w_rel > D && row0 >= 0;
        const bool act1 = own1 && w_rel < n_waves && wd1 - w_rel > D && row0 + 1 >= 0;
        const int hi = min(w_rel, cnt - 1) - 16 * zh;  // last z-step of this half that has started
        if ((act0 || act1) && hi >= 0) {
          const int c = C.NT - 1 - (C.s0 + w_rel);
          const double *col = ssm + C.ring + ((kbs % NUSI_CL_RB) * B + r) * C.SLOT + cl_mis(tpar, c, as) - as;
          const double t0 = act0 ? col[row0] : 0.0, t1 = act1 ? col[row0 + 1] : 0.0;  // an inactive diagonal adds 0 x W
          const double *Wp = ssm + C.wsh + ((kbs & 1) * B + r) * 32 + 16 * zh;
          const unsigned wp_s = (unsigned)__cvta_generic_to_shared(Wp);
          if (hi >= 15) {
#pragma unroll
            for (int s4 = 0; s4 < 16; s4 += 4) {
              const double2 wa = lds_f64x2(wp_s + (unsigned)s4 * 8u), wb = lds_f64x2(wp_s + (unsigned)(s4 + 2) * 8u);
              a0[s4] = fma(t0, wa.x, a0[s4]);
              a1[s4] = fma(t1, wa.x, a1[s4]);
              a0[s4 + 1] = fma(t0, wa.y, a0[s4 + 1]);
              a1[s4 + 1] = fma(t1, wa.y, a1[s4 + 1]);
              a0[s4 + 2] = fma(t0, wb.x, a0[s4 + 2]);
              a1[s4 + 2] = fma(t1, wb.x, a1[s4 + 2]);
              a0[s4 + 3] = fma(t0, wb.y, a0[s4 + 3]);
              a1[s4 + 3] = fma(t1, wb.y, a1[s4 + 3]);
            }
          } else {
#pragma unroll
            for (int sl = 0; sl < 16; ++sl)
              if (sl <= hi) {
                const double w = Wp[sl];
                a0[sl] = fma(t0, w, a0[sl]);
                a1[sl] = fma(t1, w, a1[sl]);
              }
          }
        }
      }
    }
    {  // hand the sums of a diagonal that is finalised in block kb+1 to the critical warp (this thread's 16 z-steps of it)
      const int lo_b = (kb + 1) * B, hi_b = (kb + 2) * B;
      if (own0 && wd0 >= lo_b && wd0 < hi_b && wd0 < n_waves) {
        const unsigned o = ash0 + (unsigned)((((kb + 1) & 1) * B + (wd0 - lo_b)) * 32) * 8u;
#pragma unroll
        for (int sl = 0; sl < 16; sl += 2) cl_st_f64x2(o + (unsigned)sl * 8u, a0[sl], a0[sl + 1]);
      }
      if (own1 && wd1 >= lo_b && wd1 < hi_b && wd1 < n_waves) {
        const unsigned o = ash0 + (unsigned)((((kb + 1) & 1) * B + (wd1 - lo_b)) * 32) * 8u;
#pragma unroll
        for (int sl = 0; sl < 16; sl += 2) cl_st_f64x2(o + (unsigned)sl * 8u, a1[sl], a1[sl + 1]);
      }
    }
    __syncwarp();
    cl_sync();
  }
}

// B waves per block; NBW bulk warps per CTA (N_bins_E + 31 <= 4 x 32 NBW); NHW helper warps
template <int B, int NBW, int NHW, int CC>
__global__ void __launch_bounds__((2 + NHW + NBW) * 32, (NBW <= 4 ? 2 : 1))
integrate_cluster_kernel(DeviceGrid G, const PointParams *__restrict__ params, const double *__restrict__ rows,
                         const double *__restrict__ tri, size_t tstride, double *__restrict__ flux_mass,
                         double *__restrict__ flux_fla, const int *__restrict__ out_index, const int *__restrict__ status_in,
                         int *__restrict__ status_out, int dbg) {
  constexpr int NTHR = (2 + NHW + NBW) * 32;  // critical | helpers | producer | bulk
  extern __shared__ __align__(16) double ssm[];
  const int tid = threadIdx.x, warp = tid >> 5;
  const int rank = (int)cl_rank();
  const int p = blockIdx.x / CC;
  const int N = G.N, NT = G.NT, S = G.Nz - 1;
  const int DPC = ((N + 31 + CC - 1) / CC + 1) & ~1;  // diagonals per CTA (even: a bulk thread owns two)
  const bool dead = !(params[p].mphi > 0);
  const ClLayout L = cl_layout(N, S, B, DPC);
  ClCtx &C = *reinterpret_cast<ClCtx *>(ssm);
  double *sdE = ssm + L.sdE, *sRdE = ssm + L.sRdE, *sCz = ssm + L.sCz, *sNd = ssm + L.sNd, *fband = ssm + L.fband;
  const int NP = (N + 1) & ~1;
  const unsigned mbar0 = (unsigned)__cvta_generic_to_shared(ssm + L.mbar);

  for (int i = NUSI_CL_CTX_DOUBLES + tid; i < L.pre; i += NTHR) ssm[i] = 0.0;  // rings, pad, W and hand-over slots
  for (int i = tid; i < S; i += NTHR) {
    sCz[i] = G.cz[i];
    sNd[i] = G.ndz[i];
  }
  for (int i = tid; i < N; i += NTHR) {
    sdE[i] = G.dE[i];
    sRdE[i] = 1.0 / G.dE[i];
    fband[i] = fband[NP + i] = fband[2 * NP + i] = 0.0;
  }
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < NUSI_CL_RB; ++i) mbar_init(mbar0 + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const PointParams P = params[p];
    C.N = N; C.NT = NT; C.S = S; C.NP = NP; C.SLOT = L.SLOT; C.DPC = DPC; C.rank = rank; C.dbg = dbg; C.csize = CC;
    C.ring = L.ring; C.near = L.near; C.wsh = L.wsh; C.ash = L.ash; C.pre = L.pre; C.sdE = L.sdE; C.sRdE = L.sRdE;
    C.sCz = L.sCz; C.sNd = L.sNd; C.fband = L.fband; C.tcoff = L.tcoff; C.fb = L.fb;
    C.mbar0 = mbar0;
    C.ring0 = (unsigned)__cvta_generic_to_shared(ssm + L.ring);
    C.near0 = (unsigned)__cvta_generic_to_shared(ssm + L.near);
    C.wsh_s = (unsigned)__cvta_generic_to_shared(ssm + L.wsh);
    C.ash_s = (unsigned)__cvta_generic_to_shared(ssm + L.ash);
    C.T = tri + (size_t)p * tstride;
    C.tpar = (unsigned)(((size_t)C.T >> 3) & 1);  // parity (in doubles) of the table's base address
    C.source = G.source;
    C.rowG = rows + ((size_t)p * 2 + 0) * NT;
    C.rowA = rows + ((size_t)p * 2 + 1) * NT;
    C.U2f[0] = P.U2f[0]; C.U2f[1] = P.U2f[1]; C.U2f[2] = P.U2f[2];
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the zeroed rings are written by the async proxy from here on
  __syncthreads();
  cl_sync();  // every CTA's shared memory is initialised before anybody writes into it remotely

  unsigned ring_phase = 0;  // bit i = parity of the next completion of ring block i (tracked identically by every thread)
  for (int s0 = 0; s0 < S && !dead; s0 += 32) {
    const int cnt = min(32, S - s0);
    const int n_waves = cnt - 1 + N, n_blocks = (n_waves + B - 1) / B;
    if (tid == 0) {
      C.s0 = s0; C.cnt = cnt; C.n_waves = n_waves; C.n_blocks = n_blocks; C.rowoff = S - 1 - s0; C.phase0 = ring_phase;
      C.a0 = S - 1 - s0 - 31 + (CC - 1 - rank) * DPC;
    }
    for (int i = tid; i < 2 * B * 32; i += NTHR) ssm[L.ash + i] = 0.0;
    __syncthreads();
    if (warp == 1 + NHW) {  // TMA producer warp of this CTA
      const ClIssue P = cl_issue_fields(C);
      for (int kb = 0; kb < 3 && kb < n_blocks && !(dbg & 8); ++kb) cl_issue_block<B>(P, kb, tid & 31);
    }
    if (rank < NUSI_CL_HR && warp >= 1 && warp <= NHW) {
      constexpr int NHT = NHW * NUSI_CL_HR;
      ClHelpIn hs0[(B + NHT - 1) / NHT];
      cl_help_load<B, NHT>(C, 0, rank * NHW + warp - 1, tid & 31, hs0);
      cl_help_block<B, NHT>(ssm, C, 0, rank * NHW + warp - 1, tid & 31, cl_mapa((unsigned)__cvta_generic_to_shared(ssm + L.pre), 0u), hs0);
    }
    __syncthreads();
    cl_sync();
    if (warp == 0) {
      if (rank == 0) {
        cl_critical<B, CC>(ssm);
      } else {
        for (int kb = 0; kb < n_blocks; ++kb) cl_sync();
      }
    } else if (warp <= NHW) {
      cl_helper<B, NHW>(ssm);
    } else if (warp == 1 + NHW) {
      cl_producer<B>(ssm);
    } else {
      cl_bulk<B, NBW, NHW>(ssm);
    }
#pragma unroll
    for (int i = 0; i < NUSI_CL_RB; ++i)
      if (i < n_blocks && (((n_blocks - 1 - i) / NUSI_CL_RB + 1) & 1)) ring_phase ^= 1u << i;
    __syncthreads();
    cl_sync();
  }

  if (rank == 0) {  // bin-width divide and flavour projection (nuSIprop.hpp:329-336) into this point's output slot
    const int q = out_index ? out_index[p] : p;
    for (int b = tid; b < N; b += NTHR) {
      const double f[3] = {fband[b], fband[NP + b], fband[2 * NP + b]};
      store_bin(G, flux_mass, flux_fla, q, b, f, sdE[b], dead);
    }
    if (status_out && tid == 0) status_out[q] = status_in[p];
  }
}

bool integrate_cluster_supports(int N, int NT) { return N > 224 && N + 31 <= NUSI_CL_C * 32 * 16 && NT > 0; }

template <int B, int NBW, int NHW, int CC>
static cudaError_t launch_cluster_t(const IntegArgs &A, cudaStream_t s, int dbg) {
  const DeviceGrid &G = A.G;
  const int DPC = ((G.N + 31 + CC - 1) / CC + 1) & ~1;
  const ClLayout L = cl_layout(G.N, G.Nz - 1, B, DPC);
  const size_t sh = (size_t)L.total * sizeof(double);
  auto kern = integrate_cluster_kernel<B, NBW, NHW, CC>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)A.cnt * CC);
  cfg.blockDim = dim3((2 + NHW + NBW) * 32);
  cfg.dynamicSmemBytes = sh;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CC;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, G, A.params, A.rows, A.tri, A.tstride, A.flux_mass, A.flux_fla, A.out_index, A.status_in, A.status_out, dbg);
}

// A band of 32 z-steps needs N_bins_E + 31 running sums per z-step: up to 1024 diagonals over a cluster of 2 CTAs (twice as
// many points in flight), up to 2048 over a cluster of 4.
// variant: 0 = the product configuration; 12 = clusters of 4 at every binning; 20 + bits = timing experiments with wrong results
// (1 no bulk update, 2 no critical chain, 4 no helper work, 8 no table copies)
cudaError_t launch_integrate_cluster(const IntegArgs &A, int variant, cudaStream_t s) {
  const int need = A.G.N + 31;  // diagonals of a band
  const int dbg = variant >= 20 ? variant - 20 : 0;
  if (variant != 12) {
    if (need <= 2 * 32 * 4) return launch_cluster_t<4, 4, 2, 2>(A, s, dbg);
    if (need <= 2 * 32 * 8) return launch_cluster_t<4, 8, 2, 2>(A, s, dbg);
    if (need <= 2 * 32 * 16) return launch_cluster_t<4, 16, 2, 2>(A, s, dbg);
  } else {
    if (need <= 4 * 32 * 4) return launch_cluster_t<4, 4, 2, 4>(A, s, dbg);
    if (need <= 4 * 32 * 8) return launch_cluster_t<4, 8, 2, 4>(A, s, dbg);
  }
  if (need <= 4 * 32 * 16) return variant == 11 ? launch_cluster_t<4, 16, 1, 4>(A, s, 0) : launch_cluster_t<4, 16, 2, 4>(A, s, dbg);
  return cudaErrorInvalidValue;
}

}  // namespace nusi
