// k_integrate_split.cu -- the implicit z-stepping integrator (reference nuSIprop.hpp:255-336), role-split wavefront, one CTA per
// point: the product kernel for 32 <= N_bins_E <= 576.
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
//
// Same wavefront as integrate_diag_kernel (cell (z-step s, bin m) on wave w = s + (N-1-m); all z-steps of a wave
// read table column c = NT-1-w), but the wave-to-wave dependency chain is carried by ONE warp instead of by two
// block-wide barriers per wave:
//
//   critical warp   lane = z-step of the band (32 lanes = 32 z-steps).  Per wave each lane finalises its cell: the
//                   flux of the previous z-step arrives from the lane below by warp shuffle, the flux-independent half
//                   of the implicit step from a slot prepared by the helpers, and the regeneration sum as
//                       (bulk part: source bins more than D waves back)  +  (the last D source bins, kept in registers).
//                   The lane then adds its new weighted flux W into its own D running sums (right-looking; the table
//                   values T[c-j][c], j = 1..D, are the same for all lanes: broadcast loads from the column ring).
//                   Nothing on this chain waits for another warp.
//   helper warps    prepare the slots (3 denominators Zdr_k, their reciprocals, source term) one block of B waves ahead.
//   bulk warps      one thread per table DIAGONAL (as in integrate_diag_kernel): for every source wave more than D
//                   waves behind a cell, acc[s] += T[row(d)][c] W_s for the 32 z-steps, one block of B waves behind the
//                   critical warp, and hand the finished sums of the next block's diagonals to the critical warp.
//   one CTA barrier per block of B waves (D = 2B-1) instead of two per wave.
//
// The table columns are streamed through a shared-memory ring by TMA bulk copies (cp.async.bulk, one per column, issued
// by one thread and completed on an mbarrier) three blocks ahead.
//
// A cell's regeneration sum is the sum of two partial sums, each accumulated in the reference's order (source bins from
// high to low): it differs from the strictly sequential sum of integrate_kernel by rounding only (<= 1e-15 relative:
// every term is a product of a non-negative table entry and a non-negative flux).
#include "nusi_launch.h"
#include "nusi_implicit.h"
#include "nusi_split.h"

namespace nusi {

#ifndef NUSI_SPLIT_MAXREG
#define NUSI_SPLIT_MAXREG 96  // 9 warps per CTA (critical, helper, 7 bulk), 2 CTAs per SM (112 registers: 5.9 instead of 3.55 ms per 8192 points)
#endif
#define NUSI_SPLIT_PF 12  // slot fields: lum, offdiag, Gam, aT, Z0, rZ0, Z1, rZ1, Z2, rZ2, dE, 1/dE
#define NUSI_SPLIT_RB 5   // ring depth in blocks of B columns (block kb-1 .. kb+3)
#define NUSI_SPLIT_PAD 16 // doubles in front of the ring: the critical warp reads up to D rows below row 0 of a column
                          // (targets that do not exist; the values are never used)

// Everything the three roles share about the point and the current band; lives at the start of dynamic shared memory
// (filled by thread 0), so that each role -- a separate, non-inlined function with its own register allocation -- reads
// what it needs with broadcast loads instead of keeping it live in registers.
struct SplitCtx {
  int N, NT, S, NP, NTP;
  int s0, cnt, n_waves, n_blocks, rowoff;
  int ring, wsh, ash, pre, sG, sA, sdE, sRdE, sCz, sNd, fband, tcoff, fb;  // offsets in doubles
  unsigned mbar0, ring0, phase0, tpar;
  int dbg, pad_;
  double U2f[3];
  const double *T, *source;
};
#define NUSI_SPLIT_CTX_DOUBLES ((int)((sizeof(SplitCtx) + 15) / 16) * 2)

struct SplitLayout {  // offsets in doubles into dynamic shared memory
  int NTP, ring, wsh, ash, pre, sG, sA, sdE, sRdE, sCz, sNd, fband, mbar, tcoff, fb, total;
};
__host__ __device__ inline SplitLayout split_layout(int N, int NT, int S, int B) {
  SplitLayout L;
  L.NTP = (NT + 3) & ~1;  // a column, possibly shifted by one double for 16-byte alignment of its global source
  int o = NUSI_SPLIT_CTX_DOUBLES + NUSI_SPLIT_PAD;
  L.ring = o; o += NUSI_SPLIT_RB * B * L.NTP;
  L.wsh = o; o += 2 * B * 32;
  L.ash = o; o += 2 * B * 32;
  L.pre = o; o += 2 * B * NUSI_SPLIT_PF * 32;
  L.sG = o; o += (NT + 1) & ~1;
  L.sA = o; o += (NT + 1) & ~1;
  L.sdE = o; o += (N + 1) & ~1;
  L.sRdE = o; o += (N + 1) & ~1;
  L.sCz = o; o += (S + 1) & ~1;
  L.sNd = o; o += (S + 1) & ~1;
  L.fband = o; o += 3 * ((N + 1) & ~1);
  L.mbar = o; o += NUSI_SPLIT_RB + 1;
  L.tcoff = o; o += B;        // 2 x B ints: per wave of a block, where row c of its column sits in shared memory
  L.fb = o; o += 4 * 32;      // results of the out-of-line general step: f0, f1, f2, W per lane
  L.total = o;
  return L;
}

// One wave of the critical warp.  ALL: every lane has a live cell (warp-uniform; the steady state), so nothing is selected.
template <int D, bool ALL>
__device__ __forceinline__ void split_wave(bool act, int sl, int cnt, int bb, int NP, double U0, double U1, double U2, double czl, double ndl,
                                           const double *__restrict__ slot, const double *__restrict__ ash, double *__restrict__ wsh,
                                           const double *__restrict__ tc, double *__restrict__ fband, double *__restrict__ fb, double (&ip)[D],
                                           double &o0, double &o1, double &o2) {
  // flux of this bin after the previous z-step: from the lane below (finalised one wave ago), or the previous band's
  // result for the first z-step of the band
  const double u0 = __shfl_up_sync(0xffffffffu, o0, 1), u1 = __shfl_up_sync(0xffffffffu, o1, 1), u2 = __shfl_up_sync(0xffffffffu, o2, 1);
  const double f0 = (sl == 0) ? fband[bb] : u0, f1 = (sl == 0) ? fband[NP + bb] : u1, f2 = (sl == 0) ? fband[2 * NP + bb] : u2;
  const double lum = slot[0], offdiag = slot[32];
  const double Z0 = slot[4 * 32], r0 = slot[5 * 32], Z1 = slot[6 * 32], r1 = slot[7 * 32], Z2 = slot[8 * 32], r2 = slot[9 * 32];
  const double dEb = slot[10 * 32], rdE = slot[11 * 32];
  const double acc_c = ash[0] + ip[0];
  const double base = czl * (ndl * acc_c);
  const double s0 = lum + base * U0, s1 = lum + base * U1, s2 = lum + base * U2;
  double v0 = quot_fast(f0 + s0, Z0, r0), v1 = quot_fast(f1 + s1, Z1, r1), v2 = quot_fast(f2 + s2, Z2, r2);
  // speculative: the 3x3 system is the identity to rounding (checked below from the high words alone)
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
  if (ALL) {
    o0 = v0;
    o1 = v1;
    o2 = v2;
    wsh[0] = W;
#pragma unroll
    for (int j = 1; j < D; ++j) ip[j - 1] = fma(tc[-j], W, ip[j]);  // table rows c-1 .. c-D of column c: the same for every lane
    ip[D - 1] = tc[-D] * W;
  } else {
    if (act) {
      o0 = v0;
      o1 = v1;
      o2 = v2;
    }
    wsh[0] = act ? W : 0.0;
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

template <int NTHR>
__device__ __forceinline__ unsigned split_parity(unsigned phase0, int kb) {  // parity on which block kb of this band completes
  const int rb = kb % NUSI_SPLIT_RB;
  return ((phase0 >> rb) & 1u) ^ (unsigned)((kb / NUSI_SPLIT_RB) & 1);
}
// packed offset c(c-1)/2 of column c is odd iff c mod 4 is 2 or 3: the column's data starts mis doubles into its slot
__device__ __forceinline__ int split_mis(unsigned tpar, int c) { return (int)((tpar ^ (unsigned)(c >> 1)) & 1u); }
// row a of the column of wave w_rel sits at ssm[split_col_off(..) + a]
template <int B>
__device__ __forceinline__ int split_col_off(const SplitCtx &C, int w_rel) {
  const int c = C.NT - 1 - (C.s0 + w_rel), kb = w_rel / B;
  return C.ring + ((kb % NUSI_SPLIT_RB) * B + (w_rel - kb * B)) * C.NTP + split_mis(C.tpar, c);
}

// ================= critical warp =================
template <int B, int NTHR>
static __device__ __noinline__ void split_critical(double *ssm) {
  constexpr int D = 2 * B - 1, PF = NUSI_SPLIT_PF;
  const SplitCtx &C = *reinterpret_cast<const SplitCtx *>(ssm);
  const int sl = threadIdx.x & 31, N = C.N, NP = C.NP, cnt = C.cnt, n_blocks = C.n_blocks;
  const double U0 = C.U2f[0], U1 = C.U2f[1], U2 = C.U2f[2];
  double *fband = ssm + C.fband, *fb = ssm + C.fb + sl;
  double ip[D];  // running sums of the next D cells of this lane (sources at most D waves back)
#pragma unroll
  for (int j = 0; j < D; ++j) ip[j] = 0.0;
  double o0 = 0.0, o1 = 0.0, o2 = 0.0;  // flux of the cell finalised in the previous wave
  const double czl = (sl < cnt) ? ssm[C.sCz + C.S - 1 - C.s0 - sl] : 0.0, ndl = (sl < cnt) ? ssm[C.sNd + C.S - 1 - C.s0 - sl] : 0.0;
  const unsigned mbar0 = C.mbar0, phase0 = C.phase0;
  const int dbg = C.dbg;
  for (int kb = 0; kb < n_blocks; ++kb) {
    mbar_wait(mbar0 + 8u * (kb % NUSI_SPLIT_RB), split_parity<NTHR>(phase0, kb));
    const double *slotB = ssm + C.pre + ((kb & 1) * B * PF) * 32 + sl;
    const double *ashB = ssm + C.ash + (kb & 1) * B * 32 + sl;
    double *wshB = ssm + C.wsh + (kb & 1) * B * 32 + sl;
    const int *tco = reinterpret_cast<const int *>(ssm + C.tcoff) + (kb & 1) * B;
    const int k0 = kb * B - sl;
    // every lane live in all B waves of the block
    const bool all = (cnt == 32) && (kb * B - 31 >= 0) && (kb * B + B - 1 <= N - 1);
    if (dbg & 2) {
      for (int r = 0; r < B; ++r) wshB[r * 32] = 1e-3;
    } else if (all) {
#pragma unroll
      for (int r = 0; r < B; ++r)
        split_wave<D, true>(true, sl, cnt, N - 1 - (k0 + r), NP, U0, U1, U2, czl, ndl, slotB + r * PF * 32, ashB + r * 32, wshB + r * 32, ssm + tco[r], fband, fb,
                            ip, o0, o1, o2);
    } else {
#pragma unroll
      for (int r = 0; r < B; ++r) {
        const int k = k0 + r;
        const bool act = (sl < cnt) && (k >= 0) && (k <= N - 1);
        split_wave<D, false>(act, sl, cnt, act ? N - 1 - k : 0, NP, U0, U1, U2, czl, ndl, slotB + r * PF * 32, ashB + r * 32, wshB + r * 32, ssm + tco[r], fband,
                             fb, ip, o0, o1, o2);
      }
    }
    block_sync<NTHR>();
  }
}

// ================= helper warps: slots of block kb+1, TMA for block kb+3 =================
// Helper warp hw owns the waves r = hw, hw + NHW, .. of a block (RH of them).  All of them are computed together,
// branch-free and for every lane (lanes without a cell work on clamped indices; their slots are never used), so the
// RH independent chains interleave; the source terms are loaded one call earlier.
template <int B, int NHW>
__device__ __forceinline__ void split_help_load(const SplitCtx &C, int kb, int hw, int lane, double (&hs)[(B + NHW - 1) / NHW]) {
  const int zi = max(C.S - 1 - C.s0 - lane, 0);
#pragma unroll
  for (int q = 0; q < (B + NHW - 1) / NHW; ++q) {
    const int r = hw + q * NHW, b = min(max(C.N - 1 - (kb * B + r - lane), 0), C.N - 1);
    hs[q] = __ldg(C.source + (size_t)zi * C.N + b);
  }
}
template <int B, int NHW>
__device__ __forceinline__ void split_help_block(double *ssm, const SplitCtx &C, int kb, int hw, int lane, const double (&hs)[(B + NHW - 1) / NHW]) {
  constexpr int RH = (B + NHW - 1) / NHW, PF = NUSI_SPLIT_PF;
  const int sl = lane, zi = max(C.S - 1 - C.s0 - sl, 0), N = C.N, NT = C.NT;
  if (hw == 0 && lane < B) {
    const int w_rel = kb * B + lane;
    reinterpret_cast<int *>(ssm + C.tcoff)[(kb & 1) * B + lane] = split_col_off<B>(C, w_rel) + (NT - 1 - (C.s0 + w_rel));
  }
  const double cz = ssm[C.sCz + zi], ndz = ssm[C.sNd + zi];
  const double U2f[3] = {C.U2f[0], C.U2f[1], C.U2f[2]};
  double Z[RH][3], rZ[RH][3], Gam[RH], aT[RH], dEb[RH], rdE[RH];
  bool ok = true;
#pragma unroll
  for (int q = 0; q < RH; ++q) {
    const int r = hw + q * NHW, w_rel = kb * B + r;
    const int b = min(max(N - 1 - (w_rel - sl), 0), N - 1), c = min(max(NT - 1 - (C.s0 + w_rel), 0), NT - 1);
    dEb[q] = ssm[C.sdE + b];
    rdE[q] = ssm[C.sRdE + b];
    Gam[q] = ndz * ssm[C.sG + c];
    aT[q] = ndz * ssm[C.sA + c];
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
      double *o = ssm + C.pre + (((kb & 1) * B + r) * PF) * 32 + sl;
      o[0 * 32] = cz * hs[q]; o[1 * 32] = offdiag; o[2 * 32] = Gam[q]; o[3 * 32] = aT[q];
      o[4 * 32] = Z[q][0]; o[5 * 32] = rZ[q][0]; o[6 * 32] = Z[q][1]; o[7 * 32] = rZ[q][1]; o[8 * 32] = Z[q][2]; o[9 * 32] = rZ[q][2];
      o[10 * 32] = dEb[q]; o[11 * 32] = rdE[q];
    }
  }
}
// column of wave w_rel: c = NT-1-(s0+w_rel); rows 0..c-1 at packed offset c(c-1)/2.  Issue block kb's B columns: called by a
// WHOLE warp, lane r (< B) copies column r, lane 0 announces the byte count first.  (No proxy fence per block: the ring slot's
// last readers are ordered before this call by the block barrier, as in any consumer-release / producer-refill pipeline.)
template <int B>
__device__ __forceinline__ void split_issue_block(const SplitCtx &C, int kb, int lane) {
  const int rb = kb % NUSI_SPLIT_RB;
  const unsigned bar = C.mbar0 + 8u * rb;
  const int r = lane, w_rel = kb * B + r, c = C.NT - 1 - (C.s0 + w_rel);
  unsigned bytes = 0;
  int mis = 0;
  if (r < B && w_rel < C.n_waves && c > 0) {
    mis = split_mis(C.tpar, c);
    bytes = (unsigned)((mis + c + 1) & ~1) * 8u;
  }
  const unsigned total = __reduce_add_sync(0xffffffffu, bytes);
  if (lane == 0) mbar_expect_tx(bar, total);
  __syncwarp();
  if (bytes) tma_bulk_g2s(C.ring0 + (unsigned)((rb * B + r) * C.NTP) * 8u, C.T + (size_t)c * (c - 1) / 2 - mis, bytes, bar);
}
template <int B, int NHW, int NTHR>
static __device__ __noinline__ void split_helper(double *ssm) {
  constexpr int RH = (B + NHW - 1) / NHW;
  const SplitCtx &C = *reinterpret_cast<const SplitCtx *>(ssm);
  const int lane = threadIdx.x & 31, hw = (threadIdx.x >> 5) - 1, n_blocks = C.n_blocks;
  double hs[RH], hsn[RH];  // source terms of the block being prepared / of the next one
  split_help_load<B, NHW>(C, 1, hw, lane, hs);
  split_help_load<B, NHW>(C, 2, hw, lane, hsn);
  for (int kb = 0; kb < n_blocks; ++kb) {
    if (hw == 0 && kb + 3 < n_blocks) split_issue_block<B>(C, kb + 3, lane);
    if (kb + 1 < n_blocks) split_help_block<B, NHW>(ssm, C, kb + 1, hw, lane, hs);
#pragma unroll
    for (int q = 0; q < RH; ++q) hs[q] = hsn[q];
    split_help_load<B, NHW>(C, kb + 3, hw, lane, hsn);  // clamped indices: harmless beyond the last block
    block_sync<NTHR>();
  }
}

// ================= bulk warps: sources of block kb-1, publish the diagonals of block kb+1 =================
// Thread u owns diagonal d = u (cells (sl, b = u + sl)), u = 0 .. N-1.  The 31 short diagonals at either end share
// threads: d = -j (cells sl >= j, finalised in the band's last waves) lives in thread u = N - j next to d' = N - j (cells
// sl <= j-1, finalised in wave j-1): their cells use disjoint slots of acc[], and d' has received its last update before the
// first z-step of d has started (wave j), so one table value per wave still serves the whole thread.
template <int B, int NHW, int NTHR>
static __device__ __noinline__ void split_bulk(double *ssm) {
  constexpr int D = 2 * B - 1;
  const SplitCtx &C = *reinterpret_cast<const SplitCtx *>(ssm);
  double acc[32];  // sums of this thread's diagonal(s) for the 32 z-steps of the band
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = 0.0;
  const int N = C.N, u = (int)threadIdx.x - (1 + NHW) * 32;
  const int jsh = N - u;                         // 1..31: this thread also owns the short diagonal d = -jsh
  const bool pair = (jsh >= 1 && jsh <= 31);
  const int n_blocks = C.n_blocks, n_waves = C.n_waves, cnt = C.cnt;
  const int wd_hi = N - 1 - u;                   // wave in which diagonal u is finalised (pair: j-1)
  const int wd_lo = N - 1 + jsh;                 // wave in which diagonal -jsh is finalised
  const int row_hi = u + C.rowoff, row_lo = -jsh + C.rowoff;
  const unsigned mbar0 = C.mbar0, phase0 = C.phase0;
  const bool skip = (C.dbg & 1) != 0, owner = (u < N);
  for (int kb = 0; kb < n_blocks; ++kb) {
    if (kb >= 1 && !skip && owner) {
      const int kbs = kb - 1;
      mbar_wait(mbar0 + 8u * (kbs % NUSI_SPLIT_RB), split_parity<NTHR>(phase0, kbs));
#pragma unroll
      for (int r = 0; r < B; ++r) {
        const int w_rel = kbs * B + r;
        const bool lo = pair && (w_rel >= jsh);  // past wave j-1 the thread works for its second diagonal
        const int wd = lo ? wd_lo : wd_hi, row = lo ? row_lo : row_hi;
        if (w_rel < n_waves && wd - w_rel > D && row >= 0) {  // distance of the diagonal's cells from the source wave
          const double tv = ssm[split_col_off<B>(C, w_rel) + row];
          const double *Wp = ssm + C.wsh + ((kbs & 1) * B + r) * 32;
          const int hi = min(w_rel, cnt - 1);  // z-steps that have started
          const unsigned wp_s = (unsigned)__cvta_generic_to_shared(Wp);
          if (hi == 31) {
            // W in batches of 8 (four 16-byte broadcast loads in flight), so that the loads are not serialised behind
            // the FMAs through a single temporary register pair
#pragma unroll
            for (int s8 = 0; s8 < 32; s8 += 8) {
              double2 w2[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) w2[q] = lds_f64x2(wp_s + (unsigned)(s8 + 2 * q) * 8u);
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                acc[s8 + 2 * q] = fma(tv, w2[q].x, acc[s8 + 2 * q]);
                acc[s8 + 2 * q + 1] = fma(tv, w2[q].y, acc[s8 + 2 * q + 1]);
              }
            }
          } else {
#pragma unroll
            for (int sl = 0; sl < 32; ++sl)
              if (sl <= hi) acc[sl] = fma(tv, Wp[sl], acc[sl]);
          }
        }
      }
    }
    // publish the diagonal finalised in block kb+1 (each of the thread's diagonals in its own time)
    if (owner) {
      const int lo_b = (kb + 1) * B, hi_b = (kb + 2) * B;
      const int wd = (pair && wd_lo >= lo_b && wd_lo < hi_b) ? wd_lo : wd_hi;
      if (wd >= lo_b && wd < hi_b && wd < n_waves) {
        double *o = ssm + C.ash + (((kb + 1) & 1) * B + (wd - lo_b)) * 32;
#pragma unroll
        for (int sl = 0; sl < 32; sl += 2) *reinterpret_cast<double2 *>(o + sl) = make_double2(acc[sl], acc[sl + 1]);
      }
    }
    block_sync<NTHR>();
  }
}

// B waves per block; NBW bulk warps (N_bins_E <= 32 NBW); NHW helper warps
template <int B, int NBW, int NHW>
__global__ void __maxnreg__(NUSI_SPLIT_MAXREG)
integrate_split_kernel(DeviceGrid G, const PointParams *__restrict__ params, const double *__restrict__ rows,
                       const double *__restrict__ tri, size_t tstride, double *__restrict__ flux_mass,
                       double *__restrict__ flux_fla, const int *__restrict__ out_index, const int *__restrict__ status_in,
                       int *__restrict__ status_out, int dbg) {
  constexpr int NTHR = (1 + NHW + NBW) * 32;
  extern __shared__ __align__(16) double ssm[];
  const int tid = threadIdx.x, warp = tid >> 5;
  const int p = blockIdx.x;
  const int N = G.N, NT = G.NT, S = G.Nz - 1;
  const bool dead = !(params[p].mphi > 0);
  const SplitLayout L = split_layout(N, NT, S, B);
  SplitCtx &C = *reinterpret_cast<SplitCtx *>(ssm);
  double *sG = ssm + L.sG, *sA = ssm + L.sA, *sdE = ssm + L.sdE, *sRdE = ssm + L.sRdE, *sCz = ssm + L.sCz, *sNd = ssm + L.sNd;
  double *fband = ssm + L.fband;
  const int NP = (N + 1) & ~1;  // stride of fband rows
  const unsigned mbar0 = (unsigned)__cvta_generic_to_shared(ssm + L.mbar);

  for (int i = tid; i < NT; i += NTHR) {
    sG[i] = rows[((size_t)p * 2 + 0) * NT + i];
    sA[i] = rows[((size_t)p * 2 + 1) * NT + i];
  }
  for (int i = tid; i < S; i += NTHR) {
    sCz[i] = G.cz[i];
    sNd[i] = G.ndz[i];
  }
  for (int i = tid; i < N; i += NTHR) {
    sdE[i] = G.dE[i];
    sRdE[i] = 1.0 / G.dE[i];
    fband[i] = fband[NP + i] = fband[2 * NP + i] = 0.0;
  }
  for (int i = NUSI_SPLIT_CTX_DOUBLES + tid; i < L.ring; i += NTHR) ssm[i] = 0.0;  // the pad in front of the ring
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the ring is written by the async proxy from here on
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < NUSI_SPLIT_RB; ++i) mbar_init(mbar0 + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const PointParams P = params[p];
    C.N = N; C.NT = NT; C.S = S; C.NP = NP; C.NTP = L.NTP;
    C.ring = L.ring; C.wsh = L.wsh; C.ash = L.ash; C.pre = L.pre; C.sG = L.sG; C.sA = L.sA; C.sdE = L.sdE; C.sRdE = L.sRdE;
    C.sCz = L.sCz; C.sNd = L.sNd; C.fband = L.fband; C.tcoff = L.tcoff; C.fb = L.fb;
    C.mbar0 = mbar0;
    C.ring0 = (unsigned)__cvta_generic_to_shared(ssm + L.ring);
    C.T = tri + (size_t)p * tstride;
    C.tpar = (unsigned)(((size_t)C.T >> 3) & 1);  // parity (in doubles) of the table's base address
    C.source = G.source;
    C.dbg = dbg;
    C.U2f[0] = P.U2f[0]; C.U2f[1] = P.U2f[1]; C.U2f[2] = P.U2f[2];
  }
  __syncthreads();

  unsigned ring_phase = 0;  // bit i = parity of the next completion of ring block i (tracked identically by every thread)
  for (int s0 = 0; s0 < S && !dead; s0 += 32) {
    const int cnt = min(32, S - s0);
    const int n_waves = cnt - 1 + N, n_blocks = (n_waves + B - 1) / B;
    // ---- prologue of the band ----
    if (tid == 0) {
      C.s0 = s0; C.cnt = cnt; C.n_waves = n_waves; C.n_blocks = n_blocks; C.rowoff = S - 1 - s0; C.phase0 = ring_phase;
    }
    for (int i = tid; i < 2 * B * 32; i += NTHR) ssm[L.ash + i] = 0.0;
    __syncthreads();
    if (warp == 1) {  // TMA producer: first helper warp, one lane per column
      for (int kb = 0; kb < 3 && kb < n_blocks; ++kb) split_issue_block<B>(C, kb, tid & 31);
    }
    if (warp >= 1 && warp <= NHW) {
      double hs0[(B + NHW - 1) / NHW];
      split_help_load<B, NHW>(C, 0, warp - 1, tid & 31, hs0);
      split_help_block<B, NHW>(ssm, C, 0, warp - 1, tid & 31, hs0);
    }
    __syncthreads();
    if (warp == 0)
      split_critical<B, NTHR>(ssm);
    else if (warp <= NHW)
      split_helper<B, NHW, NTHR>(ssm);
    else
      split_bulk<B, NHW, NTHR>(ssm);
    // ring parities after this band: ring block i was used ceil((n_blocks - i) / RB) times
#pragma unroll
    for (int i = 0; i < NUSI_SPLIT_RB; ++i)
      if (i < n_blocks && (((n_blocks - 1 - i) / NUSI_SPLIT_RB + 1) & 1)) ring_phase ^= 1u << i;
    __syncthreads();
  }

  // bin-width divide and flavour projection (nuSIprop.hpp:329-336) into this point's output slot
  const int q = out_index ? out_index[p] : p;
  for (int b = tid; b < N; b += NTHR) {
    const double f[3] = {fband[b], fband[NP + b], fband[2 * NP + b]};
    store_bin(G, flux_mass, flux_fla, q, b, f, sdE[b], dead);
  }
  if (status_out && tid == 0) status_out[q] = status_in[p];
}

// 7 / 8 bulk warps (two CTAs per SM) up to N_bins_E = 224 / 256; 12 and 18 bulk warps (one CTA per SM: 14 / 20 warps x 96 registers) up to 384 / 576
bool integrate_split_supports(int N, int NT) { return N >= 32 && N <= 576 && NT <= 704; }

template <int B, int NBW, int NHW>
static cudaError_t launch_split_t(const IntegArgs &A, cudaStream_t s, int dbg = 0) {
  const DeviceGrid &G = A.G;
  const SplitLayout L = split_layout(G.N, G.NT, G.Nz - 1, B);
  const size_t sh = (size_t)L.total * sizeof(double);
  cudaError_t e = cudaFuncSetAttribute(integrate_split_kernel<B, NBW, NHW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
  if (e != cudaSuccess) return e;
  integrate_split_kernel<B, NBW, NHW><<<A.cnt, (1 + NHW + NBW) * 32, sh, s>>>(G, A.params, A.rows, A.tri, A.tstride, A.flux_mass, A.flux_fla,
                                                                               A.out_index, A.status_in, A.status_out, dbg);
  return cudaGetLastError();
}

// variant: tuning experiments (opts.fine_variant 10..); 0 = the product configuration
cudaError_t launch_integrate_split(const IntegArgs &A, int variant, cudaStream_t s) {
  // measured against the diagonal-owned kernel on B200 (tools/gpu_wide_check.py), 4096 points: N_E = 300: 9.8 vs 16.6 ms,
  // 400: 13.8 vs 22.8, 500: 24.9 vs 60.5, 576: 28.2 vs 161.6; faster at every batch size from 8 points on
  if (A.G.N > 384) return launch_split_t<4, 18, 1>(A, s);
  if (A.G.N > 256) return launch_split_t<4, 12, 1>(A, s);
  if (A.G.N > 224) return launch_split_t<4, 8, 1>(A, s);  // 10 warps: still two CTAs per SM
  if (A.G.N <= 128 && variant == 0) return launch_split_t<4, 4, 1>(A, s);  // 6 warps, three CTAs per SM: N_E = 100: 1.71 ms per 8192 points (diagonal-owned kernel 2.89)
  switch (variant) {
    case 11: return launch_split_t<4, 7, 2>(A, s);  // two helper warps: 4.02 ms per 8192 points against 3.57 (the tenth warp costs more than it prepares)
    case 13: return launch_split_t<2, 7, 1>(A, s);
    default: return launch_split_t<4, 7, 1>(A, s, variant >= 20 ? variant - 20 : 0);  // 20 + bits: timing experiments (1 no bulk, 2 no critical)
  }
}

}  // namespace nusi
