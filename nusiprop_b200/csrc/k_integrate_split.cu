// k_integrate_split.cu -- the implicit z-stepping integrator (reference nuSIprop.hpp:255-336), role-split wavefront.
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

namespace nusi {

// ---- mbarrier / TMA bulk copy (PTX ISA 8.x, sm_90+) ------------------------------------------------------------
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
// all NTHR threads of the CTA, from any of the role loops (named barrier 1; __syncthreads() must not be used in divergent code)
template <int NTHR>
__device__ __forceinline__ void block_sync() {
  asm volatile("bar.sync 1, %0;" ::"n"(NTHR) : "memory");
}
// global -> shared bulk copy; dst, src 16-byte aligned, bytes a multiple of 16; completes on the mbarrier
__device__ __forceinline__ void tma_bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
               "r"(bar)
               : "memory");
}

#define NUSI_SPLIT_PF 10  // slot fields: lum, offdiag, Gam, aT, Z0, rZ0, Z1, rZ1, Z2, rZ2
#define NUSI_SPLIT_RB 5   // ring depth in blocks of B columns (block kb-1 .. kb+3)

struct SplitLayout {  // offsets in doubles into dynamic shared memory
  int NTP, ring, wsh, ash, pre, sG, sA, sdE, sRdE, sCz, sNd, fband, mbar, total;
};
__host__ __device__ inline SplitLayout split_layout(int N, int NT, int S, int B) {
  SplitLayout L;
  L.NTP = (NT + 3) & ~1;  // a column, possibly shifted by one double for 16-byte alignment of its global source
  int o = 0;
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
  L.mbar = o; o += NUSI_SPLIT_RB;
  L.total = o;
  return L;
}

// B waves per block; NBW bulk warps (diagonals: N + 31 <= 32 NBW); NHW helper warps
template <int B, int NBW, int NHW>
__global__ void __launch_bounds__((1 + NHW + NBW) * 32, 2)
integrate_split_kernel(DeviceGrid G, const PointParams *__restrict__ params, const double *__restrict__ rows,
                       const double *__restrict__ tri, size_t tstride, double *__restrict__ flux_mass,
                       double *__restrict__ flux_fla, const int *__restrict__ out_index, const int *__restrict__ status_in,
                       int *__restrict__ status_out) {
  constexpr int D = 2 * B - 1;
  constexpr int NTHR = (1 + NHW + NBW) * 32;
  extern __shared__ __align__(16) double ssm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int p = blockIdx.x;
  const int N = G.N, NT = G.NT, S = G.Nz - 1;
  const PointParams P = params[p];
  const bool dead = !(P.mphi > 0);
  const SplitLayout L = split_layout(N, NT, S, B);
  double *Tring = ssm + L.ring, *Wsh = ssm + L.wsh, *Ash = ssm + L.ash, *Pre = ssm + L.pre;
  double *sG = ssm + L.sG, *sA = ssm + L.sA, *sdE = ssm + L.sdE, *sRdE = ssm + L.sRdE, *sCz = ssm + L.sCz, *sNd = ssm + L.sNd;
  double *fband = ssm + L.fband;
  const int NP = (N + 1) & ~1;  // stride of fband rows
  const unsigned mbar0 = (unsigned)__cvta_generic_to_shared(ssm + L.mbar);
  const unsigned ring0 = (unsigned)__cvta_generic_to_shared(Tring);
  const double *T = tri + (size_t)p * tstride;
  const double U2f[3] = {P.U2f[0], P.U2f[1], P.U2f[2]};

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
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < NUSI_SPLIT_RB; ++i) mbar_init(mbar0 + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  unsigned ring_phase = 0;  // bit i = parity of the next completion of ring block i (tracked identically by every thread)
  for (int s0 = 0; s0 < S && !dead; s0 += 32) {
    const int cnt = min(32, S - s0);
    const int n_waves = cnt - 1 + N, n_blocks = (n_waves + B - 1) / B;
    const int rowoff = S - 1 - s0;  // table row of diagonal d in any column: d + rowoff
    // column of wave w_rel: c = NT-1-(s0+w_rel); rows 0..c-1 at packed offset c(c-1)/2.  Issue block kb's B columns.
    auto issue_block = [&](int kb) {
      const int rb = kb % NUSI_SPLIT_RB;
      const unsigned bar = mbar0 + 8u * rb;
      unsigned total = 0;
#pragma unroll
      for (int r = 0; r < B; ++r) {
        const int w_rel = kb * B + r, c = NT - 1 - (s0 + w_rel);
        if (w_rel < n_waves && c > 0) {
          const double *src = T + (size_t)c * (c - 1) / 2;
          const unsigned mis = (unsigned)(((size_t)src >> 3) & 1);
          total += ((mis + (unsigned)c + 1u) & ~1u) * 8u;
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic-proxy reads of the slots are ordered before the async writes
      mbar_expect_tx(bar, total);
#pragma unroll
      for (int r = 0; r < B; ++r) {
        const int w_rel = kb * B + r, c = NT - 1 - (s0 + w_rel);
        if (w_rel < n_waves && c > 0) {
          const double *src = T + (size_t)c * (c - 1) / 2;
          const unsigned mis = (unsigned)(((size_t)src >> 3) & 1);
          tma_bulk_g2s(ring0 + (unsigned)((rb * B + r) * L.NTP) * 8u, src - mis, ((mis + (unsigned)c + 1u) & ~1u) * 8u, bar);
        }
      }
    };
    // row a of the column of wave w_rel sits at Tring[slot * NTP + mis(w_rel) + a]
    auto col_ptr = [&](int w_rel) -> const double * {
      const int c = NT - 1 - (s0 + w_rel);
      const double *src = T + (size_t)c * (c - 1) / 2;
      const int kb = w_rel / B;
      return Tring + ((kb % NUSI_SPLIT_RB) * B + (w_rel - kb * B)) * L.NTP + (int)(((size_t)src >> 3) & 1);
    };

    // ---- helper: slots of block kb into Pre[kb & 1] ----
    auto help_block = [&](int kb, int hw) {
      const int sl = lane, zi = S - 1 - s0 - sl;
      for (int r = hw; r < B; r += NHW) {
        const int w_rel = kb * B + r, k = w_rel - sl, b = N - 1 - k;
        if (sl < cnt && k >= 0 && b >= 0 && w_rel < n_waves) {
          const int c = NT - 1 - (s0 + w_rel);
          const double cz = sCz[zi], ndz = sNd[zi], dEb = sdE[b], rdE = sRdE[b];
          const double Gam = ndz * sG[c], aT = ndz * sA[c], lum = cz * __ldg(G.source + (size_t)zi * N + b);
          double Z[3], rZ[3];
          bool ok = true;
#pragma unroll
          for (int kk = 0; kk < 3; ++kk) {
            const double x = cz * (Gam * U2f[kk] - aT * sqr(U2f[kk]));
            const double q = quot_fast(x, dEb, rdE);
            ok = ok && (fabs(q) <= DBL_MAX);
            Z[kk] = 1.0 + q;
          }
          if (!ok) {
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) Z[kk] = 1.0 + cz * (Gam * U2f[kk] - aT * sqr(U2f[kk])) / dEb;
          }
          ok = true;
#pragma unroll
          for (int kk = 0; kk < 3; ++kk) {
            rZ[kk] = recip_newton(Z[kk]);
            ok = ok && (fabs(Z[kk]) < 1e30) && (fabs(Z[kk]) > 1e-30);
          }
          if (!ok) {
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) rZ[kk] = 1.0 / Z[kk];
          }
          const double offdiag = fabs(aT) * rdE * fmax(fabs(rZ[0]), fmax(fabs(rZ[1]), fabs(rZ[2]))) * 1.00001;
          double *o = Pre + (((kb & 1) * B + r) * NUSI_SPLIT_PF) * 32 + sl;
          o[0 * 32] = lum; o[1 * 32] = offdiag; o[2 * 32] = Gam; o[3 * 32] = aT;
          o[4 * 32] = Z[0]; o[5 * 32] = rZ[0]; o[6 * 32] = Z[1]; o[7 * 32] = rZ[1]; o[8 * 32] = Z[2]; o[9 * 32] = rZ[2];
        }
      }
    };

    // ---- prologue of the band ----
    for (int i = tid; i < 2 * B * 32; i += NTHR) Ash[i] = 0.0;
    if (tid == 32) {  // TMA producer: first helper warp, lane 0
      for (int kb = 0; kb < 3 && kb < n_blocks; ++kb) issue_block(kb);
    }
    if (warp >= 1 && warp <= NHW) help_block(0, warp - 1);
    __syncthreads();

    const unsigned phase0 = ring_phase;  // parities at the start of the band; ring block i has been completed (kb_done_i) times since
    auto parity_of = [&](int kb) -> unsigned {  // parity on which block kb of this band completes
      const int rb = kb % NUSI_SPLIT_RB;
      return ((phase0 >> rb) & 1u) ^ (unsigned)((kb / NUSI_SPLIT_RB) & 1);
    };
    if (warp == 0) {
      // ================= critical warp =================
      const int sl = lane;
      double ip[D];    // running sums of the next D cells of this lane (sources at most D waves back)
      double fout[3];  // flux of the cell finalised in the previous wave
#pragma unroll
      for (int j = 0; j < D; ++j) ip[j] = 0.0;
      fout[0] = fout[1] = fout[2] = 0.0;
      const double czl = (sl < cnt) ? sCz[S - 1 - s0 - sl] : 0.0, ndl = (sl < cnt) ? sNd[S - 1 - s0 - sl] : 0.0;
      for (int kb = 0; kb < n_blocks; ++kb) {
        mbar_wait(mbar0 + 8u * (kb % NUSI_SPLIT_RB), parity_of(kb));
#pragma unroll
        for (int r = 0; r < B; ++r) {
          const int w_rel = kb * B + r;
          if (w_rel < n_waves) {
            const int k = w_rel - sl, b = N - 1 - k;
            const bool act = (sl < cnt) && (k >= 0) && (b >= 0);
            const int bb = act ? b : 0;
            // flux of this bin after the previous z-step: from the lane below (finalised one wave ago), or the
            // previous band's result for the first z-step of the band
            double f[3];
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) {
              const double up = __shfl_up_sync(0xffffffffu, fout[kk], 1);
              f[kk] = (sl == 0) ? fband[kk * NP + bb] : up;
            }
            if (k == 0) {  // this lane starts now: no regeneration into the top bin, nothing carried over
#pragma unroll
              for (int j = 0; j < D; ++j) ip[j] = 0.0;
            }
            double W = 0.0;
            if (act) {
              const double *slot = Pre + (((kb & 1) * B + r) * NUSI_SPLIT_PF) * 32 + sl;
              const double acc_c = Ash[((kb & 1) * B + r) * 32 + sl] + ip[0];
              const double dEb = sdE[bb], rdE = sRdE[bb];
              const double lum = slot[0], offdiag = slot[32];
              const double Z0 = slot[4 * 32], r0 = slot[5 * 32], Z1 = slot[6 * 32], r1 = slot[7 * 32], Z2 = slot[8 * 32], r2 = slot[9 * 32];
              double v[3], src[3], x[3];
              const double base = czl * (ndl * acc_c);
#pragma unroll
              for (int kk = 0; kk < 3; ++kk) {
                src[kk] = lum + base * U2f[kk];
                x[kk] = f[kk] + src[kk];
              }
              v[0] = quot_fast(x[0], Z0, r0);
              v[1] = quot_fast(x[1], Z1, r1);
              v[2] = quot_fast(x[2], Z2, r2);
              const int h0 = __double2hiint(v[0]) & 0x7fffffff, h1 = __double2hiint(v[1]) & 0x7fffffff, h2 = __double2hiint(v[2]) & 0x7fffffff;
              const int hmax = max(h0, max(h1, h2)), hmin = min(h0, min(h1, h2));
              bool identity = false;
              if (hmax >= 0x7ff00000) {
                slow_quotients3(x, Z0, Z1, Z2, v);
              } else {
                identity = offdiag * __hiloint2double(hmax + 1, 0) <= 6.9388939039072284e-18 /* 2^-57 */ * __hiloint2double(hmin, 0);
              }
              if (identity) {
                f[0] = v[0];
                f[1] = v[1];
                f[2] = v[2];
              } else {
                implicit_bin(U2f, czl, slot[2 * 32], slot[3 * 32], dEb, src, f);
              }
              const double dot = U2f[0] * f[0] + U2f[1] * f[1] + U2f[2] * f[2];
              W = quot_fast(dot, dEb, rdE);
              if ((__double2hiint(W) & 0x7fffffff) >= 0x7ff00000) W = slow_quotient(dot, dEb);
              fout[0] = f[0];
              fout[1] = f[1];
              fout[2] = f[2];
              if (sl == cnt - 1) {  // the band's last z-step: this bin's flux after the band
                fband[bb] = f[0];
                fband[NP + bb] = f[1];
                fband[2 * NP + bb] = f[2];
              }
            }
            Wsh[((kb & 1) * B + r) * 32 + sl] = W;
            // right-looking: this lane's W into its own next D cells (target bin b-j: table row c-j of column c)
            const int c = NT - 1 - (s0 + w_rel);
            const double *tc = col_ptr(w_rel) + c;
            if (act) {
#pragma unroll
              for (int j = 1; j < D; ++j) ip[j - 1] = fma((c - j >= 0) ? tc[-j] : 0.0, W, ip[j]);
              ip[D - 1] = ((c - D >= 0) ? tc[-D] : 0.0) * W;
            } else {
#pragma unroll
              for (int j = 1; j < D; ++j) ip[j - 1] = ip[j];
              ip[D - 1] = 0.0;
            }
          }
        }
        block_sync<NTHR>();
      }
    } else if (warp <= NHW) {
      // ================= helper warps: slots of block kb+1, TMA for block kb+3 =================
      for (int kb = 0; kb < n_blocks; ++kb) {
        if (tid == 32 && kb + 3 < n_blocks) issue_block(kb + 3);
        if (kb + 1 < n_blocks) help_block(kb + 1, warp - 1);
        block_sync<NTHR>();
      }
    } else {
      // ================= bulk warps: sources of block kb-1, publish the diagonals of block kb+1 =================
      double acc[32];  // sums of diagonal d for the 32 z-steps of the band
#pragma unroll
      for (int j = 0; j < 32; ++j) acc[j] = 0.0;
      const int d = tid - (1 + NHW) * 32 - 31;  // this thread's diagonal: cells (sl, b = d + sl)
      const int wd = N - 1 - d;                 // the wave in which it is finalised
      for (int kb = 0; kb < n_blocks; ++kb) {
        if (kb >= 1) {
          const int kbs = kb - 1;
          mbar_wait(mbar0 + 8u * (kbs % NUSI_SPLIT_RB), parity_of(kbs));
#pragma unroll
          for (int r = 0; r < B; ++r) {
            const int w_rel = kbs * B + r;
            if (w_rel < n_waves && wd - w_rel > D && d + rowoff >= 0) {  // distance of this diagonal's cells from the source wave
              const double tv = col_ptr(w_rel)[d + rowoff];
              const double *Wp = Wsh + ((kbs & 1) * B + r) * 32;
              const int hi = min(w_rel, cnt - 1);  // z-steps that have started
              if (hi == 31) {
#pragma unroll
                for (int sl = 0; sl < 32; sl += 2) {
                  const double2 w2 = *reinterpret_cast<const double2 *>(Wp + sl);
                  acc[sl] = fma(tv, w2.x, acc[sl]);
                  acc[sl + 1] = fma(tv, w2.y, acc[sl + 1]);
                }
              } else {
#pragma unroll
                for (int sl = 0; sl < 32; ++sl)
                  if (sl <= hi) acc[sl] = fma(tv, Wp[sl], acc[sl]);
              }
            }
          }
        }
        if (wd >= (kb + 1) * B && wd < (kb + 2) * B && wd < n_waves) {
          double *o = Ash + (((kb + 1) & 1) * B + (wd - (kb + 1) * B)) * 32;
#pragma unroll
          for (int sl = 0; sl < 32; sl += 2) *reinterpret_cast<double2 *>(o + sl) = make_double2(acc[sl], acc[sl + 1]);
        }
        block_sync<NTHR>();
      }
    }
    // ring parities after this band: ring block i was used ceil((n_blocks - i) / RB) times
#pragma unroll
    for (int i = 0; i < NUSI_SPLIT_RB; ++i)
      if (i < n_blocks && (((n_blocks - 1 - i) / NUSI_SPLIT_RB + 1) & 1)) ring_phase ^= 1u << i;
    block_sync<NTHR>();
  }

  // bin-width divide and flavour projection (nuSIprop.hpp:329-336) into this point's output slot
  const int q = out_index ? out_index[p] : p;
  for (int b = tid; b < N; b += NTHR) {
    const double f[3] = {fband[b], fband[NP + b], fband[2 * NP + b]};
    store_bin(G, flux_mass, flux_fla, q, b, f, sdE[b], dead);
  }
  if (status_out && tid == 0) status_out[q] = status_in[p];
}

bool integrate_split_supports(int N, int NT) { return N + 31 <= 256 && NT <= 512; }

cudaError_t launch_integrate_split(const IntegArgs &A, cudaStream_t s) {
  const DeviceGrid &G = A.G;
  constexpr int B = 4, NBW = 8, NHW = 2;
  const SplitLayout L = split_layout(G.N, G.NT, G.Nz - 1, B);
  const size_t sh = (size_t)L.total * sizeof(double);
  cudaError_t e = cudaFuncSetAttribute(integrate_split_kernel<B, NBW, NHW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
  if (e != cudaSuccess) return e;
  integrate_split_kernel<B, NBW, NHW><<<A.cnt, (1 + NHW + NBW) * 32, sh, s>>>(G, A.params, A.rows, A.tri, A.tstride, A.flux_mass, A.flux_fla,
                                                                               A.out_index, A.status_in, A.status_out);
  return cudaGetLastError();
}

}  // namespace nusi
