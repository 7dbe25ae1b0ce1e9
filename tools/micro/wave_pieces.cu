// Cycle cost of the building blocks of one integrator wave, in isolation (one CTA of 256 threads per SM):
//   nvcc -arch=sm_100a -O3 -o wave_pieces wave_pieces.cu && ./wave_pieces
//   A  rank-1 update of one diagonal: 1 LDS + 16 LDS.128 (broadcast) + 32 DFMA, by ACTIVE warps of the CTA
//   B  the same followed by 16 STS.128 (publish)
//   C  __syncthreads() round trip with all 8 warps arriving together
//   D  dependent chain shaped like the finalise step (products, compares, selects)
//   E  independent DFMAs from one warp (issue rate)
#include <cstdio>
#include <cuda_runtime.h>

__global__ void piece(int which, int active_warps, int iters, double *out, long long *cyc) {
  __shared__ __align__(16) double W[64];
  __shared__ __align__(16) double A[64];
  __shared__ double T[512];
  const int tid = threadIdx.x;
  if (tid < 64) W[tid] = 1.0 + tid * 1e-3;
  for (int i = tid; i < 512; i += blockDim.x) T[i] = 1e-3 * i;
  __syncthreads();
  double acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0;
  double v = 1.0 + tid * 1e-6, u = 0.5;
  const bool act = (tid >> 5) < active_warps;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (which == 0 || which == 1) {
      if (act) {
        const double tv = T[(tid + it) & 511];
#pragma unroll
        for (int sl = 0; sl < 32; sl += 2) {
          const double2 w2 = *reinterpret_cast<const double2 *>(W + ((it & 1) * 32) + sl);
          acc[sl] = fma(tv, w2.x, acc[sl]);
          acc[sl + 1] = fma(tv, w2.y, acc[sl + 1]);
        }
        if (which == 1 && tid == (it & 255)) {
#pragma unroll
          for (int sl = 0; sl < 32; sl += 2) *reinterpret_cast<double2 *>(A + sl) = make_double2(acc[sl], acc[sl + 1]);
        }
      }
      __syncthreads();
    } else if (which == 2) {
      __syncthreads();
    } else if (which == 3) {
      if (act) {
        // finalise-like chain
        const double a = A[tid & 31], cz = W[1], nd = W[2], lum = W[3], z = W[4], rz = W[5];
        const double base = cz * (nd * a);
        double x0 = v + (lum + base * 0.3), x1 = u + (lum + base * 0.5), x2 = v + (lum + base * 0.2);
        double q0 = x0 * rz, q1 = x1 * rz, q2 = x2 * rz;
        q0 = fma(fma(-z, q0, x0), rz, q0);
        q1 = fma(fma(-z, q1, x1), rz, q1);
        q2 = fma(fma(-z, q2, x2), rz, q2);
        const double vmax = fmax(fabs(q0), fmax(fabs(q1), fabs(q2))), vmin = fmin(fabs(q0), fmin(fabs(q1), fabs(q2)));
        if (1e-30 * vmax <= 6.9e-18 * vmin) {
          v = q0;
          u = q1 + q2 * 1e-30;
        } else {
          v = q0 / (q1 + 3.0);
          u = q2;
        }
        const double dot = 0.3 * v + 0.5 * u + 0.2 * v;
        const double r = dot * rz;
        W[32 + (tid & 31)] = fma(fma(-z, r, dot), rz, r) * 1e-3 + 1.0;
      }
      __syncthreads();
    } else if (which == 4) {
      if (act) {
#pragma unroll
        for (int sl = 0; sl < 32; ++sl) acc[sl] = fma(v, u, acc[sl]);
      }
    }
  }
  long long t1 = clock64();
  double s = v + u;
#pragma unroll
  for (int i = 0; i < 32; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + tid] = s;
  if (tid == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  double *d;
  long long *c, h;
  cudaMalloc(&d, 1 << 22);
  cudaMalloc(&c, 1 << 12);
  const char *names[] = {"A update (17 LDS + 32 DFMA) + barrier", "B update + publish (16 STS.128) + barrier", "C barrier only",
                         "D finalise-like chain + barrier", "E 32 independent DFMA"};
  const int iters = 2000;
  for (int ctas = 1; ctas <= 2; ++ctas)
    for (int which = 0; which < 5; ++which)
      for (int aw = 1; aw <= 8; aw *= 2) {
        if (which == 2 && aw > 1) continue;
        for (int rep = 0; rep < 2; ++rep) {
          piece<<<148 * ctas, 256>>>(which, aw, iters, d, c);
          cudaDeviceSynchronize();
        }
        cudaMemcpy(&h, c, sizeof h, cudaMemcpyDeviceToHost);
        printf("CTAs/SM %d  %-45s active warps %d : %.0f cycles per iteration\n", ctas, names[which], aw, (double)h / iters);
      }
  return 0;
}
