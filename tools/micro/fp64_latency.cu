// Dependent-issue latency of FP64 instructions on the device (one warp per SM, one dependent chain):
//   nvcc -arch=sm_100a -O3 -o fp64_latency fp64_latency.cu && ./fp64_latency
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void chain(double *out, long long *cyc, int iters, double a, double b) {
  double v = threadIdx.x * 1e-3 + 1.0;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (OP == 0) v = fma(v, a, b);
      if (OP == 1) v = v * a;
      if (OP == 2) v = v + b;
      if (OP == 3) v = b / v + a;   // division + add
      if (OP == 4) v = __shfl_xor_sync(0xffffffff, v, 1) + b;
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = v;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
  double *d; long long *c, h;
  cudaMalloc(&d, 1 << 20); cudaMalloc(&c, 1 << 12);
  const char *names[] = {"DFMA", "DMUL", "DADD", "DDIV+DADD", "SHFL+DADD"};
  for (int warps = 1; warps <= 8; warps *= 2)
    for (int op = 0; op < 5; ++op) {
      int iters = 2000;
      for (int rep = 0; rep < 2; ++rep) {
        if (op == 0) chain<0><<<148, 32 * warps>>>(d, c, iters, 0.999999, 1e-9);
        if (op == 1) chain<1><<<148, 32 * warps>>>(d, c, iters, 0.999999, 1e-9);
        if (op == 2) chain<2><<<148, 32 * warps>>>(d, c, iters, 0.999999, 1e-9);
        if (op == 3) chain<3><<<148, 32 * warps>>>(d, c, iters, 0.999999, 1.5);
        if (op == 4) chain<4><<<148, 32 * warps>>>(d, c, iters, 0.999999, 1e-9);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(&h, c, sizeof h, cudaMemcpyDeviceToHost);
      printf("warps/CTA %d  %-10s %.1f cycles per dependent op\n", warps, names[op], (double)h / (iters * 16.0));
    }
  return 0;
}
