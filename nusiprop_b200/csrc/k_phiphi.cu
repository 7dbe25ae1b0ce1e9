// k_phiphi.cu -- generator of the phi-phi cross-section tables (reference xsec/funcs.c + tables_phiphi.py + text_to_binary.cpp)
// (sm_100a, FP64 pipe; compiled -fmad=false: contraction only where fma() is written)
#include "nusi_launch.h"
#include "nusi_phiphi_gen.h"

namespace nusi {

// phi-phi table generator: one thread per node, log10(delta) index fastest so that the
// threads of a warp share (s, n) and take the same panels.
//   which 0: alphatilde node (|tbar_plus|_i0, l10d_i1)            out[i0*n1 + i1]
//   which 1: alpha node      (sbar_i0, n_i1, l10d_i2)             out[(i0*n1 + i1)*n2 + i2]
__global__ void __launch_bounds__(128)
phiphi_table_kernel(int which, long long first, long long count, int n1, int n2, const double *__restrict__ x0,
                    const double *__restrict__ x1, const double *__restrict__ x2, float *__restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count) return;
  const long long node = first + idx;
  double v;
  if (which == 0) {
    const long long i0 = node / n1;
    const int i1 = (int)(node - i0 * n1);
    v = pp_alphatilde_node(x0[i0], x1[i1]);
  } else {
    const long long i01 = node / n2;
    const int i2 = (int)(node - i01 * n2);
    const long long i0 = i01 / n1;
    const int i1 = (int)(i01 - i0 * n1);
    v = pp_alpha_node(x0[i0], x1[i1], x2[i2]);
  }
  out[idx] = (float)v;
}

cudaError_t launch_phiphi_table(int which, long long first, long long count, int n1, int n2, const double *x0, const double *x1,
                                const double *x2, float *out) {
  phiphi_table_kernel<<<(unsigned)((count + 127) / 128), 128>>>(which, first, count, n1, n2, x0, x1, x2, out);
  return cudaGetLastError();
}

}  // namespace nusi
