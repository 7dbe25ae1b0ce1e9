// nusi_cpasync.h -- asynchronous global->shared copies (LDGSTS) used by the scan builder and the integrators
#ifndef NUSI_CPASYNC_H
#define NUSI_CPASYNC_H
namespace nusi {
// 8-byte asynchronous global->shared copy (LDGSTS); valid == false writes zeros and reads nothing
__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gsrc, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int sz = valid ? 8 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async8s(unsigned smem_addr, const double *gsrc, bool valid) {  // 32-bit shared address
  const int sz = valid ? 8 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(smem_addr), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int PENDING>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(PENDING) : "memory");
}
}  // namespace nusi
#endif
