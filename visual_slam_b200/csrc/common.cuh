// Shared helpers for the sm_100a kernels behind include/vslam_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/vslam_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "vslam_b200 kernels are written for sm_100a only"
#endif

namespace vslam {

// Records the last CUDA failure so the Python side can print it.
void set_cuda_error(cudaError_t e, const char* where);

inline int check_launch(const char* where) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_cuda_error(e, where); return VSLAM_ERR_CUDA; }
  return VSLAM_OK;
}

#define VSLAM_CUDA_TRY(expr)                                      \
  do {                                                            \
    cudaError_t _e = (expr);                                      \
    if (_e != cudaSuccess) {                                      \
      ::vslam::set_cuda_error(_e, #expr);                         \
      return VSLAM_ERR_CUDA;                                      \
    }                                                             \
  } while (0)

int sm_count();  // cached cudaDevAttrMultiProcessorCount of the current device
// Formats a message into the string vslam_last_cuda_error() returns (argument errors that deserve an explanation).
void snprintf_last_error(const char* fmt, int a, int b);

// ---- mbarrier + TMA bulk copy (cp.async.bulk, 1-D) -------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D bulk async copy global -> shared, completion counted on `bar` (SASS: UBLKCP).
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// ---- warp reductions -------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Several sums at once, stage by stage: shuffles keep their program order, so back-to-back warp_sum() calls run one
// after the other (measured: 217 cycles each); interleaved, the shuffles of a stage are in flight together.  Each
// value sees the same additions in the same order as in warp_sum().
__device__ __forceinline__ void warp_sum3(double& a, double& b, double& c) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ta = __shfl_xor_sync(0xffffffffu, a, o), tb = __shfl_xor_sync(0xffffffffu, b, o),
                 tc = __shfl_xor_sync(0xffffffffu, c, o);
    a += ta; b += tb; c += tc;
  }
}
__device__ __forceinline__ void warp_sum4(double& a, double& b, double& c, double& d) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ta = __shfl_xor_sync(0xffffffffu, a, o), tb = __shfl_xor_sync(0xffffffffu, b, o),
                 tc = __shfl_xor_sync(0xffffffffu, c, o), td = __shfl_xor_sync(0xffffffffu, d, o);
    a += ta; b += tb; c += tc; d += td;
  }
}
__device__ __forceinline__ int warp_sum_i(int v) { return __reduce_add_sync(0xffffffffu, v); }

}  // namespace vslam
