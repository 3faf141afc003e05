// Status strings, error bookkeeping and device queries of the C ABI.
#include <cstdio>
#include <cstring>
#include "common.cuh"

namespace vslam {
static thread_local char g_err[512] = "";
void set_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
}
void snprintf_last_error(const char* fmt, int a, int b) { snprintf(g_err, sizeof(g_err), fmt, a, b); }
int sm_count() {
  static int cached = 0;
  if (cached > 0) return cached;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  cached = n;
  return n;
}
}  // namespace vslam

extern "C" const char* vslam_status_string(int s) {
  switch (s) {
    case VSLAM_OK: return "ok";
    case VSLAM_ERR_BAD_ARG: return "bad argument";
    case VSLAM_ERR_ALIGNMENT: return "buffer not aligned";
    case VSLAM_ERR_CUDA: return "CUDA error";
    case VSLAM_ERR_NUMERIC: return "numerical breakdown";
    case VSLAM_ERR_NO_DEVICE: return "no CUDA device";
    case VSLAM_ERR_TOO_LARGE: return "size exceeds the static scratch capacity of this entry point";
    default: return "unknown status";
  }
}
extern "C" const char* vslam_last_cuda_error(void) { return vslam::g_err; }
extern "C" int vslam_abi_version(void) { return 1; }
extern "C" int vslam_device_sm_count(void) { return vslam::sm_count(); }
