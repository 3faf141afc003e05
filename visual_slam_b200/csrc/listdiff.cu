// Unmatched-feature selection: the indices of the key points of the previous key
// frame that are not yet map points, reference src/v2/helper_functions.py:316-326 (`GetListDiff`,
// called at src/v2/main.py:239).  The reference returns `[i for i, x in enumerate(kp1) if x not in
// kp2]`; with numpy rows `x in kp2` is `(kp2 == x).any()`, i.e. a key point is dropped as soon as ANY
// row of kp2 shares EITHER coordinate with it (the row-wise loop above that line is dead code).  Mode 0
// reproduces that; mode 1 is the row-wise test the dead loop spells out (both coordinates equal).
// O(n1 * n2) exact comparisons, kp2 staged through shared memory; ordered compaction by one CTA.
#include "common.cuh"

namespace vslam {

constexpr int kDiffThreads = 256;
constexpr int kDiffTile = 1024;

__global__ void __launch_bounds__(kDiffThreads)
list_diff_flags(const double2* __restrict__ kp1, int n1, const double2* __restrict__ kp2, int n2, int mode,
                uint8_t* __restrict__ keep) {
  __shared__ double2 tile[kDiffTile];
  const int i = blockIdx.x * kDiffThreads + threadIdx.x;
  const double2 x = i < n1 ? kp1[i] : make_double2(0.0, 0.0);
  bool found = false;
  for (int t0 = 0; t0 < n2; t0 += kDiffTile) {
    const int m = min(kDiffTile, n2 - t0);
    __syncthreads();
    for (int j = threadIdx.x; j < m; j += kDiffThreads) tile[j] = kp2[t0 + j];
    __syncthreads();
    if (mode == 0) {
      for (int j = 0; j < m; ++j) found |= (tile[j].x == x.x) | (tile[j].y == x.y);
    } else {
      for (int j = 0; j < m; ++j) found |= (tile[j].x == x.x) & (tile[j].y == x.y);
    }
  }
  if (i < n1) keep[i] = found ? 0 : 1;
}

// Ascending indices of the set flags; one CTA, each thread owns a contiguous slice.
__global__ void __launch_bounds__(1024)
compact_flags(const uint8_t* __restrict__ keep, int n, int32_t* __restrict__ idx, int32_t* __restrict__ n_keep) {
  __shared__ int warp_tot[32];
  const int per = (n + 1023) / 1024;
  const int lo = min(n, (int)threadIdx.x * per), hi = min(n, lo + per);
  int cnt = 0;
  for (int i = lo; i < hi; ++i) cnt += keep[i];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int incl = cnt;
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) warp_tot[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    int w = warp_tot[lane];
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += v;
    }
    warp_tot[lane] = w;
  }
  __syncthreads();
  int pos = incl - cnt + (wid > 0 ? warp_tot[wid - 1] : 0);
  for (int i = lo; i < hi; ++i)
    if (keep[i]) idx[pos++] = i;
  if (threadIdx.x == 1023) *n_keep = warp_tot[31];
}

}  // namespace vslam

extern "C" int vslam_list_diff(const double* kp1, int n1, const double* kp2, int n2, int mode, uint8_t* keep,
                               int32_t* keep_idx, int32_t* n_keep, void* stream) {
  if (n1 < 0 || n2 < 0 || (mode != 0 && mode != 1) || !n_keep) return VSLAM_ERR_BAD_ARG;
  if ((n1 > 0 && (!kp1 || !keep || !keep_idx)) || (n2 > 0 && !kp2)) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (n1 == 0) {
    VSLAM_CUDA_TRY(cudaMemsetAsync(n_keep, 0, sizeof(int32_t), st));
    return VSLAM_OK;
  }
  vslam::list_diff_flags<<<(n1 + vslam::kDiffThreads - 1) / vslam::kDiffThreads, vslam::kDiffThreads, 0, st>>>(
      reinterpret_cast<const double2*>(kp1), n1, reinterpret_cast<const double2*>(kp2), n2, mode, keep);
  vslam::compact_flags<<<1, 1024, 0, st>>>(keep, n1, keep_idx, n_keep);
  return vslam::check_launch("vslam_list_diff");
}
