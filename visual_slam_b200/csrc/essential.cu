// Two-view geometry on device: cv2.undistortPoints (K2), essential-matrix RANSAC (five-point
// hypotheses, one warp per model Sampson scoring, cv2's adaptive selection), the epiline score, and
// cv2.recoverPose (decomposition + DLT triangulation + cheirality), batched over frame pairs.
// Replaces reference src/v2/helper_functions.py:47-70 (estimateEssential) and :164-195
// (estimateRelativePose, "Essential" branch).
#include "essential_math.cuh"

namespace vslam {

constexpr int kEssSample = 5;
constexpr int kEssMaxModels = 10;

__global__ void undistort_points_kernel(const float* __restrict__ uv, int64_t n, double fx, double fy,
                                        double cx, double cy, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float2 p = reinterpret_cast<const float2*>(uv)[i];
  // cv2: ifx = 1/fx; x = (x - cx) * ifx, in double, stored as float
  const double ifx = 1.0 / fx, ify = 1.0 / fy;
  const double x = __dmul_rn(__dsub_rn((double)p.x, cx), ifx);
  const double y = __dmul_rn(__dsub_rn((double)p.y, cy), ify);
  reinterpret_cast<float2*>(out)[i] = make_float2((float)x, (float)y);
}

// Harness sampling rule (oracle/pnp.py draw_samples with k = 5).
__device__ __forceinline__ void draw_sample5(const uint32_t* w, int n, int* idx) {
  for (int j = 0; j < kEssSample; ++j) {
    int v = (int)(((unsigned long long)w[j] * (unsigned long long)n) >> 32);
    bool again = true;
    while (again) {
      again = false;
      for (int i = 0; i < j; ++i)
        if (idx[i] == v) { v = (v + 1) % n; again = true; }
    }
    idx[j] = v;
  }
}

// One thread per (pair, sample): five-point solve -> up to 10 models.
__global__ void __launch_bounds__(64)
essential_hypotheses_kernel(const float* __restrict__ x1, const float* __restrict__ x2,
                            const int32_t* __restrict__ n_pts, int n_max, int H,
                            const uint32_t* __restrict__ words, double* __restrict__ models,
                            int32_t* __restrict__ n_models, int total) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  const int b = g / H;
  const int n = min(max(n_pts[b], 0), n_max);   // counts beyond the buffers are clamped
  int count = 0;
  double m[kEssMaxModels][9];
  if (n >= kEssSample) {
    int idx[kEssSample];
    draw_sample5(words + (size_t)g * kEssSample, n, idx);
    double a[5][2], c[5][2];
    for (int j = 0; j < kEssSample; ++j) {
      const float2 p = reinterpret_cast<const float2*>(x1)[(size_t)b * n_max + idx[j]];
      const float2 q = reinterpret_cast<const float2*>(x2)[(size_t)b * n_max + idx[j]];
      a[j][0] = p.x; a[j][1] = p.y; c[j][0] = q.x; c[j][1] = q.y;
    }
    count = five_point(a, c, m);
  }
  double* out = models + (size_t)g * kEssMaxModels * 9;
  for (int i = 0; i < kEssMaxModels; ++i)
    for (int q = 0; q < 9; ++q) out[9 * i + q] = (i < count) ? m[i][q] : 0.0;
  n_models[g] = count;
}

// Sampson error exactly as cv2 stores it (EMEstimatorCallback::computeError): double arithmetic in
// the source order, no FMA contraction, float32 result.
__device__ __forceinline__ float sampson_err_f32(const double* E, float2 p1, float2 p2) {
  const double x1 = p1.x, y1 = p1.y, x2 = p2.x, y2 = p2.y;
  const double Ex0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], x1), __dmul_rn(E[1], y1)), E[2]);
  const double Ex1 = __dadd_rn(__dadd_rn(__dmul_rn(E[3], x1), __dmul_rn(E[4], y1)), E[5]);
  const double Ex2 = __dadd_rn(__dadd_rn(__dmul_rn(E[6], x1), __dmul_rn(E[7], y1)), E[8]);
  const double Et0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], x2), __dmul_rn(E[3], y2)), E[6]);
  const double Et1 = __dadd_rn(__dadd_rn(__dmul_rn(E[1], x2), __dmul_rn(E[4], y2)), E[7]);
  const double s = __dadd_rn(__dadd_rn(__dmul_rn(x2, Ex0), __dmul_rn(y2, Ex1)), Ex2);
  const double den = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Ex0, Ex0), __dmul_rn(Ex1, Ex1)), __dmul_rn(Et0, Et0)),
                               __dmul_rn(Et1, Et1));
  return (float)__ddiv_rn(__dmul_rn(s, s), den);
}

// Scoring: grid (ceil(H*10 / 8), pairs), 8 warps; the pair's correspondences are staged once in
// shared memory, each warp owns one (sample, model) slot and strides over the points.
constexpr int kEssScoreWarps = 8;

__global__ void __launch_bounds__(kEssScoreWarps * 32)
essential_score_kernel(const float* __restrict__ x1, const float* __restrict__ x2,
                       const int32_t* __restrict__ n_pts, int n_max, int H,
                       const double* __restrict__ models, const int32_t* __restrict__ n_models,
                       float thr2, int32_t* __restrict__ counts) {
  extern __shared__ __align__(16) float2 sp2[];  // [n] x1 then [n] x2
  const int b = blockIdx.y;
  const int n = min(max(n_pts[b], 0), n_max);   // counts beyond the buffers are clamped
  float2* s1 = sp2;
  float2* s2 = sp2 + n_max;
  const float2* g1 = reinterpret_cast<const float2*>(x1) + (size_t)b * n_max;
  const float2* g2 = reinterpret_cast<const float2*>(x2) + (size_t)b * n_max;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { s1[i] = g1[i]; s2[i] = g2[i]; }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = blockIdx.x * kEssScoreWarps + warp;
  if (slot >= H * kEssMaxModels) return;
  const int h = slot / kEssMaxModels, m = slot % kEssMaxModels;
  const size_t g = (size_t)b * H + h;
  int cnt = 0;
  if (m < n_models[g]) {
    double E[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) E[i] = models[(g * kEssMaxModels + m) * 9 + i];
    for (int i = lane; i < n; i += 32) cnt += (sampson_err_f32(E, s1[i], s2[i]) <= thr2) ? 1 : 0;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
  }
  if (lane == 0) counts[g * kEssMaxModels + m] = cnt;
}

__device__ inline int ess_update_num_iters(double p, double ep, int model_points, int max_iters) {
  p = fmin(fmax(p, 0.0), 1.0);
  ep = fmin(fmax(ep, 0.0), 1.0);
  double num = fmax(1.0 - p, 2.2250738585072014e-308);
  double denom = 1.0 - pow(1.0 - ep, (double)model_points);
  if (denom < 2.2250738585072014e-308) return 0;
  num = log(num);
  denom = log(denom);
  return (denom >= 0 || -num >= max_iters * (-denom)) ? max_iters : (int)rint(num / denom);
}

// cv2.computeCorrespondEpilines(pts, 2, E) for float32 points: l = E^T [x,y,1], a^2+b^2 = 1,
// double arithmetic, float32 result; then the reference's error (helper_functions.py:60,66):
// (loc . l)^2 / (a^2 + b^2 + c^2) with loc = [x, y, 1] of the OTHER image.
__device__ __forceinline__ double epiline_err(const double* E, float2 src, float2 loc) {
  const double x = src.x, y = src.y;
  const double a = __dadd_rn(__dadd_rn(__dmul_rn(E[0], x), __dmul_rn(E[3], y)), E[6]);
  const double b = __dadd_rn(__dadd_rn(__dmul_rn(E[1], x), __dmul_rn(E[4], y)), E[7]);
  const double c = __dadd_rn(__dadd_rn(__dmul_rn(E[2], x), __dmul_rn(E[5], y)), E[8]);
  double nu = __dadd_rn(__dmul_rn(a, a), __dmul_rn(b, b));
  nu = (nu != 0.0) ? __ddiv_rn(1.0, __dsqrt_rn(nu)) : 1.0;
  const double fa = (double)(float)__dmul_rn(a, nu), fb = (double)(float)__dmul_rn(b, nu),
               fc = (double)(float)__dmul_rn(c, nu);
  const double d = (double)loc.x * fa + (double)loc.y * fb + fc;
  return d * d / (fa * fa + fb * fb + fc * fc);
}

constexpr int kSelThreads = 256;

// One CTA per pair: cv2's selection loop over the (sample, model) counts with the shrinking budget,
// then the inlier mask of the winner, its count and the epiline score of estimateEssential.
__global__ void __launch_bounds__(kSelThreads)
essential_select_kernel(const float* __restrict__ x1, const float* __restrict__ x2,
                        const int32_t* __restrict__ n_pts, int n_max, int H,
                        const double* __restrict__ models, const int32_t* __restrict__ n_models,
                        const int32_t* __restrict__ counts, float thr2, double confidence,
                        int32_t* __restrict__ best_out, double* __restrict__ E_out,
                        uint8_t* __restrict__ mask_out, int32_t* __restrict__ n_inl_out,
                        double* __restrict__ score_out) {
  extern __shared__ int32_t s_counts[];  // [H][10], -1 for absent models
  __shared__ int s_best;
  __shared__ double s_red[kSelThreads / 32];
  __shared__ int s_redi[kSelThreads / 32];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int n = min(max(n_pts[b], 0), n_max);   // counts beyond the buffers are clamped
  for (int i = tid; i < H * kEssMaxModels; i += kSelThreads) {
    const int h = i / kEssMaxModels, m = i % kEssMaxModels;
    s_counts[i] = (m < n_models[(size_t)b * H + h]) ? counts[((size_t)b * H + h) * kEssMaxModels + m] : -1;
  }
  __syncthreads();
  if (tid == 0) {
    int niters = H, best = -1, best_count = 0;
    for (int it = 0; it < niters && it < H; ++it) {
      for (int m = 0; m < kEssMaxModels; ++m) {
        const int c = s_counts[it * kEssMaxModels + m];
        if (c > max(best_count, kEssSample - 1)) {
          best = it * kEssMaxModels + m; best_count = c;
          niters = ess_update_num_iters(confidence, (double)(n - c) / (double)n, kEssSample, niters);
        }
      }
    }
    s_best = best;
  }
  __syncthreads();
  const int best = s_best;
  uint8_t* mask = mask_out + (size_t)b * n_max;
  if (best < 0) {
    for (int i = tid; i < n_max; i += kSelThreads) mask[i] = 0;
    if (tid == 0) {
      best_out[b] = -1; n_inl_out[b] = 0; score_out[b] = 0.0;
      for (int i = 0; i < 9; ++i) E_out[9 * b + i] = 0.0;
    }
    return;
  }
  double E[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) E[i] = models[((size_t)b * H * kEssMaxModels + best) * 9 + i];
  const float2* g1 = reinterpret_cast<const float2*>(x1) + (size_t)b * n_max;
  const float2* g2 = reinterpret_cast<const float2*>(x2) + (size_t)b * n_max;
  int cnt = 0;
  double sc = 0.0;
  for (int i = tid; i < n_max; i += kSelThreads) {
    bool in = false;
    if (i < n) {
      const float2 p1 = g1[i], p2 = g2[i];
      in = sampson_err_f32(E, p1, p2) <= thr2;
      if (in) {
        // both line sets use whichImage = 2 in the reference (helper_functions.py:57,64)
        sc += fmax(4.0 - epiline_err(E, p2, p1), 0.0) + fmax(4.0 - epiline_err(E, p1, p2), 0.0);
      }
    }
    mask[i] = in ? 1 : 0;
    cnt += in ? 1 : 0;
  }
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  sc = warp_sum(sc);
  if ((tid & 31) == 0) { s_redi[tid >> 5] = cnt; s_red[tid >> 5] = sc; }
  __syncthreads();
  if (tid == 0) {
    int c = 0; double s = 0.0;
    for (int w = 0; w < kSelThreads / 32; ++w) { c += s_redi[w]; s += s_red[w]; }
    best_out[b] = best; n_inl_out[b] = c; score_out[b] = s;
    for (int i = 0; i < 9; ++i) E_out[9 * b + i] = E[i];
  }
}

// The iteration budget cv2's RANSAC loop is left with after the first H hypotheses of a run that started with
// max_iters: lets a caller generate hypotheses in rounds and stop as soon as budget <= H, with the result of
// the full run (essential_select_kernel on that prefix picks the same model).  One thread per pair.
__global__ void essential_budget_kernel(const int32_t* __restrict__ n_pts, int n_max, int batch, int H, int max_iters,
                                        const int32_t* __restrict__ n_models, const int32_t* __restrict__ counts,
                                        double confidence, int32_t* __restrict__ budget) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const int n = min(max(n_pts[b], 0), n_max);
  int niters = max_iters, best_count = 0;
  for (int it = 0; it < niters && it < H; ++it) {
    const int nm = n_models[(size_t)b * H + it];
    for (int m = 0; m < nm && m < kEssMaxModels; ++m) {
      const int c = counts[((size_t)b * H + it) * kEssMaxModels + m];
      if (c > max(best_count, kEssSample - 1)) {
        best_count = c;
        niters = ess_update_num_iters(confidence, (double)(n - c) / (double)n, kEssSample, niters);
      }
    }
  }
  budget[b] = niters;
}

// cv2.recoverPose: one CTA per pair.  Pass 1 counts, for each of the four (R, t) candidates, the
// points that triangulate in front of both cameras and closer than `dist`; the candidate with the
// most survivors wins (cv2's order of preference on ties); pass 2 writes its mask (0 / 255) and the
// homogeneous triangulated points (4 x N like cv2, unit columns, sign free).
constexpr int kPoseThreads = 512;

__global__ void __launch_bounds__(kPoseThreads)
recover_pose_kernel(const double* __restrict__ E_in, const float* __restrict__ p1, const float* __restrict__ p2,
                    const int32_t* __restrict__ n_pts, int n_max, double fx, double fy, double cx, double cy,
                    double dist, double* __restrict__ R_out, double* __restrict__ t_out,
                    uint8_t* __restrict__ mask_out, double* __restrict__ tri_out,
                    int32_t* __restrict__ good_out /*[batch][5]: chosen count + the four counts*/) {
  __shared__ double s_R[2][9];
  __shared__ double s_t[3];
  __shared__ int s_cnt[kPoseThreads / 32][4];
  __shared__ int s_sel;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int n = min(max(n_pts[b], 0), n_max);   // counts beyond the buffers are clamped
  if (tid == 0) decompose_essential(E_in + 9 * b, s_R[0], s_R[1], s_t);
  __syncthreads();
  const float2* g1 = reinterpret_cast<const float2*>(p1) + (size_t)b * n_max;
  const float2* g2 = reinterpret_cast<const float2*>(p2) + (size_t)b * n_max;
  auto candidate = [&](int k, const double*& R, double* t) {
    R = s_R[k & 1];
    const double sgn = (k >= 2) ? -1.0 : 1.0;
    t[0] = sgn * s_t[0]; t[1] = sgn * s_t[1]; t[2] = sgn * s_t[2];
  };
  // pass 1: one work item per (point, candidate) -- the 4x4 DLT decomposition is a long dependent chain, so the
  // items are spread over the threads instead of giving a thread all four candidates of its points
  int cnt[4] = {0, 0, 0, 0};
  for (int w = tid; w < 4 * n; w += kPoseThreads) {
    const int i = w >> 2, k = w & 3;
    const float2 a = g1[i], c = g2[i];
    const double x1 = ((double)a.x - cx) / fx, y1 = ((double)a.y - cy) / fy;
    const double x2 = ((double)c.x - cx) / fx, y2 = ((double)c.y - cy) / fy;
    const double* R; double t[3], Q[4];
    candidate(k, R, t);
    triangulate_dlt(R, t, x1, y1, x2, y2, Q);
    const int ok = cheirality_ok(R, t, Q, dist) ? 1 : 0;
    cnt[0] += (k == 0) ? ok : 0; cnt[1] += (k == 1) ? ok : 0; cnt[2] += (k == 2) ? ok : 0; cnt[3] += (k == 3) ? ok : 0;
  }
  for (int k = 0; k < 4; ++k) cnt[k] = __reduce_add_sync(0xffffffffu, cnt[k]);
  if ((tid & 31) == 0)
    for (int k = 0; k < 4; ++k) s_cnt[tid >> 5][k] = cnt[k];
  __syncthreads();
  if (tid == 0) {
    int g[4] = {0, 0, 0, 0};
    for (int w = 0; w < kPoseThreads / 32; ++w)
      for (int k = 0; k < 4; ++k) g[k] += s_cnt[w][k];
    int sel = 3;
    if (g[0] >= g[1] && g[0] >= g[2] && g[0] >= g[3]) sel = 0;
    else if (g[1] >= g[0] && g[1] >= g[2] && g[1] >= g[3]) sel = 1;
    else if (g[2] >= g[0] && g[2] >= g[1] && g[2] >= g[3]) sel = 2;
    s_sel = sel;
    good_out[5 * b] = g[sel];
    for (int k = 0; k < 4; ++k) good_out[5 * b + 1 + k] = g[k];
    const double* R; double t[3];
    candidate(sel, R, t);
    for (int i = 0; i < 9; ++i) R_out[9 * b + i] = R[i];
    for (int i = 0; i < 3; ++i) t_out[3 * b + i] = t[i];
  }
  __syncthreads();
  const double* R; double t[3];
  candidate(s_sel, R, t);
  uint8_t* mask = mask_out + (size_t)b * n_max;
  double* tri = tri_out + (size_t)b * 4 * n_max;
  for (int i = tid; i < n_max; i += kPoseThreads) {
    double Q[4] = {0.0, 0.0, 0.0, 0.0};
    bool ok = false;
    if (i < n) {
      const float2 a = g1[i], c = g2[i];
      const double x1 = ((double)a.x - cx) / fx, y1 = ((double)a.y - cy) / fy;
      const double x2 = ((double)c.x - cx) / fx, y2 = ((double)c.y - cy) / fy;
      triangulate_dlt(R, t, x1, y1, x2, y2, Q);
      ok = cheirality_ok(R, t, Q, dist);
    }
    mask[i] = ok ? 255 : 0;
    for (int r = 0; r < 4; ++r) tri[(size_t)r * n_max + i] = Q[r];
  }
}

}  // namespace vslam

using namespace vslam;

extern "C" int vslam_undistort_points(const float* uv, int64_t n, const double* K, float* out, void* stream) {
  if (n < 0) return VSLAM_ERR_BAD_ARG;
  if (n == 0) return VSLAM_OK;
  if (!uv || !K || !out) return VSLAM_ERR_BAD_ARG;
  undistort_points_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      uv, n, K[0], K[1], K[2], K[3], out);
  return check_launch("undistort_points_kernel");
}

extern "C" int vslam_essential_hypotheses(const float* x1, const float* x2, const int32_t* n_pts, int n_max,
                                          int batch, const uint32_t* words, int H, double* models,
                                          int32_t* n_models, void* stream) {
  if (batch < 0 || H < 0 || n_max < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0 || H == 0) return VSLAM_OK;
  if (!x1 || !x2 || !n_pts || !words || !models || !n_models) return VSLAM_ERR_BAD_ARG;
  const int total = batch * H;
  essential_hypotheses_kernel<<<(total + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(
      x1, x2, n_pts, n_max, H, words, models, n_models, total);
  return check_launch("essential_hypotheses_kernel");
}

extern "C" int vslam_essential_score(const float* x1, const float* x2, const int32_t* n_pts, int n_max,
                                     int batch, const double* models, const int32_t* n_models, int H,
                                     double thr, int32_t* counts, void* stream) {
  if (batch < 0 || H < 0 || n_max < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0 || H == 0) return VSLAM_OK;
  if (!x1 || !x2 || !n_pts || !models || !n_models || !counts) return VSLAM_ERR_BAD_ARG;
  const size_t smem = (size_t)n_max * 2 * sizeof(float2);
  if (smem > 200 * 1024) return VSLAM_ERR_BAD_ARG;  // n_max <= 12800 correspondences per pair
  VSLAM_CUDA_TRY(cudaFuncSetAttribute(essential_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((H * kEssMaxModels + kEssScoreWarps - 1) / kEssScoreWarps, batch);
  essential_score_kernel<<<grid, kEssScoreWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
      x1, x2, n_pts, n_max, H, models, n_models, (float)(thr * thr), counts);
  return check_launch("essential_score_kernel");
}

extern "C" int vslam_essential_select(const float* x1, const float* x2, const int32_t* n_pts, int n_max,
                                      int batch, const double* models, const int32_t* n_models,
                                      const int32_t* counts, int H, double thr, double confidence,
                                      int32_t* best, double* E, uint8_t* mask, int32_t* n_inliers,
                                      double* score, void* stream) {
  if (batch < 0 || H < 0 || n_max < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!x1 || !x2 || !n_pts || !models || !n_models || !counts || !best || !E || !mask || !n_inliers || !score)
    return VSLAM_ERR_BAD_ARG;
  const size_t smem = (size_t)H * kEssMaxModels * sizeof(int32_t);
  if (smem > 200 * 1024) return VSLAM_ERR_BAD_ARG;  // H <= 5120 samples
  VSLAM_CUDA_TRY(cudaFuncSetAttribute(essential_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  essential_select_kernel<<<batch, kSelThreads, smem, static_cast<cudaStream_t>(stream)>>>(
      x1, x2, n_pts, n_max, H, models, n_models, counts, (float)(thr * thr), confidence, best, E, mask,
      n_inliers, score);
  return check_launch("essential_select_kernel");
}

extern "C" int vslam_essential_budget(const int32_t* n_pts, int n_max, int batch, const int32_t* n_models,
                                      const int32_t* counts, int H, int max_iters, double confidence,
                                      int32_t* budget, void* stream) {
  if (batch < 0 || H < 0 || n_max < 0 || max_iters < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!n_pts || !budget || (H > 0 && (!n_models || !counts))) return VSLAM_ERR_BAD_ARG;
  essential_budget_kernel<<<(batch + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(
      n_pts, n_max, batch, H, max_iters, n_models, counts, confidence, budget);
  return check_launch("essential_budget_kernel");
}

extern "C" int vslam_recover_pose(const double* E, const float* p1, const float* p2, const int32_t* n_pts,
                                  int n_max, int batch, const double* K, double distance_thresh, double* R,
                                  double* t, uint8_t* mask, double* tri, int32_t* good, void* stream) {
  if (batch < 0 || n_max < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!E || !p1 || !p2 || !n_pts || !K || !R || !t || !mask || !tri || !good) return VSLAM_ERR_BAD_ARG;
  recover_pose_kernel<<<batch, kPoseThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      E, p1, p2, n_pts, n_max, K[0], K[1], K[2], K[3], distance_thresh, R, t, mask, tri, good);
  return check_launch("recover_pose_kernel");
}
