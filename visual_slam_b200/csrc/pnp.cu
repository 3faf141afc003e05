// Batched RANSAC PnP: hypothesis generation (P3P+1), scoring (one warp per hypothesis), selection
// with cv2's adaptive iteration budget, and the final iterative refinement on the inliers.
// Replaces cv2.solvePnPRansac as called at reference src/v2/main.py:196-197.
#include "pnp_math.cuh"

namespace vslam {

constexpr int kSampleSize = 4;

// Harness sampling rule (oracle/pnp.py draw_samples): idx_j = (word_j * n) >> 32, duplicates step
// to the next free index modulo n.
__device__ __forceinline__ void draw_sample(const uint32_t* w, int n, int* idx) {
#pragma unroll
  for (int j = 0; j < kSampleSize; ++j) {
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

// One thread per (pair, hypothesis).
__global__ void __launch_bounds__(128)
pnp_hypotheses_kernel(const float* __restrict__ X, const float* __restrict__ uv,
                      const int32_t* __restrict__ n_pts, int n_max, int H,
                      const uint32_t* __restrict__ words, double4 K, double* __restrict__ hyps,
                      uint8_t* __restrict__ valid, int total) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  const int b = g / H;
  const int n = min(max(n_pts[b], 0), n_max);   // counts beyond the buffers are clamped
  double* out = hyps + (size_t)g * 12;
  bool ok = false;
  if (n >= kSampleSize) {
    int idx[kSampleSize];
    draw_sample(words + (size_t)g * kSampleSize, n, idx);
    double X4[4][3], uv4[4][2];
    for (int j = 0; j < kSampleSize; ++j) {
      const float* xp = X + ((size_t)b * n_max + idx[j]) * 3;
      const float* up = uv + ((size_t)b * n_max + idx[j]) * 2;
      X4[j][0] = xp[0]; X4[j][1] = xp[1]; X4[j][2] = xp[2];
      uv4[j][0] = up[0]; uv4[j][1] = up[1];
    }
    double Rt[12];
    ok = p3p_plus1(X4, uv4, K, Rt);
    if (ok) {
      for (int i = 0; i < 12; ++i) ok = ok && isfinite(Rt[i]);
    }
    if (ok) {
      for (int i = 0; i < 12; ++i) out[i] = Rt[i];
    }
  }
  if (!ok) {
    for (int i = 0; i < 12; ++i) out[i] = 0.0;
  }
  valid[g] = ok ? 1 : 0;
}

// Scoring: grid (ceil(H / 8), pairs), 8 warps; the pair's correspondences are staged once in shared
// memory (coalesced 128-bit loads), each warp owns one hypothesis and strides over the points.
constexpr int kScoreWarps = 8;

__global__ void __launch_bounds__(kScoreWarps * 32)
pnp_score_kernel(const float* __restrict__ X, const float* __restrict__ uv,
                 const int32_t* __restrict__ n_pts, int n_max, int H, const double* __restrict__ hyps,
                 const uint8_t* __restrict__ valid, double4 K, float thr2, int32_t* __restrict__ counts) {
  extern __shared__ __align__(16) float sp[];  // [n][3] then [n][2]
  const int b = blockIdx.y;
  const int n = min(max(n_pts[b], 0), n_max);   // counts beyond the buffers are clamped
  float* sX = sp;
  float* sU = sp + (size_t)n_max * 3;
  const float* gX = X + (size_t)b * n_max * 3;
  const float* gU = uv + (size_t)b * n_max * 2;
  for (int i = threadIdx.x; i < n * 3; i += blockDim.x) sX[i] = gX[i];
  for (int i = threadIdx.x; i < n * 2; i += blockDim.x) sU[i] = gU[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x * kScoreWarps + warp;
  if (h >= H) return;
  const size_t g = (size_t)b * H + h;
  int cnt = 0;
  if (valid[g]) {
    double Rt[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) Rt[i] = hyps[g * 12 + i];
    for (int i = lane; i < n; i += 32) {
      const float e = reproj_err_f32(Rt, sX[3 * i], sX[3 * i + 1], sX[3 * i + 2], sU[2 * i], sU[2 * i + 1], K);
      cnt += (e <= thr2) ? 1 : 0;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
  }
  if (lane == 0) counts[g] = cnt;
}

// cv::RANSACUpdateNumIters
__device__ inline int ransac_update_num_iters(double p, double ep, int model_points, int max_iters) {
  p = fmin(fmax(p, 0.0), 1.0);
  ep = fmin(fmax(ep, 0.0), 1.0);
  double num = fmax(1.0 - p, 2.2250738585072014e-308);
  double denom = 1.0 - pow(1.0 - ep, (double)model_points);
  if (denom < 2.2250738585072014e-308) return 0;
  num = log(num);
  denom = log(denom);
  return (denom >= 0 || -num >= max_iters * (-denom)) ? max_iters : (int)rint(num / denom);
}

constexpr int kRefineThreads = 128;
constexpr int kRefineVals = 28;  // 21 JtJ upper + 6 Jtr + cost

__device__ inline bool solve6(const double* Hu /*21 upper*/, const double* g, double lam, double* d) {
  double A[36];
  int k = 0;
  for (int i = 0; i < 6; ++i)
    for (int j = i; j < 6; ++j) { A[6 * i + j] = Hu[k]; A[6 * j + i] = Hu[k]; ++k; }
  for (int i = 0; i < 6; ++i) A[7 * i] += lam * A[7 * i];
  // Cholesky A = L L^T
  for (int j = 0; j < 6; ++j) {
    double s = A[7 * j];
    for (int q = 0; q < j; ++q) s -= A[6 * j + q] * A[6 * j + q];
    if (!(s > 0.0)) return false;
    s = sqrt(s);
    A[7 * j] = s;
    for (int i = j + 1; i < 6; ++i) {
      double t = A[6 * i + j];
      for (int q = 0; q < j; ++q) t -= A[6 * i + q] * A[6 * j + q];
      A[6 * i + j] = t / s;
    }
  }
  double y[6];
  for (int i = 0; i < 6; ++i) {
    double t = -g[i];
    for (int q = 0; q < i; ++q) t -= A[6 * i + q] * y[q];
    y[i] = t / A[7 * i];
  }
  for (int i = 5; i >= 0; --i) {
    double t = y[i];
    for (int q = i + 1; q < 6; ++q) t -= A[6 * q + i] * d[q];
    d[i] = t / A[7 * i];
  }
  return true;
}

// One CTA per pair: (1) sequential selection over the hypothesis counts with cv2's shrinking budget,
// (2) inlier mask of the winner, (3) Levenberg-Marquardt on the inliers from the caller's guess
// (cv2 Stage B: solvePnP(ITERATIVE, useExtrinsicGuess=True)); left-multiplicative rotation update.
__global__ void __launch_bounds__(kRefineThreads)
pnp_select_refine_kernel(const float* __restrict__ X, const float* __restrict__ uv,
                         const int32_t* __restrict__ n_pts, int n_max, int H,
                         const double* __restrict__ hyps, const uint8_t* __restrict__ valid,
                         const int32_t* __restrict__ counts, double4 K, float thr2, double confidence,
                         const double* __restrict__ guess, int max_refine_iters,
                         int32_t* __restrict__ best_out, uint32_t* __restrict__ inl_bits,
                         int32_t* __restrict__ n_inl_out, double* __restrict__ pose_out,
                         int32_t* __restrict__ status_out) {
  __shared__ int s_best;
  __shared__ double s_red[(kRefineThreads / 32) * kRefineVals];
  __shared__ double s_pose[12];   // R (9) + t (3), current
  __shared__ double s_trial[12];
  __shared__ double s_sys[kRefineVals];
  __shared__ int s_flag;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int n = min(max(n_pts[b], 0), n_max);   // counts beyond the buffers are clamped
  const int words = (n_max + 31) / 32;
  const float* gX = X + (size_t)b * n_max * 3;
  const float* gU = uv + (size_t)b * n_max * 2;
  uint32_t* bits = inl_bits + (size_t)b * words;
  if (tid == 0) {
    int niters = H, best = -1, best_count = 0;
    for (int it = 0; it < niters && it < H; ++it) {
      const size_t g = (size_t)b * H + it;
      if (!valid[g]) continue;
      const int c = counts[g];
      if (c > max(best_count, kSampleSize - 1)) {
        best = it; best_count = c;
        niters = ransac_update_num_iters(confidence, (double)(n - c) / (double)n, kSampleSize, niters);
      }
    }
    s_best = best;
  }
  __syncthreads();
  const int best = s_best;
  if (best < 0) {
    for (int w = tid; w < words; w += kRefineThreads) bits[w] = 0u;
    if (tid == 0) {
      best_out[b] = -1; n_inl_out[b] = 0; status_out[b] = 1;
      for (int i = 0; i < 6; ++i) pose_out[6 * b + i] = guess[6 * b + i];
    }
    return;
  }
  double Rt[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) Rt[i] = hyps[((size_t)b * H + best) * 12 + i];
  // inlier mask of the winning hypothesis: each warp ballots 32 consecutive points
  for (int base = (tid >> 5) * 32; base < words * 32; base += kRefineThreads) {
    const int i = base + (tid & 31);
    bool in = false;
    if (i < n) in = reproj_err_f32(Rt, gX[3 * i], gX[3 * i + 1], gX[3 * i + 2], gU[2 * i], gU[2 * i + 1], K) <= thr2;
    const unsigned m = __ballot_sync(0xffffffffu, in);
    if ((tid & 31) == 0) bits[base >> 5] = m;
  }
  if (tid == 0) {
    best_out[b] = best;
    n_inl_out[b] = counts[(size_t)b * H + best];
    rodrigues(guess + 6 * b, s_pose);
    s_pose[9] = guess[6 * b + 3]; s_pose[10] = guess[6 * b + 4]; s_pose[11] = guess[6 * b + 5];
  }
  __syncthreads();

  // accumulate the normal equations of sum |proj - uv|^2 over the inliers at pose P
  auto accumulate = [&](const double* P, bool with_jac) {
    double acc[kRefineVals];
#pragma unroll
    for (int i = 0; i < kRefineVals; ++i) acc[i] = 0.0;
    for (int i = tid; i < n; i += kRefineThreads) {
      if (!((bits[i >> 5] >> (i & 31)) & 1u)) continue;
      const double Xw = gX[3 * i], Yw = gX[3 * i + 1], Zw = gX[3 * i + 2];
      const double px = P[0] * Xw + P[1] * Yw + P[2] * Zw + P[9];
      const double py = P[3] * Xw + P[4] * Yw + P[5] * Zw + P[10];
      const double pz = P[6] * Xw + P[7] * Yw + P[8] * Zw + P[11];
      const double iz = 1.0 / pz;
      const double r0 = K.x * px * iz + K.z - (double)gU[2 * i];
      const double r1 = K.y * py * iz + K.w - (double)gU[2 * i + 1];
      acc[27] += r0 * r0 + r1 * r1;
      if (!with_jac) continue;
      // d pc = dw x pc + dt  (columns 0..2 rotation, 3..5 translation)
      const double a = K.x * iz, c = K.y * iz;
      const double J0[6] = {-a * px * py * iz, a * (pz + px * px * iz), -a * py, a, 0.0, -a * px * iz};
      const double J1[6] = {-c * (pz + py * py * iz), c * px * py * iz, c * px, 0.0, c, -c * py * iz};
      int k = 0;
#pragma unroll
      for (int p = 0; p < 6; ++p) {
#pragma unroll
        for (int q = p; q < 6; ++q) acc[k++] += J0[p] * J0[q] + J1[p] * J1[q];
      }
#pragma unroll
      for (int p = 0; p < 6; ++p) acc[21 + p] += J0[p] * r0 + J1[p] * r1;
    }
    // deterministic block reduction
#pragma unroll
    for (int i = 0; i < kRefineVals; ++i) acc[i] = warp_sum(acc[i]);
    if ((tid & 31) == 0) {
#pragma unroll
      for (int i = 0; i < kRefineVals; ++i) s_red[(tid >> 5) * kRefineVals + i] = acc[i];
    }
    __syncthreads();
    if (tid < kRefineVals) {
      double s = 0.0;
      for (int w = 0; w < kRefineThreads / 32; ++w) s += s_red[w * kRefineVals + tid];
      s_sys[tid] = s;
    }
    __syncthreads();
  };

  accumulate(s_pose, true);
  double lam = 1e-3;
  double cost = s_sys[27];
  int status = 0;
  for (int it = 0; it < max_refine_iters; ++it) {
    if (tid == 0) {
      double d[6];
      int flag = 0;
      if (!solve6(s_sys, s_sys + 21, lam, d)) flag = 2;  // damped system not SPD: raise lambda
      else {
        double dR[9];
        rodrigues(d, dR);
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j)
            s_trial[3 * i + j] = dR[3 * i] * s_pose[j] + dR[3 * i + 1] * s_pose[3 + j] + dR[3 * i + 2] * s_pose[6 + j];
        for (int i = 0; i < 3; ++i)  // T <- [exp(dw) | dt] T, consistent with d pc = dw x pc + dt
          s_trial[9 + i] = dR[3 * i] * s_pose[9] + dR[3 * i + 1] * s_pose[10] + dR[3 * i + 2] * s_pose[11] + d[3 + i];
        const double nd = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2] + d[3] * d[3] + d[4] * d[4] + d[5] * d[5]);
        if (nd < 1e-14) flag = 1;  // converged
      }
      s_flag = flag;
    }
    __syncthreads();
    const int flag = s_flag;
    if (flag == 2) { lam *= 10.0; if (lam > 1e12) { status = 2; break; } __syncthreads(); continue; }
    // evaluate the trial pose (cost only), then relinearise if accepted
    double saved = 0.0;
    if (tid < kRefineVals) saved = s_sys[tid];
    __syncthreads();
    accumulate(s_trial, false);
    const double trial_cost = s_sys[27];
    __syncthreads();
    if (trial_cost <= cost && isfinite(trial_cost)) {
      if (tid < 12) s_pose[tid] = s_trial[tid];
      __syncthreads();
      cost = trial_cost;
      lam = fmax(lam * 0.1, 1e-12);
      if (flag == 1) break;
      accumulate(s_pose, true);
    } else {
      if (tid < kRefineVals) s_sys[tid] = saved;
      __syncthreads();
      lam *= 10.0;
      if (flag == 1 || lam > 1e12) break;
    }
  }
  if (tid == 0) {
    double r[3];
    rodrigues_inv(s_pose, r);
    double* o = pose_out + 6 * b;
    o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = s_pose[9]; o[4] = s_pose[10]; o[5] = s_pose[11];
    status_out[b] = status;
  }
}

}  // namespace vslam

using namespace vslam;

static inline double4 k4(const double* K) { return make_double4(K[0], K[1], K[2], K[3]); }

extern "C" int vslam_pnp_hypotheses(const float* X, const float* uv, const int32_t* n_pts, int n_max,
                                    int batch, const uint32_t* words, int H, const double* K,
                                    double* hyps, uint8_t* valid, void* stream) {
  if (batch < 0 || H < 0 || n_max < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0 || H == 0) return VSLAM_OK;
  if (!X || !uv || !n_pts || !words || !K || !hyps || !valid) return VSLAM_ERR_BAD_ARG;
  const int total = batch * H;
  pnp_hypotheses_kernel<<<(total + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      X, uv, n_pts, n_max, H, words, k4(K), hyps, valid, total);
  return check_launch("pnp_hypotheses_kernel");
}

extern "C" int vslam_pnp_score(const float* X, const float* uv, const int32_t* n_pts, int n_max, int batch,
                               const double* hyps, const uint8_t* valid, int H, const double* K,
                               double thr, int32_t* counts, void* stream) {
  if (batch < 0 || H < 0 || n_max < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0 || H == 0) return VSLAM_OK;
  if (!X || !uv || !n_pts || !hyps || !valid || !K || !counts) return VSLAM_ERR_BAD_ARG;
  const size_t smem = (size_t)n_max * 5 * sizeof(float);
  if (smem > 200 * 1024) return VSLAM_ERR_BAD_ARG;  // n_max <= 10240 correspondences per pair
  VSLAM_CUDA_TRY(cudaFuncSetAttribute(pnp_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((H + kScoreWarps - 1) / kScoreWarps, batch);
  pnp_score_kernel<<<grid, kScoreWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
      X, uv, n_pts, n_max, H, hyps, valid, k4(K), (float)(thr * thr), counts);
  return check_launch("pnp_score_kernel");
}

extern "C" int vslam_pnp_select_refine(const float* X, const float* uv, const int32_t* n_pts, int n_max,
                                       int batch, const double* hyps, const uint8_t* valid,
                                       const int32_t* counts, int H, const double* K, double thr,
                                       double confidence, const double* guess, int max_refine_iters,
                                       int32_t* best, uint32_t* inlier_bits, int32_t* n_inliers,
                                       double* pose, int32_t* status, void* stream) {
  if (batch < 0 || H < 0 || n_max < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!X || !uv || !n_pts || !hyps || !valid || !counts || !K || !guess || !best || !inlier_bits ||
      !n_inliers || !pose || !status)
    return VSLAM_ERR_BAD_ARG;
  pnp_select_refine_kernel<<<batch, kRefineThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      X, uv, n_pts, n_max, H, hyps, valid, counts, k4(K), (float)(thr * thr), confidence, guess,
      max_refine_iters > 0 ? max_refine_iters : 50, best, inlier_bits, n_inliers, pose, status);
  return check_launch("pnp_select_refine_kernel");
}

// Correspondence gather between matching and PnP (reference src/v2/main.py:187-188 and
// frame.py:36-37): X[b][i] = map_xyz[b][keep_q[b][i]], uv[b][i] = kp_cur[b][keep_t[b][i]].
namespace vslam {
__global__ void gather_corr_kernel(const int32_t* __restrict__ keep_q, const int32_t* __restrict__ keep_t,
                                   const int32_t* __restrict__ n_keep, int nq, int nt,
                                   const float* __restrict__ map_xyz, const float* __restrict__ kp_cur,
                                   float* __restrict__ X, float* __restrict__ uv) {
  const int b = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  float x = 0.f, y = 0.f, z = 0.f, u = 0.f, v = 0.f;
  if (i < n_keep[b]) {
    const int q = keep_q[(size_t)b * nq + i], t = keep_t[(size_t)b * nq + i];
    const float* xp = map_xyz + ((size_t)b * nq + q) * 3;
    const float* up = kp_cur + ((size_t)b * nt + t) * 2;
    x = xp[0]; y = xp[1]; z = xp[2]; u = up[0]; v = up[1];
  }
  float* xo = X + ((size_t)b * nq + i) * 3;
  float* uo = uv + ((size_t)b * nq + i) * 2;
  xo[0] = x; xo[1] = y; xo[2] = z; uo[0] = u; uo[1] = v;
}
}  // namespace vslam

extern "C" int vslam_gather_correspondences(const int32_t* keep_q, const int32_t* keep_t,
                                            const int32_t* n_keep, int nq, int nt, int batch,
                                            const float* map_xyz, const float* kp_cur, float* X,
                                            float* uv, void* stream) {
  if (batch < 0 || nq < 0 || nt < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0 || nq == 0) return VSLAM_OK;
  if (!keep_q || !keep_t || !n_keep || !map_xyz || !kp_cur || !X || !uv) return VSLAM_ERR_BAD_ARG;
  dim3 grid((nq + 255) / 256, batch);
  vslam::gather_corr_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      keep_q, keep_t, n_keep, nq, nt, map_xyz, kp_cur, X, uv);
  return vslam::check_launch("gather_corr_kernel");
}
