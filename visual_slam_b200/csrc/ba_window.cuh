// Device-resident tracking window: the local map of the reference's tracking loop kept in HBM and
// appended to as frames arrive, with the motion-only bundle adjustment that runs after EVERY tracked frame
// (reference src/v2/main.py:208-214, src/v2/LocalBA.py:195-229) as ONE kernel launch.
//
// The reference rebuilds a g2o graph from its Map objects for each of those calls (O(points x frames) dict probes,
// one pybind call per vertex and edge, a fresh optimiser).  Here the window is three arrays:
//   pts    [P][4]            the map points the last key frame sees (fixed in motion-only BA), padded to 32 B
//   cams   [F][7]            (qx qy qz qw tx ty tz) camera->world; row 0 = the key frame (fixed), rows 1.. = tracked frames
//   obs    [F][slot]         per frame: point row + measured pixel, n_obs[F] valid entries -- a frame's observations
//                            arrive together, so appending a frame IS camera-major order: nothing is ever sorted
// Motion-only BA has no free points, hence no Schur complement: the normal equations are one independent 6x6 block per
// free frame, coupled only through the Levenberg-Marquardt controller (one lambda, one cost, one accept / reject for
// all frames, as g2o's OptimizationAlgorithmLevenberg does).  One CTA per frame; per LM trial each CTA linearises
// its frame if the state changed, solves its 6x6 system, applies SBACam::update and evaluates the trial cost; ONE grid
// barrier per trial publishes (cost, d.(lambda d + b), solver status) and every CTA takes the same decision from
// the same numbers summed in frame order.  Fixed-order reductions everywhere: bit-reproducible.
#pragma once
#include "ba_kernels.cuh"
#include "ba_pcg.cuh"

namespace vslam {

constexpr int kWinThreads = 256;
constexpr int kWinPart = 4;     // per frame and trial: cost, scale term, solver ok, max |diag|

struct WinView {
  int n_frames, slot, max_iterations;
  double4 K;
  const double* pts4;
  double* cams;                 // [F][7], updated in place at the end
  const uint8_t* fixed;
  const int* obs_pt; const double2* obs_uv; const int* n_obs;
  double* part;                 // [3][F][kWinPart]: first linearisation, then two trial buffers (parity)
  unsigned* bar;
  double* stats;                // [8] like vslam_ba_optimize
};

__global__ void window_reset_kernel(const double* __restrict__ pts, const float* __restrict__ kf_uv,
                                    const double* __restrict__ kf_cam7, int P, int slot, double* __restrict__ pts4,
                                    double* __restrict__ cams, uint8_t* __restrict__ fixed, int* __restrict__ obs_pt,
                                    double2* __restrict__ obs_uv, int* __restrict__ n_obs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P) {
    *reinterpret_cast<double4*>(pts4 + 4 * (size_t)i) = make_double4(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], 0.0);
    if (i < slot) { obs_pt[i] = i; obs_uv[i] = make_double2((double)kf_uv[2 * i], (double)kf_uv[2 * i + 1]); }
  }
  if (i < 7) cams[i] = kf_cam7[i];
  if (i == 0) { fixed[0] = 1; n_obs[0] = min(P, slot); }
}

// main.py:208-210: the new frame's pose and one observation per ratio match (point row keep_q[i] of the key frame's
// list, pixel kp_cur[keep_t[i]]), straight from the device outputs of matching / PnP / pose plumbing.
__global__ void window_append_kernel(int f, int slot, const double* __restrict__ cam7, const int* __restrict__ keep_q,
                                     const int* __restrict__ keep_t, const int* __restrict__ n_keep,
                                     const float* __restrict__ kp_cur, int n_pts, double* __restrict__ cams,
                                     uint8_t* __restrict__ fixed, int* __restrict__ obs_pt,
                                     double2* __restrict__ obs_uv, int* __restrict__ n_obs) {
  const int n = min(max(*n_keep, 0), slot);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const int q = keep_q[i], t = keep_t[i];
    const bool ok = q >= 0 && q < n_pts && t >= 0;
    obs_pt[(size_t)f * slot + i] = ok ? q : -1;                 // a row outside the point list is ignored by BA
    obs_uv[(size_t)f * slot + i] = ok ? make_double2((double)kp_cur[2 * (size_t)t], (double)kp_cur[2 * (size_t)t + 1])
                                      : make_double2(0.0, 0.0);
  }
  if (i < 7) cams[7 * (size_t)f + i] = cam7[i];
  if (i == 0) { fixed[f] = 0; n_obs[f] = n; }
}

__device__ __forceinline__ void win_update_cam(const double* q, const double* t, const double* d, double* qo, double* to) {
  to[0] = t[0] + d[0]; to[1] = t[1] + d[1]; to[2] = t[2] + d[2];
  const double vx = d[3], vy = d[4], vz = d[5];
  const double vw = sqrt(1.0 - (vx * vx + vy * vy + vz * vz));
  const double x = q[3] * vx + q[0] * vw + q[1] * vz - q[2] * vy;
  const double y = q[3] * vy - q[0] * vz + q[1] * vw + q[2] * vx;
  const double z = q[3] * vz + q[0] * vy - q[1] * vx + q[2] * vw;
  const double w = q[3] * vw - q[0] * vx - q[1] * vy - q[2] * vz;
  const double inv = 1.0 / sqrt(x * x + y * y + z * z + w * w);
  qo[0] = x * inv; qo[1] = y * inv; qo[2] = z * inv; qo[3] = w * inv;
}

__global__ void __launch_bounds__(kWinThreads)
window_motion_ba_kernel(WinView V) {
  __shared__ double s_red[(kWinThreads / 32) * 28];
  __shared__ double s_sys[28];          // Hpp upper (21), bp (6), cost
  __shared__ double s_q[2][4], s_t[2][3], s_R[2][9];   // [0] current, [1] trial
  __shared__ double s_d[6];
  __shared__ double s_dec[4];           // decision inputs broadcast: tmp, scale, all ok
  const int f = blockIdx.x, tid = threadIdx.x, F = V.n_frames;
  const bool is_fixed = V.fixed[f] != 0;
  const int n = V.n_obs[f];
  const int* __restrict__ opt = V.obs_pt + (size_t)f * V.slot;
  const double2* __restrict__ ouv = V.obs_uv + (size_t)f * V.slot;
  if (tid < 4) s_q[0][tid] = V.cams[7 * (size_t)f + tid];
  if (tid < 3) s_t[0][tid] = V.cams[7 * (size_t)f + 4 + tid];
  __syncthreads();
  if (tid == 0) quat_to_R(s_q[0], s_R[0]);
  __syncthreads();

  // one pass over the frame's observations at state `which`; with_jac: the 28-value system, else the cost only
  auto pass = [&](int which, bool with_jac) {
    double v[28];
#pragma unroll
    for (int i = 0; i < 28; ++i) v[i] = 0.0;
    for (int o = tid; o < n; o += kWinThreads) {
      const int p = opt[o];
      if (p < 0) continue;
      double X[3];
      load_point(V.pts4, p, X);
      double e[2], rho, w, Jp[2][3], Jr[2][3];
      if (!eval_obs(s_R[which], s_t[which], X, ouv[o], V.K, with_jac, e, rho, w, Jp, Jr)) continue;
      v[27] += rho;
      if (!with_jac) continue;
      // camera Jacobian: columns 0..2 = -Jp (translation), 3..5 = Jr (rotation); H += w J^T J, b += -w J^T e
      double J0[6] = {-Jp[0][0], -Jp[0][1], -Jp[0][2], Jr[0][0], Jr[0][1], Jr[0][2]};
      double J1[6] = {-Jp[1][0], -Jp[1][1], -Jp[1][2], Jr[1][0], Jr[1][1], Jr[1][2]};
      int k = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
        const double wa0 = w * J0[a], wa1 = w * J1[a];
#pragma unroll
        for (int b = a; b < 6; ++b) v[k++] += wa0 * J0[b] + wa1 * J1[b];
        v[21 + a] -= wa0 * e[0] + wa1 * e[1];
      }
    }
    block_reduce<28, kWinThreads>(v, s_red);
    if (tid == 0) {
      if (with_jac) {
#pragma unroll
        for (int i = 0; i < 28; ++i) s_sys[i] = v[i];
      } else {
        s_sys[27] = v[27];
      }
    }
    __syncthreads();
  };

  unsigned target = 0;
  double* part0 = V.part;
  // ---- first linearisation: cost of the state as given, computeLambdaInit = tau * max |diag H| over the free frames
  pass(0, !is_fixed);
  const double my_lin_cost = s_sys[27];
  if (tid == 0) {
    double md = 0.0;
    if (!is_fixed) {
      int k = 0;
      for (int a = 0; a < 6; ++a) { md = fmax(md, fabs(s_sys[k])); k += 6 - a; }
    }
    part0[(size_t)f * kWinPart] = my_lin_cost; part0[(size_t)f * kWinPart + 3] = md;
  }
  grid_barrier(V.bar, target, F);
  double cur_cost = 0.0, max_diag = 0.0;
  for (int g = 0; g < F; ++g) {   // every thread of every CTA: the same numbers in the same order
    cur_cost += __ldcg(part0 + (size_t)g * kWinPart);
    max_diag = fmax(max_diag, __ldcg(part0 + (size_t)g * kWinPart + 3));
  }
  const double initial_cost = cur_cost;
  double lambda = 1e-5 * max_diag, ni = 2.0;
  int iters = 0, trials = 0, q = 0, rejected = 0, breakdowns = 0;
  bool done = V.max_iterations <= 0, dirty = false;
  double frame_cost = my_lin_cost;           // this frame's share of cur_cost
  int slot_i = 0;
  while (!done) {
    if (dirty && !is_fixed) pass(0, true);
    dirty = false;
    // ---- solve (Hpp + lambda I) d = bp, trial camera, this frame's computeScale term
    if (tid == 0) {
      double sc = 0.0, okf = 1.0;
      if (!is_fixed) {
        double A[36];
        int k = 0;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
#pragma unroll
          for (int b = a; b < 6; ++b) { A[6 * a + b] = s_sys[k]; A[6 * b + a] = s_sys[k]; ++k; }
          A[7 * a] += lambda;
        }
        double d[6] = {0, 0, 0, 0, 0, 0};
        if (spd6_inverse(A)) {
#pragma unroll
          for (int a = 0; a < 6; ++a) {
            double s = 0.0;
#pragma unroll
            for (int b = 0; b < 6; ++b) s = fma(A[6 * a + b], s_sys[21 + b], s);
            d[a] = s;
          }
        } else {
          okf = 0.0;
        }
#pragma unroll
        for (int a = 0; a < 6; ++a) { s_d[a] = d[a]; sc += d[a] * (lambda * d[a] + s_sys[21 + a]); }
        win_update_cam(s_q[0], s_t[0], d, s_q[1], s_t[1]);
        quat_to_R(s_q[1], s_R[1]);
      }
      s_dec[0] = sc; s_dec[1] = okf;
    }
    __syncthreads();
    double trial_cost = frame_cost;
    if (!is_fixed) { pass(1, false); trial_cost = s_sys[27]; }
    double* part = V.part + (size_t)(1 + (slot_i & 1)) * F * kWinPart;
    if (tid == 0) {
      part[(size_t)f * kWinPart] = trial_cost; part[(size_t)f * kWinPart + 1] = s_dec[0];
      part[(size_t)f * kWinPart + 2] = s_dec[1];
    }
    grid_barrier(V.bar, target, F);
    ++slot_i;
    double tmp = 0.0, scale = 0.0, all_ok = 1.0;
    for (int g = 0; g < F; ++g) {
      tmp += __ldcg(part + (size_t)g * kWinPart);
      scale += __ldcg(part + (size_t)g * kWinPart + 1);
      all_ok = fmin(all_ok, __ldcg(part + (size_t)g * kWinPart + 2));
    }
    // ---- OptimizationAlgorithmLevenberg::solve, one trial (same statements as trial_controller in ba_kernels.cuh)
    if (all_ok == 0.0) { tmp = 1.7976931348623157e308; breakdowns += 1; }
    const double rho = (cur_cost - tmp) / (scale + 1e-3);
    trials += 1;
    const bool accepted = rho > 0.0 && isfinite(tmp);
    if (accepted) {
      double alpha = 1.0 - pow(2.0 * rho - 1.0, 3.0);
      alpha = fmin(alpha, 2.0 / 3.0);
      lambda *= fmax(1.0 / 3.0, alpha);
      ni = 2.0;
      cur_cost = tmp;
      frame_cost = trial_cost;
      __syncthreads();
      if (!is_fixed) {
        if (tid < 4) s_q[0][tid] = s_q[1][tid];
        if (tid < 3) s_t[0][tid] = s_t[1][tid];
        if (tid < 9) s_R[0][tid] = s_R[1][tid];
      }
      __syncthreads();
      dirty = true;
    } else {
      lambda *= ni;
      ni *= 2.0;
      rejected += 1;
    }
    const bool lam_bad = !accepted && !isfinite(lambda);
    if (!lam_bad) q += 1;
    if (!lam_bad && rho < 0.0 && q < 10) continue;     // next trial of the same outer iteration
    iters += 1;
    const bool stop = (q == 10) || rho == 0.0 || !isfinite(lambda);
    q = 0;
    if (stop || iters >= V.max_iterations) done = true;
  }
  if (!is_fixed) {
    if (tid < 4) V.cams[7 * (size_t)f + tid] = s_q[0][tid];
    if (tid < 3) V.cams[7 * (size_t)f + 4 + tid] = s_t[0][tid];
  }
  if (f == 0 && tid == 0) {
    V.stats[0] = initial_cost; V.stats[1] = cur_cost; V.stats[2] = iters; V.stats[3] = trials;
    V.stats[4] = lambda; V.stats[5] = 0.0; V.stats[6] = rejected; V.stats[7] = breakdowns;
  }
}

struct Window {
  int max_frames = 0, slot = 0, max_pts = 0, n_frames = 0, n_pts = 0;
  double4 K = make_double4(0, 0, 0, 0);
  double* pts4 = nullptr; double* cams = nullptr; uint8_t* fixed = nullptr;
  int* obs_pt = nullptr; double2* obs_uv = nullptr; int* n_obs = nullptr;
  double* part = nullptr; unsigned* bar = nullptr; double* stats = nullptr;
  double* stats_host = nullptr;
  ~Window() {
    cudaFree(pts4); cudaFree(cams); cudaFree(fixed); cudaFree(obs_pt); cudaFree(obs_uv); cudaFree(n_obs);
    cudaFree(part); cudaFree(bar); cudaFree(stats);
    if (stats_host) cudaFreeHost(stats_host);
  }
};

}  // namespace vslam
