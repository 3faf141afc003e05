// Bundle-adjustment kernels: linearisation (fused residual + Jacobian + Huber + block accumulation),
// Schur complement by sorted segments, back-substitution, state update, cost.
// Semantics: SURVEY.md 8a-3 (g2o BlockSolver<6,3>); call sites reference src/v2/LocalBA.py:143-229.
//
// Two static orderings of the observations are kept in HBM (built once per problem):
//   cm  camera-major  (cam, point)  -> per-camera reductions (Hpp, bp, E*bl), Hpl / E storage
//   pm  point-major   (point, cam)  -> per-point reductions (Hll, bl, back-substitution)
// Every reduction is a fixed-order tree (warp butterfly -> per-CTA -> ordered chunk sum), so
// a BA run is bit-reproducible: no floating-point atomics anywhere.
#pragma once
#include "ba_math.cuh"

namespace vslam {

constexpr int kCamChunk = 256;    // observations of one camera handled by one CTA (one per thread)
constexpr int kCamThreads = 256;
constexpr int kCamVals = 28;      // 21 Hpp upper + 6 bp + 1 cost
constexpr int kUnitItems = 256;   // Schur items per warp work unit

// Levenberg-Marquardt controller state, resident in HBM: every kernel of an LM step reads it, the
// two controller kernels (ba_lin_controller / ba_trial_controller) are the only writers.  A whole
// optimize() call is enqueued without host round trips (one CUDA graph launch per trial slot).
struct LMState {
  int cur;             // index of the CURRENT state buffers (the trial state is cur ^ 1)
  int dirty;           // 1: the system must be rebuilt from the current state before the next trial
  int done;            // 1: the optimisation has terminated; every kernel returns at once
  int manual;          // 1: stage API (vslam_ba_linearize/trial): controllers only record results
  int first;           // 1 until lambda has been initialised (computeLambdaInit)
  int iters, trials, q, max_iterations;
  int last_pcg_status, last_pcg_iters, pad_;
  double lambda, ni, cur_cost, initial_cost, max_diag;
  double pcg_total, rejected, breakdowns;
  double last_cost, last_scale, last_rho;
};

struct BAView {
  int C, P, N, Cf, Pf, E, max_chunks;
  double4 K;
  const LMState* st;
  double* cams[2]; double* pts[2]; CamRT* camrt[2];
  const int* cam_col; const int* pt_col;
  const int* cm_start; const int* cm_pt; const int* cm_pm; const double2* cm_uv;
  const int* pt_start; const int* pm_cam; const double2* pm_uv;
};

__global__ void ba_prepare_cams(const double* __restrict__ cams, CamRT* __restrict__ out, int C) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  CamRT r;
  quat_to_R(cams + 7 * c, r.R);
  r.t[0] = cams[7 * c + 4]; r.t[1] = cams[7 * c + 5]; r.t[2] = cams[7 * c + 6];
  out[c] = r;
}

__device__ __forceinline__ void load18(const double* p, double (&v)[18]) {
  const double2* s = reinterpret_cast<const double2*>(p);
#pragma unroll
  for (int i = 0; i < 9; ++i) { double2 t = s[i]; v[2 * i] = t.x; v[2 * i + 1] = t.y; }
}
__device__ __forceinline__ void store18(double* p, const double (&v)[18]) {
  double2* s = reinterpret_cast<double2*>(p);
#pragma unroll
  for (int i = 0; i < 9; ++i) s[i] = make_double2(v[2 * i], v[2 * i + 1]);
}

// Sum over the 32 lanes of 32 per-lane values in 31 shuffles: after the call lane L holds the warp
// total of value L in v[0].  Each step halves the live values (lanes with bit s keep the upper half
// and ship the lower one).  The combination tree is fixed, so the result is deterministic.
__device__ __forceinline__ void warp_transpose_sum32(double (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const double send = up ? v[i] : v[i + s];
      const double keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
}

// Camera-major pass, one thread per observation.  COST_ONLY = evaluate sum(rho) for a trial state;
// otherwise also accumulate Hpp (upper 21) / bp per camera and write the Hpl blocks
// (w Jc^T Jp, 6x3 row-major).  Values 0..20 Hpp upper, 21..26 bp, 27 cost.
template <bool COST_ONLY>
__global__ void __launch_bounds__(kCamThreads)
ba_cam_pass(BAView V, double* __restrict__ Hpl, double* __restrict__ camp) {
  if (V.st->done || (!COST_ONLY && !V.st->dirty)) return;
  const int sb = COST_ONLY ? (V.st->cur ^ 1) : V.st->cur;   // trial state for the cost-only pass
  const CamRT* __restrict__ camrt = V.camrt[sb];
  const double* __restrict__ pts = V.pts[sb];
  __shared__ double red[(kCamThreads / 32) * 32];
  __shared__ __align__(16) double stage[COST_ONLY ? 2 : kCamThreads * 18];  // Hpl blocks of the CTA
  const int c = blockIdx.y, chunk = blockIdx.x;
  const int start = V.cm_start[c] + chunk * kCamChunk;
  const int end = min(V.cm_start[c + 1], start + kCamChunk);
  if (start >= end) return;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const CamRT cam = camrt[c];
  const bool cam_free = V.cam_col[c] >= 0;
  double v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = 0.0;
  const int o = start + threadIdx.x;
  double h[18];
#pragma unroll
  for (int i = 0; i < 18; ++i) h[i] = 0.0;
  if (o < end) {
    const int p = V.cm_pt[o];
    const double X[3] = {pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]};
    const double2 uv = V.cm_uv[o];
    double pc[3], e[2], rho, w;
    to_camera(cam, X, pc);
    if (fabs(pc[2]) > 1e-12) {  // g2o aborts on z = 0; we drop the observation
      residual(pc, V.K, uv.x, uv.y, e);
      huber(e[0] * e[0] + e[1] * e[1], rho, w);
      v[27] = rho;
      if (!COST_ONLY && cam_free) {
        double Jc[2][6], Jp[2][3], W0[6], W1[6];
        jacobians(cam, pc, V.K, Jc, Jp);
#pragma unroll
        for (int i = 0; i < 6; ++i) { W0[i] = w * Jc[0][i]; W1[i] = w * Jc[1][i]; }
        int k = 0;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
#pragma unroll
          for (int j = i; j < 6; ++j) v[k++] = W0[i] * Jc[0][j] + W1[i] * Jc[1][j];
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) v[21 + i] = -(W0[i] * e[0] + W1[i] * e[1]);
#pragma unroll
        for (int i = 0; i < 6; ++i) {
#pragma unroll
          for (int j = 0; j < 3; ++j) h[3 * i + j] = W0[i] * Jp[0][j] + W1[i] * Jp[1][j];
        }
      }
    }
  }
  if (!COST_ONLY && cam_free) store18(stage + threadIdx.x * 18, h);  // 144 B stride: conflict-free STS.128
  if (!COST_ONLY && cam_free) {
    // the CTA's blocks are contiguous in HBM (camera-major): write them with coalesced 128-bit stores
    __syncthreads();
    const int n_blk = end - start;
    double2* dst = reinterpret_cast<double2*>(Hpl + (size_t)start * 18);
    const double2* src = reinterpret_cast<const double2*>(stage);
    for (int i = threadIdx.x; i < n_blk * 9; i += kCamThreads) dst[i] = src[i];
  }
  if (COST_ONLY) {
    double s = warp_sum(v[27]);
    if (lane == 0) red[wid] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w8 = 0; w8 < kCamThreads / 32; ++w8) t += red[w8];
      camp[((size_t)c * V.max_chunks + chunk) * kCamVals + 27] = t;
    }
    return;
  }
  warp_transpose_sum32(v, lane);
  red[wid * 32 + lane] = v[0];
  __syncthreads();
  if (threadIdx.x < kCamVals) {
    double t = 0.0;
#pragma unroll
    for (int w8 = 0; w8 < kCamThreads / 32; ++w8) t += red[w8 * 32 + threadIdx.x];
    camp[((size_t)c * V.max_chunks + chunk) * kCamVals + threadIdx.x] = t;
  }
}

// Ordered sum of the chunk partials of each camera -> Hpp [Cf][36] (full symmetric), bp [Cf][6]
// (+ scale-edge part), per-camera cost.
__device__ __forceinline__ void atomic_max_nonneg(unsigned long long* bits, double v) {
  atomicMax(bits, (unsigned long long)__double_as_longlong(fabs(v)));  // exact and order-independent
}

__global__ void ba_cam_finalize(BAView V, const double* __restrict__ camp,
                                const double* __restrict__ bp_extra, const int* __restrict__ diag_extra,
                                const double* __restrict__ extra, double* __restrict__ Hpp,
                                double* __restrict__ bp, double* __restrict__ cam_cost,
                                unsigned long long* __restrict__ maxdiag_bits) {
  const int c = blockIdx.x, i = threadIdx.x;
  if (V.st->done || (Hpp != nullptr && !V.st->dirty)) return;
  if (i >= kCamVals) return;
  const int n_obs = V.cm_start[c + 1] - V.cm_start[c];
  const int n_chunks = (n_obs + kCamChunk - 1) / kCamChunk;
  double s = 0.0;
  for (int k = 0; k < n_chunks; ++k) s += camp[((size_t)c * V.max_chunks + k) * kCamVals + i];
  if (i == 27) { cam_cost[c] = s; return; }
  const int col = V.cam_col[c];
  if (col < 0 || Hpp == nullptr) return;
  if (i >= 21) { bp[6 * col + (i - 21)] = s + bp_extra[6 * col + (i - 21)]; return; }
  int r = 0, k = i;
  while (k >= 6 - r) { k -= 6 - r; ++r; }
  const int cc = r + k;
  Hpp[36 * col + 6 * r + cc] = s;
  Hpp[36 * col + 6 * cc + r] = s;
  if (r == cc) {  // computeLambdaInit: diagonal of the full camera block (projection + scale edges)
    const int ex = diag_extra[col];
    atomic_max_nonneg(maxdiag_bits, s + (ex >= 0 ? extra[36 * (size_t)ex + 7 * r] : 0.0));
  }
}

// Point-major pass: a quad of 4 lanes per point; lane q takes observations q, q+4, ... of the point,
// the quad is reduced by two butterfly steps (fixed order -> deterministic).
constexpr int kPointLanes = 4;

__global__ void __launch_bounds__(256)
ba_point_pass(BAView V, double* __restrict__ Hll, double* __restrict__ bl,
              unsigned long long* __restrict__ maxdiag_bits) {
  if (V.st->done || !V.st->dirty) return;
  const CamRT* __restrict__ camrt = V.camrt[V.st->cur];
  const double* __restrict__ pts = V.pts[V.st->cur];
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  const int p = g / kPointLanes, q = g % kPointLanes;
  const int col = (p < V.P) ? V.pt_col[p] : -1;
  double acc[9];
#pragma unroll
  for (int j = 0; j < 9; ++j) acc[j] = 0.0;
  if (col >= 0) {
    const double X[3] = {pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]};
    const int o_end = V.pt_start[p + 1];
    for (int o = V.pt_start[p] + q; o < o_end; o += kPointLanes) {
      const CamRT cam = camrt[V.pm_cam[o]];
      const double2 uv = V.pm_uv[o];
      double pc[3], e[2], rho, w;
      to_camera(cam, X, pc);
      if (!(fabs(pc[2]) > 1e-12)) continue;
      residual(pc, V.K, uv.x, uv.y, e);
      huber(e[0] * e[0] + e[1] * e[1], rho, w);
      double Jc[2][6], Jp[2][3];
      jacobians(cam, pc, V.K, Jc, Jp);
      const double w0[3] = {w * Jp[0][0], w * Jp[0][1], w * Jp[0][2]};
      const double w1[3] = {w * Jp[1][0], w * Jp[1][1], w * Jp[1][2]};
      acc[0] += w0[0] * Jp[0][0] + w1[0] * Jp[1][0];
      acc[1] += w0[0] * Jp[0][1] + w1[0] * Jp[1][1];
      acc[2] += w0[0] * Jp[0][2] + w1[0] * Jp[1][2];
      acc[3] += w0[1] * Jp[0][1] + w1[1] * Jp[1][1];
      acc[4] += w0[1] * Jp[0][2] + w1[1] * Jp[1][2];
      acc[5] += w0[2] * Jp[0][2] + w1[2] * Jp[1][2];
#pragma unroll
      for (int j = 0; j < 3; ++j) acc[6 + j] -= w0[j] * e[0] + w1[j] * e[1];
    }
  }
#pragma unroll
  for (int j = 0; j < 9; ++j) {
    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 2);
  }
  double m = 0.0;
  if (col >= 0 && q == 0) {
#pragma unroll
    for (int j = 0; j < 6; ++j) Hll[6 * (size_t)col + j] = acc[j];
#pragma unroll
    for (int j = 0; j < 3; ++j) bl[3 * (size_t)col + j] = acc[6 + j];
    m = fmax(fabs(acc[0]), fmax(fabs(acc[3]), fabs(acc[5])));
  }
  // computeLambdaInit: max |diag Hll|; CTA max, then ONE order-independent atomic per CTA
  // (same-address atomics serialise in L2: one per warp cost more than the whole pass)
  __shared__ double smax[8];
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) m = fmax(m, smax[w]);
    if (m > 0.0) atomic_max_nonneg(maxdiag_bits, m);
  }
}

__global__ void ba_point_invert(const LMState* __restrict__ st, int Pf, const double* __restrict__ Hll,
                                double* __restrict__ Hinv) {
  if (st->done) return;
  const double lambda = st->lambda;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= Pf) return;
  double h[6], o[6];
#pragma unroll
  for (int j = 0; j < 6; ++j) h[j] = Hll[6 * (size_t)p + j];
  h[0] += lambda; h[3] += lambda; h[5] += lambda;
  sym3_inverse(h, o);
#pragma unroll
  for (int j = 0; j < 6; ++j) Hinv[6 * (size_t)p + j] = o[j];
}

// E = Hpl * Hinv per observation (camera-major) and the per-camera partial of sum E * bl.
__global__ void __launch_bounds__(kCamThreads)
ba_compute_E(BAView V, const double* __restrict__ Hpl, const double* __restrict__ Hinv,
             const double* __restrict__ bl, double* __restrict__ Eout, double* __restrict__ camp) {
  __shared__ double red[(kCamThreads / 32) * 6];
  if (V.st->done) return;
  const int c = blockIdx.y, chunk = blockIdx.x;
  const int start = V.cm_start[c] + chunk * kCamChunk;
  const int end = min(V.cm_start[c + 1], start + kCamChunk);
  if (start >= end || V.cam_col[c] < 0) return;
  double acc[6] = {0, 0, 0, 0, 0, 0};
  for (int o = start + threadIdx.x; o < end; o += kCamThreads) {
    const int pcol = V.pt_col[V.cm_pt[o]];
    if (pcol < 0) continue;
    double h[18], hi[6], b[3], e[18];
    load18(Hpl + (size_t)o * 18, h);
#pragma unroll
    for (int j = 0; j < 6; ++j) hi[j] = Hinv[6 * (size_t)pcol + j];
#pragma unroll
    for (int j = 0; j < 3; ++j) b[j] = bl[3 * (size_t)pcol + j];
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      sym3_mul(hi, &h[3 * r], &e[3 * r]);  // row r of Hpl times symmetric Hinv
      acc[r] += e[3 * r] * b[0] + e[3 * r + 1] * b[1] + e[3 * r + 2] * b[2];
    }
    store18(Eout + (size_t)o * 18, e);
  }
  block_reduce<6, kCamThreads>(acc, red);
  if (threadIdx.x == 0) {
    double* dst = camp + ((size_t)c * V.max_chunks + chunk) * kCamVals;
#pragma unroll
    for (int i = 0; i < 6; ++i) dst[i] = acc[i];
  }
}

// bhat = bp - sum_j E_j bl_j   (ordered chunk sum)
__global__ void ba_bhat_finalize(BAView V, const double* __restrict__ camp,
                                 const double* __restrict__ bp, double* __restrict__ bhat) {
  const int c = blockIdx.x, i = threadIdx.x;
  const int col = V.cam_col[c];
  if (V.st->done || i >= 6 || col < 0) return;
  const int n_obs = V.cm_start[c + 1] - V.cm_start[c];
  const int n_chunks = (n_obs + kCamChunk - 1) / kCamChunk;
  double s = 0.0;
  for (int k = 0; k < n_chunks; ++k) s += camp[((size_t)c * V.max_chunks + k) * kCamVals + i];
  bhat[6 * col + i] = bp[6 * col + i] - s;
}

// Schur work units: one warp reduces <= kUnitItems products E_a * Hpl_b^T of one (cam a, cam b) key.
__global__ void __launch_bounds__(128)
ba_schur_units(const LMState* __restrict__ st, const double* __restrict__ E, const double* __restrict__ Hpl,
               const int* __restrict__ item_a, const int* __restrict__ item_b,
               const int* __restrict__ unit_start, const int* __restrict__ unit_len, int n_units,
               double* __restrict__ partial) {
  const int u = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (st->done || u >= n_units) return;
  const int s = unit_start[u], len = unit_len[u];
  double acc[36];
#pragma unroll
  for (int i = 0; i < 36; ++i) acc[i] = 0.0;
  for (int i = lane; i < len; i += 32) {
    double ea[18], hb[18];
    load18(E + (size_t)item_a[s + i] * 18, ea);
    load18(Hpl + (size_t)item_b[s + i] * 18, hb);
#pragma unroll
    for (int r = 0; r < 6; ++r) {
#pragma unroll
      for (int c = 0; c < 6; ++c)
        acc[6 * r + c] += ea[3 * r] * hb[3 * c] + ea[3 * r + 1] * hb[3 * c + 1] +
                          ea[3 * r + 2] * hb[3 * c + 2];
    }
  }
#pragma unroll
  for (int i = 0; i < 36; ++i) acc[i] = warp_sum(acc[i]);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < 36; ++i) partial[(size_t)u * 36 + i] = acc[i];
  }
}

// S block (a,b) = [a==b](Hpp_a + lambda I) + scale-edge block - sum of its units; mirrored.
__global__ void ba_schur_finalize(const LMState* __restrict__ st, int n_seg, int n, const int* __restrict__ seg_a,
                                  const int* __restrict__ seg_b,
                                  const int* __restrict__ seg_unit_start,
                                  const double* __restrict__ partial,
                                  const double* __restrict__ Hpp,
                                  const double* __restrict__ extra,
                                  const int* __restrict__ seg_extra_idx, double* __restrict__ S) {
  if (st->done) return;
  const double lambda = st->lambda;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int seg = t / 36, e = t % 36;
  if (seg >= n_seg) return;
  const int a = seg_a[seg], b = seg_b[seg], r = e / 6, c = e % 6;
  const int ex = seg_extra_idx[seg];
  double v = ex >= 0 ? extra[(size_t)ex * 36 + e] : 0.0;
  if (a == b) v += Hpp[36 * a + e] + (r == c ? lambda : 0.0);
  double s = 0.0;
  for (int u = seg_unit_start[seg]; u < seg_unit_start[seg + 1]; ++u) s += partial[(size_t)u * 36 + e];
  v -= s;
  S[(size_t)(6 * a + r) * n + 6 * b + c] = v;
  if (a != b) S[(size_t)(6 * b + c) * n + 6 * a + r] = v;
}

// Back-substitution, camera-major half: v_o = Hpl_o^T dp_cam(o), scattered to the point-major slot.
__global__ void __launch_bounds__(kCamThreads)
ba_backsub_cam(BAView V, const double* __restrict__ Hpl, const double* __restrict__ dp,
               double* __restrict__ tmp3) {
  const int c = blockIdx.y, chunk = blockIdx.x;
  const int start = V.cm_start[c] + chunk * kCamChunk;
  const int end = min(V.cm_start[c + 1], start + kCamChunk);
  const int col = V.cam_col[c];
  if (V.st->done || start >= end || col < 0) return;
  double d[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) d[i] = dp[6 * col + i];
  for (int o = start + threadIdx.x; o < end; o += kCamThreads) {
    if (V.pt_col[V.cm_pt[o]] < 0) continue;
    double h[18];
    load18(Hpl + (size_t)o * 18, h);
    double v[3] = {0, 0, 0};
#pragma unroll
    for (int r = 0; r < 6; ++r) {
#pragma unroll
      for (int j = 0; j < 3; ++j) v[j] += h[3 * r + j] * d[r];
    }
    double* dst = tmp3 + (size_t)V.cm_pm[o] * 3;
    dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2];
  }
}

// Back-substitution, point half: dl = Hinv (bl - sum_o v_o); trial point = point + dl.
// Also the per-CTA partial of sum dl (lambda dl + bl)  (computeScale, landmark part).
__global__ void __launch_bounds__(256)
ba_backsub_pt(BAView V, const double* __restrict__ tmp3, const double* __restrict__ Hinv,
              const double* __restrict__ bl, double* __restrict__ dl_out,
              double* __restrict__ scale_partial) {
  __shared__ double red[8];
  if (V.st->done) return;
  const double lambda = V.st->lambda;
  const double* __restrict__ pts = V.pts[V.st->cur];
  double* __restrict__ pts_trial = V.pts[V.st->cur ^ 1];
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  double sc[1] = {0.0};
  if (p < V.P) {
    const int col = V.pt_col[p];
    double X[3] = {pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]};
    if (col >= 0) {
      double r[3] = {bl[3 * (size_t)col], bl[3 * (size_t)col + 1], bl[3 * (size_t)col + 2]};
      const double b0 = r[0], b1 = r[1], b2 = r[2];
      for (int o = V.pt_start[p]; o < V.pt_start[p + 1]; ++o) {
        if (V.cam_col[V.pm_cam[o]] < 0) continue;
        r[0] -= tmp3[3 * (size_t)o]; r[1] -= tmp3[3 * (size_t)o + 1]; r[2] -= tmp3[3 * (size_t)o + 2];
      }
      double hi[6], d[3];
#pragma unroll
      for (int j = 0; j < 6; ++j) hi[j] = Hinv[6 * (size_t)col + j];
      sym3_mul(hi, r, d);
      sc[0] = d[0] * (lambda * d[0] + b0) + d[1] * (lambda * d[1] + b1) + d[2] * (lambda * d[2] + b2);
      X[0] += d[0]; X[1] += d[1]; X[2] += d[2];
      dl_out[3 * (size_t)col] = d[0]; dl_out[3 * (size_t)col + 1] = d[1]; dl_out[3 * (size_t)col + 2] = d[2];
    }
    pts_trial[3 * p] = X[0]; pts_trial[3 * p + 1] = X[1]; pts_trial[3 * p + 2] = X[2];
  }
  block_reduce<1, 256>(sc, red);
  if (threadIdx.x == 0) scale_partial[blockIdx.x] = sc[0];
}

// SBACam::update on every free camera (t += d[0:3]; q <- normalize(q * (sqrt(1-|v|^2), v))).
__global__ void ba_update_cams(BAView V, const double* __restrict__ dp) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (V.st->done || c >= V.C) return;
  const double* __restrict__ cams = V.cams[V.st->cur];
  double* __restrict__ cams_trial = V.cams[V.st->cur ^ 1];
  CamRT* __restrict__ camrt_trial = V.camrt[V.st->cur ^ 1];
  double q[4] = {cams[7 * c], cams[7 * c + 1], cams[7 * c + 2], cams[7 * c + 3]};
  double t[3] = {cams[7 * c + 4], cams[7 * c + 5], cams[7 * c + 6]};
  const int col = V.cam_col[c];
  if (col >= 0) {
    const double* d = dp + 6 * col;
    t[0] += d[0]; t[1] += d[1]; t[2] += d[2];
    const double vx = d[3], vy = d[4], vz = d[5];
    const double vw = sqrt(1.0 - (vx * vx + vy * vy + vz * vz));
    const double x = q[3] * vx + q[0] * vw + q[1] * vz - q[2] * vy;
    const double y = q[3] * vy - q[0] * vz + q[1] * vw + q[2] * vx;
    const double z = q[3] * vz + q[0] * vy - q[1] * vx + q[2] * vw;
    const double w = q[3] * vw - q[0] * vx - q[1] * vy - q[2] * vz;
    const double inv = 1.0 / sqrt(x * x + y * y + z * z + w * w);
    q[0] = x * inv; q[1] = y * inv; q[2] = z * inv; q[3] = w * inv;
  }
  double* o = cams_trial + 7 * c;
  o[0] = q[0]; o[1] = q[1]; o[2] = q[2]; o[3] = q[3]; o[4] = t[0]; o[5] = t[1]; o[6] = t[2];
  CamRT r;
  quat_to_R(q, r.R);
  r.t[0] = t[0]; r.t[1] = t[1]; r.t[2] = t[2];
  camrt_trial[c] = r;
}

}  // namespace vslam
