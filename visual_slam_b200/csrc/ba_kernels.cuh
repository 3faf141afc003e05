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
constexpr int kPointLanes = 4;    // lanes cooperating on one point in the point role

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

// ---- fused linearisation ---------------------------------------------------------------------
// ONE launch builds the whole system: camera-role CTAs (chunks of kLinChunk observations of one
// camera, kLinObs per thread, next observation prefetched) produce the Hpl blocks and the per-chunk
// Hpp / bp / cost partials; point-role CTAs (a quad of lanes per point) produce Hll / bl.  The two
// roles are interleaved in the grid so that the write-heavy camera pass and the latency-bound point
// pass overlap.  The last chunk of a camera to finish (ticket counter) folds that camera's partials in
// chunk order; the last CTA of the grid runs the controller (cost, computeLambdaInit).  All sums are
// fixed-order trees -> bit-reproducible, no floating-point atomics.
constexpr int kLinThreads = 128;
constexpr int kLinObs = 8;
constexpr int kPtItemsPerWarp = 8;   // point-role work items (<= 32 observations each) per warp
constexpr int kCamSmemMax = 340;     // cameras kept in shared memory by the point role (96 B each)
constexpr int kLinChunk = kLinThreads * kLinObs;

struct LinArgs {
  int nA, nB;                       // camera-role / point-role CTAs
  const int* role;                  // [nA + nB]: >= 0 camera chunk index, < 0 point block -1 - b
  const int* wk_cam; const int* wk_start; const int* wk_chunk;   // per camera chunk
  const int* cam_chunk_off;         // [C + 1] first partial row of each camera
  // point role, regular items: 32 padded slots each (whole points, <= 32 observations in total)
  int n_pt_items; const int* px_pt; const int* px_cam; const double2* px_uv;   // [n_pt_items * 32], pt -1 = pad
  // points with more than 32 observations: (first point-major observation, count), strided path
  int n_long; const int2* long_item; const int* pm_pt;
  int cams_in_smem;
  unsigned* cam_ticket;             // [C], zero between launches
  unsigned* lin_ticket;             // [1]
  double* Hpl; double* camp; double* Hpp; double* bp; double* cam_cost; double* cam_maxd;
  double* Hll; double* bl; double* pt_maxd;
  const double* bp_extra; const int* diag_extra; const double* extra; const double* scale_cost;
  LMState* st_rw;
};

// Controller after buildSystem(): total cost of the current state, computeLambdaInit on the first
// linearisation (lambda = tau * max |diag H|).  Runs in the last CTA of the fused kernel.
__device__ __forceinline__ void lin_controller(const BAView& V, const LinArgs& A, double* red /*[8]*/) {
  double v[2] = {0.0, 0.0};
  for (int i = threadIdx.x; i < V.C; i += kLinThreads) {
    v[0] += __ldcg(A.cam_cost + i);
    v[1] = fmax(v[1], __ldcg(A.cam_maxd + i));
  }
  for (int i = threadIdx.x; i < A.nB; i += kLinThreads) v[1] = fmax(v[1], __ldcg(A.pt_maxd + i));
  v[0] = warp_sum(v[0]);
  for (int o = 16; o > 0; o >>= 1) v[1] = fmax(v[1], __shfl_xor_sync(0xffffffffu, v[1], o));
  if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = v[0]; red[4 + (threadIdx.x >> 5)] = v[1]; }
  __syncthreads();
  if (threadIdx.x == 0) {
    LMState* st = A.st_rw;
    const double cost = ((red[0] + red[1]) + (red[2] + red[3])) + __ldcg(A.scale_cost);
    const double md = fmax(fmax(red[4], red[5]), fmax(red[6], red[7]));
    st->cur_cost = cost;
    st->max_diag = md;
    if (st->first && !st->manual) {
      st->initial_cost = cost;
      st->lambda = 1e-5 * md;   // tau (OptimizationAlgorithmLevenberg default)
      st->ni = 2.0;
      st->first = 0;
    }
    st->dirty = 0;
  }
}

template <bool COST_ONLY>
__global__ void __launch_bounds__(kLinThreads, 4)
ba_linearize_kernel(BAView V, LinArgs A) {
  __shared__ __align__(16) double stage[COST_ONLY ? 2 : (kLinThreads / 32) * 32 * 18];
  __shared__ double red[(kLinThreads / 32) * 32];
  __shared__ CamRT s_cam;
  __shared__ int s_flag;
  if (V.st->done || (!COST_ONLY && !V.st->dirty)) return;
  const int sb = COST_ONLY ? (V.st->cur ^ 1) : V.st->cur;   // the cost-only pass evaluates the trial state
  const CamRT* __restrict__ camrt = V.camrt[sb];
  const double* __restrict__ pts = V.pts[sb];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int role = COST_ONLY ? (int)blockIdx.x : A.role[blockIdx.x];
  bool take_ticket = false;

  if (role >= 0) {
    // ------------------------------------------------ camera role
    const int c = A.wk_cam[role], chunk = A.wk_chunk[role];
    const int start = A.wk_start[role];
    const int end = min(V.cm_start[c + 1], start + kLinChunk);
    if (tid < 12) reinterpret_cast<double*>(&s_cam)[tid] = reinterpret_cast<const double*>(camrt + c)[tid];
    const bool cam_free = V.cam_col[c] >= 0;
    double v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = 0.0;
    // two-stage software pipeline (warps issue in order, so a dependent address must never be
    // computed right after its index load): the point INDEX is loaded two observations ahead, the
    // point itself one observation ahead, and both are in flight under the current arithmetic.
    int o = start + tid;
    double X[3] = {0.0, 0.0, 1.0};
    double2 uv = make_double2(0.0, 0.0);
    int p_next = 0;
    if (o + kLinThreads < end) p_next = V.cm_pt[o + kLinThreads];
    if (o < end) {
      const int p = V.cm_pt[o];
      uv = V.cm_uv[o];
      X[0] = pts[3 * p]; X[1] = pts[3 * p + 1]; X[2] = pts[3 * p + 2];
    }
    __syncthreads();
#pragma unroll 1
    for (int k = 0; k < kLinObs; ++k) {
      const bool valid = o < end;
      const double Xc[3] = {X[0], X[1], X[2]};
      const double2 uvc = uv;
      const int o_next = o + kLinThreads;
      const int p_cur = p_next;
      if (k + 2 < kLinObs && o_next + kLinThreads < end) p_next = V.cm_pt[o_next + kLinThreads];
      if (k + 1 < kLinObs && o_next < end) {   // p_cur arrived during the previous observation
        uv = V.cm_uv[o_next];
        X[0] = pts[3 * p_cur]; X[1] = pts[3 * p_cur + 1]; X[2] = pts[3 * p_cur + 2];
      }
      double h[18];
#pragma unroll
      for (int i = 0; i < 18; ++i) h[i] = 0.0;
      if (valid) {
        double pc[3], e[2], rho, w;
        to_camera(s_cam, Xc, pc);
        if (fabs(pc[2]) > 1e-12) {  // g2o aborts on z = 0; we drop the observation
          residual(pc, V.K, uvc.x, uvc.y, e);
          huber(e[0] * e[0] + e[1] * e[1], rho, w);
          v[27] += rho;
          if (!COST_ONLY && cam_free) {
            double Jc[2][6], Jp[2][3], W0[6], W1[6];
            jacobians(s_cam, pc, V.K, Jc, Jp);
#pragma unroll
            for (int i = 0; i < 6; ++i) { W0[i] = w * Jc[0][i]; W1[i] = w * Jc[1][i]; }
            int q = 0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
#pragma unroll
              for (int j = i; j < 6; ++j) { v[q] += W0[i] * Jc[0][j] + W1[i] * Jc[1][j]; ++q; }
            }
#pragma unroll
            for (int i = 0; i < 6; ++i) v[21 + i] -= W0[i] * e[0] + W1[i] * e[1];
#pragma unroll
            for (int i = 0; i < 6; ++i) {
#pragma unroll
              for (int j = 0; j < 3; ++j) h[3 * i + j] = W0[i] * Jp[0][j] + W1[i] * Jp[1][j];
            }
          }
        }
      }
      if (!COST_ONLY && cam_free) {
        // the warp's 32 blocks are contiguous in HBM (camera-major): stage them in the warp's own
        // slice of shared memory (144 B stride: conflict-free STS.128), write with coalesced STG.128
        double* ws = stage + wid * (32 * 18);
        const int base = o - lane;                       // first observation of this warp's row
        const int n_blk = max(0, min(32, end - base));
        __syncwarp();
        store18(ws + lane * 18, h);
        __syncwarp();
        double2* dst = reinterpret_cast<double2*>(A.Hpl + (size_t)base * 18);
        const double2* src = reinterpret_cast<const double2*>(ws);
        for (int i = lane; i < n_blk * 9; i += 32) dst[i] = src[i];
      }
      o = o_next;
    }
    // chunk partial: 28 values
    if (COST_ONLY) {
      double s = warp_sum(v[27]);
      if (lane == 0) red[wid] = s;
      __syncthreads();
      if (tid == 0) A.camp[(size_t)(A.cam_chunk_off[c] + chunk) * kCamVals + 27] = (red[0] + red[1]) + (red[2] + red[3]);
    } else {
      warp_transpose_sum32(v, lane);
      red[wid * 32 + lane] = v[0];
      __syncthreads();
      if (tid < kCamVals)
        A.camp[(size_t)(A.cam_chunk_off[c] + chunk) * kCamVals + tid] =
            (red[tid] + red[32 + tid]) + (red[64 + tid] + red[96 + tid]);
    }
    // last chunk of this camera folds the partials (chunk order -> deterministic)
    __threadfence();
    __syncthreads();
    const int n_chunks = A.cam_chunk_off[c + 1] - A.cam_chunk_off[c];
    if (tid == 0) {
      const unsigned t = atomicAdd(A.cam_ticket + c, 1u);
      s_flag = (t == (unsigned)(n_chunks - 1));
      if (s_flag) A.cam_ticket[c] = 0u;
    }
    __syncthreads();
    if (s_flag) {
      __threadfence();
      if (tid < kCamVals) {
        double s = 0.0;
        const double* base = A.camp + (size_t)A.cam_chunk_off[c] * kCamVals + tid;
        for (int k = 0; k < n_chunks; ++k) s += __ldcg(base + (size_t)k * kCamVals);
        red[tid] = s;
      }
      __syncthreads();
      if (tid == 27) A.cam_cost[c] = red[27];
      if (!COST_ONLY) {
        const int col = V.cam_col[c];
        if (col >= 0 && tid < 27) {
          if (tid >= 21) {
            A.bp[6 * col + (tid - 21)] = red[tid] + A.bp_extra[6 * col + (tid - 21)];
          } else {
            int r = 0, k = tid;
            while (k >= 6 - r) { k -= 6 - r; ++r; }
            const int cc = r + k;
            A.Hpp[36 * col + 6 * r + cc] = red[tid];
            A.Hpp[36 * col + 6 * cc + r] = red[tid];
          }
        }
        if (tid == 0) {   // computeLambdaInit: diagonal of the full camera block (projection + scale edges)
          double m = 0.0;
          if (col >= 0) {
            const int ex = A.diag_extra[col];
            const int dpos[6] = {0, 6, 11, 15, 18, 20};
            for (int r = 0; r < 6; ++r)
              m = fmax(m, fabs(red[dpos[r]] + (ex >= 0 ? A.extra[36 * (size_t)ex + 7 * r] : 0.0)));
          }
          A.cam_maxd[c] = m;
        }
      }
      take_ticket = !COST_ONLY;
    }
  } else {
    // ------------------------------------------------ point role (never reached when COST_ONLY)
    // One lane per observation (point-major order): a warp item is a run of whole points with at
    // most 32 observations (or one long point, strided); per-point sums by a segmented scan over the
    // lanes -- fixed shape, deterministic.  Cameras come from shared memory when they fit.
    extern __shared__ __align__(16) double s_cams[];   // [C][12] when A.cams_in_smem
    const int b = -1 - role;
    if (A.cams_in_smem) {
      const double2* src = reinterpret_cast<const double2*>(camrt);
      double2* dst = reinterpret_cast<double2*>(s_cams);
      for (int i = tid; i < V.C * 6; i += kLinThreads) dst[i] = src[i];
      __syncthreads();
    }
    const CamRT* __restrict__ cam_tab = A.cams_in_smem ? reinterpret_cast<const CamRT*>(s_cams) : camrt;
    double m = 0.0;
    auto accumulate = [&](const CamRT& cam, const double* X, double2 uv, double* acc) {
      double pc[3], e[2], rho, wgt;
      to_camera(cam, X, pc);
      if (!(fabs(pc[2]) > 1e-12)) return;
      residual(pc, V.K, uv.x, uv.y, e);
      huber(e[0] * e[0] + e[1] * e[1], rho, wgt);
      double Jc[2][6], Jp[2][3];
      jacobians(cam, pc, V.K, Jc, Jp);
      const double w0[3] = {wgt * Jp[0][0], wgt * Jp[0][1], wgt * Jp[0][2]};
      const double w1[3] = {wgt * Jp[1][0], wgt * Jp[1][1], wgt * Jp[1][2]};
      acc[0] += w0[0] * Jp[0][0] + w1[0] * Jp[1][0];
      acc[1] += w0[0] * Jp[0][1] + w1[0] * Jp[1][1];
      acc[2] += w0[0] * Jp[0][2] + w1[0] * Jp[1][2];
      acc[3] += w0[1] * Jp[0][1] + w1[1] * Jp[1][1];
      acc[4] += w0[1] * Jp[0][2] + w1[1] * Jp[1][2];
      acc[5] += w0[2] * Jp[0][2] + w1[2] * Jp[1][2];
#pragma unroll
      for (int j = 0; j < 3; ++j) acc[6 + j] -= w0[j] * e[0] + w1[j] * e[1];
    };
    auto store_point = [&](int key, const double* acc) {
      const int col = V.pt_col[key];
#pragma unroll
      for (int j = 0; j < 6; ++j) A.Hll[6 * (size_t)col + j] = acc[j];
#pragma unroll
      for (int j = 0; j < 3; ++j) A.bl[3 * (size_t)col + j] = acc[6 + j];
      m = fmax(m, fmax(fabs(acc[0]), fmax(fabs(acc[3]), fabs(acc[5]))));
    };
    // regular items, software-pipelined like the camera role: slot indices two items ahead, the
    // point coordinates one item ahead (slots are at fixed addresses: no header load in the chain)
    const int item0 = (b * (kLinThreads / 32) + wid) * kPtItemsPerWarp;
    const int n_it = max(0, min(kPtItemsPerWarp, A.n_pt_items - item0));
    int p_c = -1, ci_c = 0, p_n = -1, ci_n = 0;
    double2 uv_c = make_double2(0.0, 0.0), uv_n = uv_c;
    double Xc[3] = {0.0, 0.0, 1.0};
    if (n_it > 0) {
      const size_t s0 = (size_t)item0 * 32 + lane;
      p_c = A.px_pt[s0]; ci_c = A.px_cam[s0]; uv_c = A.px_uv[s0];
      if (n_it > 1) { p_n = A.px_pt[s0 + 32]; ci_n = A.px_cam[s0 + 32]; uv_n = A.px_uv[s0 + 32]; }
      if (p_c >= 0) { Xc[0] = pts[3 * p_c]; Xc[1] = pts[3 * p_c + 1]; Xc[2] = pts[3 * p_c + 2]; }
    }
#pragma unroll 1
    for (int it = 0; it < n_it; ++it) {
      // issue: coordinates of item it+1 (its indices arrived during the previous item), indices of it+2
      double Xn[3] = {0.0, 0.0, 1.0};
      if (p_n >= 0) { Xn[0] = pts[3 * p_n]; Xn[1] = pts[3 * p_n + 1]; Xn[2] = pts[3 * p_n + 2]; }
      int p_2 = -1, ci_2 = 0;
      double2 uv_2 = make_double2(0.0, 0.0);
      if (it + 2 < n_it) {
        const size_t s2 = (size_t)(item0 + it + 2) * 32 + lane;
        p_2 = A.px_pt[s2]; ci_2 = A.px_cam[s2]; uv_2 = A.px_uv[s2];
      }
      double acc[9];
#pragma unroll
      for (int j = 0; j < 9; ++j) acc[j] = 0.0;
      const int key = p_c;
      if (key >= 0) accumulate(cam_tab[ci_c], Xc, uv_c, acc);
      // inclusive segmented scan; the last lane of a segment holds the point's sums
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int ku = __shfl_up_sync(0xffffffffu, key, d);
        const bool take = (lane >= d) && (ku == key) && (key >= 0);
#pragma unroll
        for (int j = 0; j < 9; ++j) {
          const double u = __shfl_up_sync(0xffffffffu, acc[j], d);
          if (take) acc[j] += u;
        }
      }
      const int kn = __shfl_down_sync(0xffffffffu, key, 1);
      if ((key >= 0) && (lane == 31 || kn != key)) store_point(key, acc);
      p_c = p_n; ci_c = ci_n; uv_c = uv_n; Xc[0] = Xn[0]; Xc[1] = Xn[1]; Xc[2] = Xn[2];
      p_n = p_2; ci_n = ci_2; uv_n = uv_2;
    }
    // long points (> 32 observations): one warp each, strided, plain butterfly
    for (int li = b * (kLinThreads / 32) + wid; li < A.n_long; li += A.nB * (kLinThreads / 32)) {
      const int2 w = A.long_item[li];
      double acc[9];
#pragma unroll
      for (int j = 0; j < 9; ++j) acc[j] = 0.0;
      const int key = A.pm_pt[w.x];
      const double X[3] = {pts[3 * key], pts[3 * key + 1], pts[3 * key + 2]};
      for (int o = w.x + lane; o < w.x + w.y; o += 32) accumulate(cam_tab[V.pm_cam[o]], X, V.pm_uv[o], acc);
#pragma unroll
      for (int j = 0; j < 9; ++j) acc[j] = warp_sum(acc[j]);
      if (lane == 0) store_point(key, acc);
    }
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) red[wid] = m;
    __syncthreads();
    if (tid == 0) A.pt_maxd[b] = fmax(fmax(red[0], red[1]), fmax(red[2], red[3]));
    take_ticket = true;
  }

  if (!COST_ONLY) {
    // the last CTA of the grid (camera finalisers + point blocks) runs the controller
    if (!take_ticket) return;     // uniform per CTA
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      const unsigned t = atomicAdd(A.lin_ticket, 1u);
      s_flag = (t == (unsigned)(V.C + A.nB - 1));
      if (s_flag) *A.lin_ticket = 0u;
    }
    __syncthreads();
    if (s_flag) {
      __threadfence();
      lin_controller(V, A, red);
    }
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
