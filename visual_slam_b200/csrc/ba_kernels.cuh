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
constexpr int kUnitItemsSmall = 64;    // ... when it is a local window
constexpr int kUnitItemsLarge = 352;   // ... when the problem is large enough to keep every warp busy with those
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

// ---- scale edges (EdgeSBAScale + DCS, LocalBA.py:115-131) -----------------------------------
// One CTA: (1) thread per edge evaluates residual / weight / direction, (2) thread per extra S block
// and per free camera gathers the contributions of its incident edges in edge order (CSR built on the
// host) -> deterministic.
struct ScaleView {
  int E, n_extra, Cf;
  const LMState* st; const double* cams2[2];
  const int* sa; const int* sb; const double* meas; const int* cam_col;
  const int* inc_start; const int* inc_edge;   // per free camera: incident edge ids (ascending)
  const int* extra_a; const int* extra_b;      // free-camera columns of each extra block (a <= b)
  double* edge_w; double* edge_es; double* edge_u;  // [E], [E], [E][3]
  double* extra; double* bp_extra; double* scale_cost;
  double* extra_diag;                          // [Cf][3] translation diagonal of each camera's own extra block
};

template <int THREADS>
__device__ __forceinline__ void scale_edges_body(const ScaleView& V, bool cost_only, double* red /*[THREADS/32]*/,
                                                 bool trial_state = false) {
  const double* __restrict__ cams = V.cams2[(cost_only || trial_state) ? (V.st->cur ^ 1) : V.st->cur];
  double c[1] = {0.0};
  for (int k = threadIdx.x; k < V.E; k += THREADS) {
    const double* ta = cams + 7 * V.sa[k] + 4;
    const double* tb = cams + 7 * V.sb[k] + 4;
    const double d[3] = {tb[0] - ta[0], tb[1] - ta[1], tb[2] - ta[2]};
    const double nrm = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    const double es = V.meas[k] - nrm;
    double rho, w;
    dcs(es * es, rho, w);
    c[0] += rho;
    if (!cost_only) {
      V.edge_w[k] = w; V.edge_es[k] = es;
      V.edge_u[3 * k] = d[0] / nrm; V.edge_u[3 * k + 1] = d[1] / nrm; V.edge_u[3 * k + 2] = d[2] / nrm;
    }
  }
  block_reduce<1, THREADS>(c, red);
  if (threadIdx.x == 0) *V.scale_cost = c[0];
  if (cost_only) return;
  __threadfence_block();
  __syncthreads();
  for (int x = threadIdx.x; x < V.n_extra; x += THREADS) {
    const int a = V.extra_a[x], b = V.extra_b[x];
    double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int q = V.inc_start[a]; q < V.inc_start[a + 1]; ++q) {
      const int k = V.inc_edge[q];
      const int ca = V.cam_col[V.sa[k]], cb = V.cam_col[V.sb[k]];
      double sgn;
      if (a == b) sgn = 1.0;                                         // diagonal: every incident edge
      else if ((ca == a && cb == b) || (ca == b && cb == a)) sgn = -1.0;  // off-diagonal: edges a-b
      else continue;
      const double w = V.edge_w[k];
      const double* u = V.edge_u + 3 * k;
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) h[3 * i + j] += sgn * w * u[i] * u[j];
    }
    double* o = V.extra + 36 * (size_t)x;
    for (int i = 0; i < 36; ++i) o[i] = 0.0;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) o[6 * i + j] = h[3 * i + j];
    if (a == b) { V.extra_diag[3 * a] = h[0]; V.extra_diag[3 * a + 1] = h[4]; V.extra_diag[3 * a + 2] = h[8]; }
  }
  for (int col = threadIdx.x; col < V.Cf; col += THREADS) {
    double g[3] = {0, 0, 0};
    for (int q = V.inc_start[col]; q < V.inc_start[col + 1]; ++q) {
      const int k = V.inc_edge[q];
      // d e / d t_a = +u, d e / d t_b = -u ;  b += -w J e
      const double sgn = (V.cam_col[V.sa[k]] == col) ? -1.0 : 1.0;
      const double f = sgn * V.edge_w[k] * V.edge_es[k];
      g[0] += f * V.edge_u[3 * k]; g[1] += f * V.edge_u[3 * k + 1]; g[2] += f * V.edge_u[3 * k + 2];
    }
    double* o = V.bp_extra + 6 * col;
    o[0] = g[0]; o[1] = g[1]; o[2] = g[2]; o[3] = 0.0; o[4] = 0.0; o[5] = 0.0;
  }
}

__global__ void __launch_bounds__(256) ba_scale_edges(ScaleView V, int cost_only) {
  __shared__ double red[8];
  if (V.st->done || (!cost_only && !V.st->dirty)) return;
  scale_edges_body<256>(V, cost_only != 0, red);
}

// ---- fused linearisation ---------------------------------------------------------------------
// ONE launch builds the whole system.  The grid mixes three kinds of CTA (work descriptors built on
// the host, interleaved so that all kinds are resident together):
//   camera chunks  kLinChunk observations of one camera in camera-major order, kLinObs per thread, the
//                  next observation prefetched: the 6x3 Hpl blocks (staged per warp in shared memory,
//                  written with ONE bulk async store per 32 blocks) and the chunk's Hpp / bp / cost
//                  partial (28 values, transposed warp reduction);
//   point blocks   one LANE per point over an ELL layout (points sorted by degree, 32 per item, row j
//                  of an item = the j-th observation of each of its points, coalesced): Hll / bl are
//                  plain per-lane sums, no cross-lane traffic; cameras come from shared memory;
//   scale edges    one CTA (EdgeSBAScale blocks).
// Every observation is evaluated once per role; the translation half of the camera Jacobian is -Jp,
// so Hpp_tt, Hpl_t and Hll share the six values of w Jp^T Jp.  The last chunk of a camera to finish
// (ticket counter) folds that camera's partials in chunk order; the last CTA of the grid runs the
// controller (scale-edge terms, cost, computeLambdaInit).  All sums are fixed-order -> bit-reproducible,
// no floating-point atomics.
constexpr int kLinThreads = 128;
constexpr int kLinWarps = kLinThreads / 32;
constexpr int kLinObs = 8;
constexpr int kLinChunk = kLinThreads * kLinObs;
constexpr int kStageDoubles = 32 * 18;          // one warp's 32 Hpl blocks (4608 B)
constexpr bool kBulkStore = true;               // Hpl stage -> HBM by bulk async store (else LDS + STG.128)
constexpr int kHplStages = 1;                   // Hpl stages per warp
constexpr int kLinInSmem = kLinObs * kLinThreads * (4 + 16);              // chunk inputs: indices + measurements
constexpr int kLinCamSmem = kLinWarps * kHplStages * kStageDoubles * 8 + kLinInSmem;   // camera-role bytes
constexpr int kLinCostSmem = kLinCamSmem;        // same layout in the cost-only launch (the stage stays unused)   // camera-role bytes
constexpr int kPtRowsPerWarp = 8;               // ELL rows a point-role warp aims for
constexpr int kCamTabStride = 14;               // doubles per camera in the point role's table (112 B)
constexpr int kCamSmemMax = 1800;               // cameras kept in shared memory by the point role

struct __align__(16) LinWork {   // one per CTA of the fused grid
  int kind;        // 0 camera chunk, 1 point block, 2 scale edges
  int cam;         // camera (kind 0) / point block index (kind 1)
  int start, end;  // camera-major observation range of the chunk
  int row;         // row of the chunk's partial in camp
  int n_chunks;    // chunks of this camera
  int first_row;   // first partial row of this camera
  int col;         // free-camera column or -1
};

struct TrialArgs {   // what the LM controller needs besides the trial cost (cost-only launch)
  int n_sp, n;
  const double* scale_partial; const double* dp; const double* bp; const double* bp_extra;
  const int* pcg_status; const int* pcg_iters;
};

struct LinArgs {
  TrialArgs trial;
  const LinWork* work;              // fused grid: [n_ctas]
  const LinWork* cam_work;          // camera chunks only (cost-only grid)
  int n_pt_ctas, n_pt_warps, n_tickets, cams_in_smem, has_scale;
  // point role (ELL): warp w owns items [pw_first[w], pw_first[w+1]); item it owns rows
  // [item_row[it], item_row[it+1]) of 32 slots; slot = (camera or -1, measurement)
  const int* pw_first; const int* item_row; const int* ell_pt; const int* ell_cam; const double2* ell_uv;
  unsigned* cam_ticket;             // [C], zero between launches
  unsigned* lin_ticket;             // [1]
  double* Hpl; double* camp; double* Hpp; double* bp; double* cam_cost;
  double* Hll; double* bl; double* pt_maxd;
  const double* extra_diag; const double* scale_cost;   // scale-edge terms (bp_extra is added by the consumers of bp)
  LMState* st_rw;
  long long* trace;                 // optional [n_ctas][2] globaltimer stamps (profiling), else null
  int force;                        // profiling: rebuild even when the system is not dirty
  int spec;                         // speculative: build the system of the TRIAL state and let the LM controller decide
};

__device__ __forceinline__ long long globaltimer_ns();
__global__ void ba_stamp(long long* dst);
__device__ __forceinline__ long long globaltimer_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void ba_stamp(long long* dst) { *dst = globaltimer_ns(); }

// bulk async copy shared -> global (SASS UBLKCP), bulk-group completion
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void cp_async4(void* dst_smem, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst_smem, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// One point = one 32-byte sector of the padded state array [P][4]: a single 256-bit load (LDG.E.256)
// instead of three 8-byte loads that each cost a pass through the L1 data pipe per distinct line.
__device__ __forceinline__ void load_point(const double* __restrict__ pts4, int p, double (&X)[3]) {
  double pad;
  asm volatile("ld.global.v4.f64 {%0, %1, %2, %3}, [%4];"
               : "=d"(X[0]), "=d"(X[1]), "=d"(X[2]), "=d"(pad) : "l"(pts4 + 4 * (size_t)p));
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// Programmatic dependent launch: every kernel of an LM step is launched with the programmatic-stream-
// serialization attribute, so its CTAs may become resident while the previous kernel is still draining.
// pdl_wait() returns once the previous grid has completed and its writes are visible (a no-op for a plain
// launch) -- nothing another kernel of the step writes may be read before it; pdl_trigger() lets the NEXT
// kernel be scheduled.  Waiting first keeps the look-ahead at one kernel.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Branch-free FP64 reciprocal / reciprocal square root for finite normal arguments: hardware seed
// (MUFU.RCP64H / RSQ64H, ~2^-20) + two Newton steps (error ~1 ulp).  The IEEE division / sqrt of the
// compiler expands to the same seed plus a slow-path call that costs branches in every warp.
__device__ __forceinline__ double fast_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}
__device__ __forceinline__ double fast_rsqrt(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double h = 0.5 * x;
  double e = fma(-h * y, y, 0.5);
  y = fma(y, e, y);
  e = fma(-h * y, y, 0.5);
  return fma(y, e, y);
}

// Residual, Huber weight and the reduced Jacobians of one observation.  Returns false when the point
// is on the camera plane (g2o aborts there; we drop the observation).
//   Jp[2][3] = d e / d X (= -translation columns of the camera Jacobian), Jr[2][3] = rotation columns.
__device__ __forceinline__ bool eval_obs(const double* __restrict__ R, const double* __restrict__ t,
                                         const double* X, double2 uv, const double4 K, bool want_jr,
                                         double* e, double& rho, double& w, double (&Jp)[2][3],
                                         double (&Jr)[2][3]) {
  const double dx = X[0] - t[0], dy = X[1] - t[1], dz = X[2] - t[2];
  const double px = R[0] * dx + R[3] * dy + R[6] * dz;
  const double py = R[1] * dx + R[4] * dy + R[7] * dz;
  const double pz = R[2] * dx + R[5] * dy + R[8] * dz;
  if (!(fabs(pz) > 1e-12)) return false;
  const double iz = fast_rcp(pz);
  e[0] = K.x * px * iz + K.z - uv.x;
  e[1] = K.y * py * iz + K.w - uv.y;
  const double chi2 = e[0] * e[0] + e[1] * e[1];
  // Huber (robust_kernel_impl): rho = chi2, w = 1 inside; rho = 2 d sqrt(chi2) - d^2, w = d / sqrt(chi2)
  const double d = sqrt(kHuberDelta2);
  const double r = fast_rsqrt(fmax(chi2, kHuberDelta2));
  const bool inl = chi2 <= kHuberDelta2;
  w = inl ? 1.0 : d * r;
  rho = inl ? chi2 : 2.0 * d * (chi2 * r) - kHuberDelta2;
  const double ipz2 = iz * iz;
  const double a = ipz2 * K.x, b = ipz2 * K.y;
#pragma unroll
  for (int k = 0; k < 3; ++k) {   // w2n column k = row k of R
    Jp[0][k] = (pz * R[3 * k + 0] - px * R[3 * k + 2]) * a;
    Jp[1][k] = (pz * R[3 * k + 1] - py * R[3 * k + 2]) * b;
  }
  if (want_jr) {
    // dR/dk (X - t) = A_k pc: A_x pc = (0, 2pz, -2py); A_y pc = (-2pz, 0, 2px); A_z pc = (2py, -2px, 0)
    const double a2 = 2.0 * a, b2 = 2.0 * b, pz2 = pz * pz;
    Jr[0][0] = px * py * a2;           Jr[1][0] = (pz2 + py * py) * b2;
    Jr[0][1] = -(pz2 + px * px) * a2;  Jr[1][1] = -(py * px) * b2;
    Jr[0][2] = pz * py * a2;           Jr[1][2] = -(pz * px) * b2;
  }
  return true;
}

// Controller after buildSystem(): scale-edge terms of bp, total cost of the current state,
// computeLambdaInit on the first linearisation (lambda = tau * max |diag H|).  Last CTA of the grid.
__device__ __forceinline__ void lin_controller(const BAView& V, const LinArgs& A, double* red /*[8]*/) {
  // read-only loops first (independent loads pipeline freely), the bp update last
  if (A.trace && threadIdx.x == 0) A.trace[2 * gridDim.x + 4] = globaltimer_ns();
  // every load of a batch is issued before the first use: the values were just written by other SMs
  // and each L2 round trip costs ~0.5 us, so the loops must not serialise them
  double v[2] = {0.0, 0.0};
  constexpr int U = 10;
  const int n6 = 6 * V.Cf;
  const double sc = A.has_scale ? __ldcg(A.scale_cost) : 0.0;
  const int n_max = max(max(V.C, n6), A.n_pt_ctas);
  for (int i0 = 0; i0 < n_max; i0 += U * kLinThreads) {
    double a[U], x[U], cc[U], pm[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = i0 + u * kLinThreads + threadIdx.x;
      const int col = i / 6, r = i - 6 * col;
      a[u] = i < n6 ? __ldcg(A.Hpp + 36 * (size_t)col + 7 * r) : 0.0;
      x[u] = (i < n6 && A.has_scale && r < 3) ? __ldcg(A.extra_diag + 3 * col + r) : 0.0;
      cc[u] = i < V.C ? __ldcg(A.cam_cost + i) : 0.0;
      pm[u] = i < A.n_pt_ctas ? __ldcg(A.pt_maxd + i) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      v[0] += cc[u];
      v[1] = fmax(v[1], fmax(fabs(a[u] + x[u]), pm[u]));
    }
  }
  if (A.trace && threadIdx.x == 0) A.trace[2 * gridDim.x + 5] = globaltimer_ns() + (v[0] > 1e300) + (v[1] > 1e300);
  if (A.trace && threadIdx.x == 0) A.trace[2 * gridDim.x + 6] = globaltimer_ns();
  v[0] = warp_sum(v[0]);
  for (int o = 16; o > 0; o >>= 1) v[1] = fmax(v[1], __shfl_xor_sync(0xffffffffu, v[1], o));
  if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = v[0]; red[4 + (threadIdx.x >> 5)] = v[1]; }
  __syncthreads();
  if (threadIdx.x == 0) {
    LMState* st = A.st_rw;
    const double cost = ((red[0] + red[1]) + (red[2] + red[3])) + sc;
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

// The Levenberg-Marquardt accept / reject logic of g2o (OptimizationAlgorithmLevenberg::solve), one thread.  tmp = cost
// of the trial state, scale_sum = sum dx (lambda dx + b).  system_of_trial: the system in memory was built at the trial
// state (speculative linearisation), max_diag is its largest diagonal entry.
__device__ __forceinline__ void lm_decide(LMState* st, double tmp, double scale_sum, int status, int its,
                                          bool system_of_trial, double max_diag) {
  const double lambda = st->lambda;
  st->last_cost = tmp; st->last_scale = scale_sum; st->last_pcg_status = status; st->last_pcg_iters = its;
  if (st->manual) return;
  const bool ok = status != 1;
  st->pcg_total += (double)its;
  if (!ok) { tmp = 1.7976931348623157e308; st->breakdowns += 1.0; }
  const double scale = scale_sum + 1e-3;
  const double rho = (st->cur_cost - tmp) / scale;
  st->last_rho = rho;
  st->trials += 1;
  const bool accepted = rho > 0.0 && isfinite(tmp);
  if (accepted) {
    double alpha = 1.0 - pow(2.0 * rho - 1.0, 3.0);
    alpha = fmin(alpha, 2.0 / 3.0);
    st->lambda = lambda * fmax(1.0 / 3.0, alpha);
    st->ni = 2.0;
    st->cur_cost = tmp;
    st->cur ^= 1;          // discardTop(): the trial state becomes current
    st->dirty = system_of_trial ? 0 : 1;
    if (system_of_trial) st->max_diag = max_diag;
  } else {
    st->lambda = lambda * st->ni;   // pop(): current state untouched; the system stays valid unless it was rebuilt
    st->ni *= 2.0;                  // at the trial state
    st->rejected += 1.0;
    if (system_of_trial) st->dirty = 1;
  }
  const bool lam_bad = !accepted && !isfinite(st->lambda);
  int q = st->q;
  if (!lam_bad) q += 1;
  if (!lam_bad && rho < 0.0 && q < 10) { st->q = q; return; }   // next trial of the same outer iteration
  st->iters += 1;
  const bool stop = (q == 10) || rho == 0.0 || !isfinite(st->lambda);
  st->q = 0;
  if (!accepted) st->dirty = 1;   // g2o rebuilds the system at the start of every outer iteration
  if (stop || st->iters >= st->max_iterations) st->done = 1;
}

// Controller after a trial (last CTA of the cost-only launch): cost of the trial state, computeScale,
// then the Levenberg-Marquardt accept / reject logic of g2o (OptimizationAlgorithmLevenberg::solve)
// entirely on the device.
__device__ __forceinline__ void trial_controller(const BAView& V, const LinArgs& A, const ScaleView& SV,
                                                 double* red /*[2 * kLinWarps]*/) {
  LMState* st = A.st_rw;
  const TrialArgs& T = A.trial;
  // every load that does not depend on the scale edges is issued first, so that its L2 round trip runs under theirs
  double v[2] = {0.0, 0.0};
  for (int i = threadIdx.x; i < V.C; i += kLinThreads) v[0] += __ldcg(A.cam_cost + i);
  for (int i = threadIdx.x; i < T.n_sp; i += kLinThreads) v[1] += __ldcg(T.scale_partial + i);   // points, then cameras
  int status = 0, its = 0;
  if (threadIdx.x == 0 && T.n) { status = __ldcg(T.pcg_status); its = __ldcg(T.pcg_iters); }
  if (A.has_scale) {   // cost of the scale edges at the trial state (EdgeSBAScale + DCS)
    scale_edges_body<kLinThreads>(SV, true, red);
    __syncthreads();
  }
  block_reduce<2, kLinThreads>(v, red);
  if (threadIdx.x != 0) return;
  const double tmp = v[0] + (A.has_scale ? *A.scale_cost : 0.0);
  lm_decide(st, tmp, v[1], status, its, false, 0.0);
}

// Controller of a SPECULATIVE linearisation (last CTA): the launch has built the system of the trial state (cost
// included), so the accept / reject decision needs no separate cost pass, and an accepted trial -- the usual case --
// leaves the next LM iteration's system ready (its linearisation launch returns at once).  A rejected trial has
// overwritten the system of the current state: it is marked dirty and rebuilt, bit for bit the same, by the next
// launch.  The state sequence is that of the cost-pass form; only what is evaluated when differs.
__device__ __forceinline__ void spec_controller(const BAView& V, const LinArgs& A, double* red /*[3 * kLinWarps]*/) {
  LMState* st = A.st_rw;
  const TrialArgs& T = A.trial;
  double cost = 0.0, md = 0.0, sct = 0.0;
  const int n6 = 6 * V.Cf;
  const int n_max = max(max(max(V.C, n6), A.n_pt_ctas), T.n_sp);
  for (int i = threadIdx.x; i < n_max; i += kLinThreads) {
    const int col = i / 6, r = i - 6 * col;
    const double a = i < n6 ? __ldcg(A.Hpp + 36 * (size_t)col + 7 * r) : 0.0;
    const double x = (i < n6 && A.has_scale && r < 3) ? __ldcg(A.extra_diag + 3 * col + r) : 0.0;
    const double pm = i < A.n_pt_ctas ? __ldcg(A.pt_maxd + i) : 0.0;
    if (i < V.C) cost += __ldcg(A.cam_cost + i);
    if (i < T.n_sp) sct += __ldcg(T.scale_partial + i);
    md = fmax(md, fmax(fabs(a + x), pm));
  }
  int status = 0, its = 0;
  if (threadIdx.x == 0 && T.n) { status = __ldcg(T.pcg_status); its = __ldcg(T.pcg_iters); }
  const double sc = A.has_scale ? __ldcg(A.scale_cost) : 0.0;
  cost = warp_sum(cost); sct = warp_sum(sct);
  for (int o = 16; o > 0; o >>= 1) md = fmax(md, __shfl_xor_sync(0xffffffffu, md, o));
  const int wid = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { red[wid] = cost; red[kLinWarps + wid] = sct; red[2 * kLinWarps + wid] = md; }
  __syncthreads();
  if (threadIdx.x != 0) return;
  double c2 = 0.0, s2 = 0.0, m2 = 0.0;
  for (int w = 0; w < kLinWarps; ++w) { c2 += red[w]; s2 += red[kLinWarps + w]; m2 = fmax(m2, red[2 * kLinWarps + w]); }
  lm_decide(st, c2 + sc, s2, status, its, true, m2);
}

__device__ __forceinline__ constexpr int hpp_u(int i, int j) { return i * 6 - (i * (i - 1)) / 2 + (j - i); }

template <bool COST_ONLY>
__global__ void __launch_bounds__(kLinThreads, 4)
ba_linearize_kernel(BAView V, LinArgs A, ScaleView SV) {
  extern __shared__ __align__(128) double dyn_smem[];   // camera role: Hpl stage; point role: camera table
  __shared__ double red[kLinWarps * 32];
  __shared__ __align__(16) CamRT s_cam;
  __shared__ int s_flag;
  const int4* wsrc = reinterpret_cast<const int4*>((COST_ONLY ? A.cam_work : A.work) + blockIdx.x);
  const int4 w0 = __ldg(wsrc), w1 = __ldg(wsrc + 1);   // (static work list: may be read before the wait)
  pdl_wait();
  pdl_trigger();
  const bool spec = !COST_ONLY && A.spec != 0;
  if (V.st->done || (!COST_ONLY && !spec && !V.st->dirty && !A.force)) return;
  const int sb = (COST_ONLY || spec) ? (V.st->cur ^ 1) : V.st->cur;   // cost pass / speculative launch: the trial state
  const CamRT* __restrict__ camrt = V.camrt[sb];
  const double* __restrict__ pts = V.pts[sb];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int kind = w0.x;
  bool take_ticket = false;
  if (!COST_ONLY && A.trace && tid == 0) A.trace[2 * blockIdx.x] = globaltimer_ns();

  if (kind == 0) {
    // ------------------------------------------------ camera chunk
    const int c = w0.y, start = w0.z, end = w0.w, row = w1.x, n_chunks = w1.y, first_row = w1.z, col = w1.w;
    if (tid < 12) reinterpret_cast<double*>(&s_cam)[tid] = reinterpret_cast<const double*>(camrt + c)[tid];
    const bool cam_free = col >= 0;
    double v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = 0.0;
    // Inputs: the chunk's point indices and measurements (contiguous in camera-major order) are
    // fetched up front by cp.async into shared memory, all kLinObs observations of a thread in flight
    // at once, so the streams from HBM cost ONE exposed round trip per chunk; the points themselves
    // (random 32-byte sectors, mostly L2 hits, one 256-bit load each) run one observation ahead in
    // registers.
    const int o0 = start + tid;
    int* s_idx = reinterpret_cast<int*>(dyn_smem + kLinWarps * kHplStages * kStageDoubles);   // [kLinObs][kLinThreads]
    double2* s_uv = reinterpret_cast<double2*>(s_idx + kLinObs * kLinThreads);                 // [kLinObs][kLinThreads]
#pragma unroll
    for (int k = 0; k < kLinObs; ++k) {
      const int o = o0 + k * kLinThreads;
      if (o < end) {
        cp_async4(s_idx + k * kLinThreads + tid, V.cm_pt + o);
        cp_async16(s_uv + k * kLinThreads + tid, V.cm_uv + o);
      }
    }
    cp_async_commit();
    double* ws0 = dyn_smem + wid * (kHplStages * kStageDoubles);
    cp_async_wait<0>();                                  // each thread reads back only what it fetched
    double X[3] = {0.0, 0.0, 1.0};
    if (o0 < end) load_point(pts, s_idx[tid], X);
    __syncthreads();
    int o = o0;
    const int n_iter = min(kLinObs, (end - start + kLinThreads - 1) / kLinThreads);   // the host sizes chunks to the problem
#pragma unroll 1
    for (int k = 0; k < n_iter; ++k) {
      const bool valid = o < end;
      const double Xc[3] = {X[0], X[1], X[2]};
      double2 uvc = make_double2(0.0, 0.0);
      if (valid) uvc = s_uv[k * kLinThreads + tid];
      const int o_next = o + kLinThreads;
      if (k + 1 < kLinObs && o_next < end) load_point(pts, s_idx[(k + 1) * kLinThreads + tid], X);
      const bool want_hpl = !COST_ONLY && cam_free;
      double* ws = ws0 + (k % kHplStages) * kStageDoubles;
      double2* wsl = reinterpret_cast<double2*>(ws + lane * 18);
      if (want_hpl) {
        if (kBulkStore && lane == 0) {                   // the last store from this stage has left it
          if (kHplStages == 1) bulk_wait_read0(); else bulk_wait_read1();
        }
        __syncwarp();
      }
      bool ok = false;
      if (valid) {
        double e[2], rho, w, Jp[2][3], Jr[2][3];
        ok = eval_obs(s_cam.R, s_cam.t, Xc, uvc, V.K, !COST_ONLY, e, rho, w, Jp, Jr);
        if (ok) {
          v[27] += rho;
          if (want_hpl) {
            // blocks go to the warp's stage as soon as they exist (keeps the live set small)
            double wp0[3], wp1[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) { wp0[i] = w * Jp[0][i]; wp1[i] = w * Jp[1][i]; }
            double a[6];   // (w Jp^T Jp) upper: 00 01 02 11 12 22
            a[0] = wp0[0] * Jp[0][0] + wp1[0] * Jp[1][0];
            a[1] = wp0[0] * Jp[0][1] + wp1[0] * Jp[1][1];
            a[2] = wp0[0] * Jp[0][2] + wp1[0] * Jp[1][2];
            a[3] = wp0[1] * Jp[0][1] + wp1[1] * Jp[1][1];
            a[4] = wp0[1] * Jp[0][2] + wp1[1] * Jp[1][2];
            a[5] = wp0[2] * Jp[0][2] + wp1[2] * Jp[1][2];
            v[hpp_u(0, 0)] += a[0]; v[hpp_u(0, 1)] += a[1]; v[hpp_u(0, 2)] += a[2];
            v[hpp_u(1, 1)] += a[3]; v[hpp_u(1, 2)] += a[4]; v[hpp_u(2, 2)] += a[5];
            // Hpl translation rows = -(w Jp^T Jp)
            wsl[0] = make_double2(-a[0], -a[1]); wsl[1] = make_double2(-a[2], -a[1]);
            wsl[2] = make_double2(-a[3], -a[4]); wsl[3] = make_double2(-a[2], -a[4]);
            const double h8 = -a[5];
#pragma unroll
            for (int i = 0; i < 3; ++i) v[21 + i] = fma(wp0[i], e[0], fma(wp1[i], e[1], v[21 + i]));   // bp = -Jc^T w e
#pragma unroll
            for (int i = 0; i < 3; ++i) {
#pragma unroll
              for (int j = 0; j < 3; ++j)
                v[hpp_u(i, 3 + j)] = fma(-wp0[i], Jr[0][j], fma(-wp1[i], Jr[1][j], v[hpp_u(i, 3 + j)]));
            }
            double wr0[3], wr1[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) { wr0[i] = w * Jr[0][i]; wr1[i] = w * Jr[1][i]; }
#pragma unroll
            for (int i = 0; i < 3; ++i) {
#pragma unroll
              for (int j = i; j < 3; ++j)
                v[hpp_u(3 + i, 3 + j)] = fma(wr0[i], Jr[0][j], fma(wr1[i], Jr[1][j], v[hpp_u(3 + i, 3 + j)]));
              v[24 + i] = fma(-wr0[i], e[0], fma(-wr1[i], e[1], v[24 + i]));
            }
            // Hpl rotation rows
            double r[9];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
#pragma unroll
              for (int j = 0; j < 3; ++j) r[3 * i + j] = wr0[i] * Jp[0][j] + wr1[i] * Jp[1][j];
            }
            wsl[4] = make_double2(h8, r[0]); wsl[5] = make_double2(r[1], r[2]);
            wsl[6] = make_double2(r[3], r[4]); wsl[7] = make_double2(r[5], r[6]);
            wsl[8] = make_double2(r[7], r[8]);
          }
        }
      }
      if (want_hpl) {
        // the warp's 32 blocks are contiguous in HBM (camera-major): staged in the warp's slice of
        // shared memory (144 B stride: conflict-free STS.128), written by one bulk async store
        if (!ok) {
#pragma unroll
          for (int i = 0; i < 9; ++i) wsl[i] = make_double2(0.0, 0.0);
        }
        const int base = o - lane;                       // first observation of this warp's row
        const int n_blk = max(0, min(32, end - base));
        if (kBulkStore) {
          fence_async_smem();
          __syncwarp();
          if (lane == 0 && n_blk > 0) bulk_s2g(A.Hpl + (size_t)base * 18, ws, (uint32_t)n_blk * 144u);
        } else {
          __syncwarp();
          double2* dst = reinterpret_cast<double2*>(A.Hpl + (size_t)base * 18);
          const double2* srcp = reinterpret_cast<const double2*>(ws);
          for (int i = lane; i < n_blk * 9; i += 32) dst[i] = srcp[i];
        }
      }
      o = o_next;
    }
    // chunk partial: 28 values
    if (COST_ONLY) {
      double s = warp_sum(v[27]);
      if (lane == 0) red[wid] = s;
      __syncthreads();
      if (tid == 0) A.camp[(size_t)row * kCamVals + 27] = (red[0] + red[1]) + (red[2] + red[3]);
    } else {
      warp_transpose_sum32(v, lane);
      red[wid * 32 + lane] = v[0];
      __syncthreads();
      if (tid < kCamVals)
        A.camp[(size_t)row * kCamVals + tid] = (red[tid] + red[32 + tid]) + (red[64 + tid] + red[96 + tid]);
      if (lane == 0 && cam_free) bulk_wait_read0();      // the stage is free; the kernel boundary publishes Hpl
    }
    // last chunk of this camera folds the partials (chunk order -> deterministic)
    __syncthreads();
    if (tid == 0) {
      fence_acq_rel_gpu();   // cumulative over the CTA's writes (ordered before it by the barrier)
      const unsigned t = atomicAdd(A.cam_ticket + c, 1u);
      s_flag = (t == (unsigned)(n_chunks - 1));
      if (s_flag) A.cam_ticket[c] = 0u;
    }
    __syncthreads();
    if (s_flag) {
      fence_acq_rel_gpu();
      // the camera's chunk partials are summed by all four warps (warp g takes chunks g, g + 4, ...; the four sums
      // are added in warp order): a camera of a local window has ~24 chunks, and one thread per column walking them
      // one after the other was six dependent L2 round trips on the critical path of every linearisation / trial cost
      {
        double s = 0.0;
        if (lane < kCamVals) {
          const double* base = A.camp + (size_t)first_row * kCamVals + lane;
          for (int k = wid; k < n_chunks; k += kLinWarps) s += __ldcg(base + (size_t)k * kCamVals);
        }
        red[wid * 32 + lane] = s;
      }
      __syncthreads();
      if (tid < kCamVals) {
        const double s = (red[tid] + red[32 + tid]) + (red[64 + tid] + red[96 + tid]);
        if (tid == 27) A.cam_cost[c] = s;
        if (!COST_ONLY && col >= 0 && tid < 27) {
          if (tid >= 21) {
            A.bp[6 * col + (tid - 21)] = s;
          } else {
            int r = 0, k = tid;
            while (k >= 6 - r) { k -= 6 - r; ++r; }
            const int cc = r + k;
            A.Hpp[36 * col + 6 * r + cc] = s;
            A.Hpp[36 * col + 6 * cc + r] = s;
          }
        }
      }
      take_ticket = true;
    }
  } else if (kind == 1) {
    // ------------------------------------------------ point block (never reached when COST_ONLY)
    const int b = w0.y;
    if (A.cams_in_smem) {
      for (int i = tid; i < V.C * 6; i += kLinThreads) {
        const int c = i / 6, k = i - 6 * c;
        reinterpret_cast<double2*>(dyn_smem + (size_t)c * kCamTabStride)[k] =
            reinterpret_cast<const double2*>(camrt + c)[k];
      }
      __syncthreads();
    }
    double m = 0.0;
    const int wg = b * kLinWarps + wid;
    if (wg < A.n_pt_warps) {
      const int it0 = A.pw_first[wg], it1 = A.pw_first[wg + 1];
      int row_n = 0, row_n1 = 0, p_n = -1;
      if (it0 < it1) { row_n = A.item_row[it0]; row_n1 = A.item_row[it0 + 1]; p_n = A.ell_pt[(size_t)it0 * 32 + lane]; }
#pragma unroll 1
      for (int it = it0; it < it1; ++it) {
        const int row0 = row_n, row1 = row_n1, p = p_n;
        if (it + 1 < it1) {   // header of the next item
          row_n = row1; row_n1 = A.item_row[it + 2]; p_n = A.ell_pt[(size_t)(it + 1) * 32 + lane];
        }
        double X[3] = {0.0, 0.0, 1.0};
        if (p >= 0) load_point(pts, p, X);
        double acc[9];
#pragma unroll
        for (int j = 0; j < 9; ++j) acc[j] = 0.0;
        auto accumulate = [&](int c, double2 uv) {
          if (c < 0) return;
          double Rt[12];
          if (A.cams_in_smem) {
            const double2* s = reinterpret_cast<const double2*>(dyn_smem + (size_t)c * kCamTabStride);
#pragma unroll
            for (int k = 0; k < 6; ++k) { const double2 q = s[k]; Rt[2 * k] = q.x; Rt[2 * k + 1] = q.y; }
          } else {
            const double2* s = reinterpret_cast<const double2*>(camrt + c);
#pragma unroll
            for (int k = 0; k < 6; ++k) { const double2 q = __ldg(s + k); Rt[2 * k] = q.x; Rt[2 * k + 1] = q.y; }
          }
          double e[2], rho, w, Jp[2][3], Jr[2][3];
          if (!eval_obs(Rt, Rt + 9, X, uv, V.K, false, e, rho, w, Jp, Jr)) return;
          double wp0[3], wp1[3];
#pragma unroll
          for (int i = 0; i < 3; ++i) { wp0[i] = w * Jp[0][i]; wp1[i] = w * Jp[1][i]; }
          acc[0] = fma(wp0[0], Jp[0][0], fma(wp1[0], Jp[1][0], acc[0]));
          acc[1] = fma(wp0[0], Jp[0][1], fma(wp1[0], Jp[1][1], acc[1]));
          acc[2] = fma(wp0[0], Jp[0][2], fma(wp1[0], Jp[1][2], acc[2]));
          acc[3] = fma(wp0[1], Jp[0][1], fma(wp1[1], Jp[1][1], acc[3]));
          acc[4] = fma(wp0[1], Jp[0][2], fma(wp1[1], Jp[1][2], acc[4]));
          acc[5] = fma(wp0[2], Jp[0][2], fma(wp1[2], Jp[1][2], acc[5]));
#pragma unroll
          for (int j = 0; j < 3; ++j) acc[6 + j] = fma(-wp0[j], e[0], fma(-wp1[j], e[1], acc[6 + j]));
        };
        // four register sets in rotation (no loop-carried copy behind a load): the loads of rows
        // r+4 .. r+7 are in flight under the arithmetic of rows r .. r+3
        const size_t sl = (size_t)row0 * 32 + lane;
        int cq[4];
        double2 uq[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          cq[u] = -1; uq[u] = make_double2(0.0, 0.0);
          if (row0 + u < row1) { cq[u] = A.ell_cam[sl + 32 * u]; uq[u] = A.ell_uv[sl + 32 * u]; }
        }
#pragma unroll 1
        for (int r = row0; r < row1; r += 4) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int c0 = cq[u];
            const double2 uv0 = uq[u];
            cq[u] = -1;
            if (r + 4 + u < row1) {
              const size_t s2 = (size_t)(r + 4 + u) * 32 + lane;
              cq[u] = A.ell_cam[s2]; uq[u] = A.ell_uv[s2];
            }
            accumulate(c0, uv0);
          }
        }
        if (p >= 0) {
          const int pc = V.pt_col[p];
#pragma unroll
          for (int j = 0; j < 6; ++j) A.Hll[6 * (size_t)pc + j] = acc[j];
#pragma unroll
          for (int j = 0; j < 3; ++j) A.bl[3 * (size_t)pc + j] = acc[6 + j];
          m = fmax(m, fmax(fabs(acc[0]), fmax(fabs(acc[3]), fabs(acc[5]))));
        }
      }
    }
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) red[wid] = m;
    __syncthreads();
    if (tid == 0) A.pt_maxd[b] = fmax(fmax(red[0], red[1]), fmax(red[2], red[3]));
    take_ticket = true;
  } else {
    // ------------------------------------------------ scale edges
    scale_edges_body<kLinThreads>(SV, false, red, spec);
    take_ticket = true;
  }

  if (!COST_ONLY && A.trace && tid == 0) A.trace[2 * blockIdx.x + 1] = globaltimer_ns();
  // the last CTA of the grid (camera finalisers + point blocks + scale edges; camera finalisers only in
  // the cost-only launch) runs the controller
  if (!take_ticket) return;     // uniform per CTA
  __syncthreads();
  if (tid == 0) {
    fence_acq_rel_gpu();
    const unsigned t = atomicAdd(A.lin_ticket, 1u);
    s_flag = (t == (unsigned)((COST_ONLY ? V.C : A.n_tickets) - 1));
    if (s_flag) *A.lin_ticket = 0u;
  }
  __syncthreads();
  if (s_flag) {
    if (!COST_ONLY && A.trace && tid == 0) A.trace[2 * gridDim.x + 3] = globaltimer_ns();
    fence_acq_rel_gpu();
    if (COST_ONLY) trial_controller(V, A, SV, red);
    else if (spec) spec_controller(V, A, red);
    else lin_controller(V, A, red);
    if (!COST_ONLY && A.trace && tid == 0) A.trace[2 * gridDim.x] = globaltimer_ns();
  }
}

// Per free point: Hinv = (Hll + lambda I)^-1 (back-substitution) and the upper-triangular factor W
// with W W^T = Hinv (W = C^-T, C the Cholesky factor of Hll + lambda I), stored (00 01 02 11 12 22).
__global__ void ba_point_invert(const LMState* __restrict__ st, int Pf, const double* __restrict__ Hll,
                                double* __restrict__ Hinv, double* __restrict__ Wf) {
  pdl_wait();
  pdl_trigger();
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
  // Cholesky h = C C^T (lower), then Cinv; W = Cinv^T
  const double c00 = sqrt(h[0]);
  const double i00 = 1.0 / c00;
  const double c10 = h[1] * i00, c20 = h[2] * i00;
  const double c11 = sqrt(h[3] - c10 * c10);
  const double i11 = 1.0 / c11;
  const double c21 = (h[4] - c20 * c10) * i11;
  const double c22 = sqrt(h[5] - c20 * c20 - c21 * c21);
  const double i22 = 1.0 / c22;
  // Cinv (lower): rows (i00, 0, 0), (-c10 i00 i11, i11, 0), ((c10 c21 i11 - c20) i00 i22, -c21 i11 i22, i22)
  const double m10 = -c10 * i00 * i11;
  const double m21 = -c21 * i11 * i22;
  const double m20 = -(c20 * i00 + c21 * m10) * i22;
  double* w = Wf + 6 * (size_t)p;    // W[k][j] = Cinv[j][k], j >= k
  w[0] = i00; w[1] = m10; w[2] = m20; w[3] = i11; w[4] = m21; w[5] = i22;
}

// G = Hpl * W per observation (camera-major), so that Hpl Hinv Hpl^T = G G^T, and the per-camera
// partial of sum Hpl Hinv bl = sum G (W^T bl).
__global__ void __launch_bounds__(kCamThreads)
ba_compute_G(BAView V, const double* __restrict__ Hpl, const double* __restrict__ Wf,
             const double* __restrict__ bl, double* __restrict__ Gout, double* __restrict__ camp,
             const double* __restrict__ Hll /* non-null: W is formed here from Hll + lambda I (no ba_point_invert launch) */) {
  __shared__ double red[(kCamThreads / 32) * 6];
  __shared__ __align__(128) double stage_all[(kCamThreads / 32) * 32 * 18];   // a warp's 32 blocks (4608 B)
  static_assert(kCamChunk == kCamThreads, "one observation per thread and chunk");
  // static structure before the wait (see ba_schur_finalize): chunk bounds and this thread's point column
  const int c = blockIdx.y, chunk = blockIdx.x;
  const int start = __ldg(V.cm_start + c) + chunk * kCamChunk;
  const int end = min(__ldg(V.cm_start + c + 1), start + kCamChunk);
  const int ccol = __ldg(V.cam_col + c);
  const int o_mine = start + threadIdx.x;
  const int pcol_mine = o_mine < end ? __ldg(V.pt_col + __ldg(V.cm_pt + o_mine)) : -1;
  pdl_wait();
  pdl_trigger();
  if (V.st->done) return;
  if (start >= end || ccol < 0) return;
  // The warp's 32 blocks are contiguous in HBM (camera-major).  A lane reading ITS block straight from global
  // memory makes every 128-bit load of the warp touch 32 different lines (144-byte stride); instead the warp copies
  // the 4608 bytes with consecutive lanes on consecutive 16-byte pieces (cp.async) into its slice of shared memory,
  // every lane works on its block there (144-byte stride: conflict-free 128-bit accesses), G goes back into the
  // same slice and leaves by ONE bulk async store.
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double* stage = stage_all + wid * (32 * 18);
  const int base = start + wid * 32;                     // first observation of this warp
  const int n_blk = max(0, min(32, end - base));
  for (int i = lane; i < n_blk * 9; i += 32) cp_async16(stage + 2 * i, Hpl + (size_t)base * 18 + 2 * i);
  cp_async_commit();
  double acc[6] = {0, 0, 0, 0, 0, 0};
  double w[6], b[3];
  const int pcol = pcol_mine;
  if (pcol >= 0) {                                       // (the gathers run under the copy)
    if (Hll != nullptr) {
      // small problems: the upper factor W (W W^T = (Hll + lambda I)^-1) per observation, by the statements of
      // ba_point_invert -- one launch fewer in a chain where a launch costs as much as the arithmetic
      double h[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) h[j] = Hll[6 * (size_t)pcol + j];
      const double lambda = V.st->lambda;
      h[0] += lambda; h[3] += lambda; h[5] += lambda;
      const double c00 = sqrt(h[0]);
      const double i00 = 1.0 / c00;
      const double c10 = h[1] * i00, c20 = h[2] * i00;
      const double c11 = sqrt(h[3] - c10 * c10);
      const double i11 = 1.0 / c11;
      const double c21 = (h[4] - c20 * c10) * i11;
      const double c22 = sqrt(h[5] - c20 * c20 - c21 * c21);
      const double i22 = 1.0 / c22;
      const double m10 = -c10 * i00 * i11;
      const double m21 = -c21 * i11 * i22;
      const double m20 = -(c20 * i00 + c21 * m10) * i22;
      w[0] = i00; w[1] = m10; w[2] = m20; w[3] = i11; w[4] = m21; w[5] = i22;
    } else {
#pragma unroll
      for (int j = 0; j < 6; ++j) w[j] = Wf[6 * (size_t)pcol + j];
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) b[j] = bl[3 * (size_t)pcol + j];
  }
  cp_async_wait<0>();
  __syncwarp();
  if (pcol >= 0) {
    double h[18], g[18];
    load18(stage + lane * 18, h);
    // y = W^T bl  (W upper: 00 01 02 11 12 22)
    const double y0 = w[0] * b[0];
    const double y1 = w[1] * b[0] + w[3] * b[1];
    const double y2 = w[2] * b[0] + w[4] * b[1] + w[5] * b[2];
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const double h0 = h[3 * r], h1 = h[3 * r + 1], h2 = h[3 * r + 2];
      g[3 * r] = h0 * w[0];
      g[3 * r + 1] = h0 * w[1] + h1 * w[3];
      g[3 * r + 2] = h0 * w[2] + h1 * w[4] + h2 * w[5];
      acc[r] += g[3 * r] * y0 + g[3 * r + 1] * y1 + g[3 * r + 2] * y2;
    }
    store18(stage + lane * 18, g);                       // (rows of fixed points keep the copied block: never read)
  }
  fence_async_smem();
  __syncwarp();
  if (lane == 0 && n_blk > 0) bulk_s2g(Gout + (size_t)base * 18, stage, (uint32_t)n_blk * 144u);
  block_reduce<6, kCamThreads>(acc, red);
  if (threadIdx.x == 0) {
    double* dst = camp + ((size_t)c * V.max_chunks + chunk) * kCamVals;
#pragma unroll
    for (int i = 0; i < 6; ++i) dst[i] = acc[i];
  }
  if (lane == 0) bulk_wait_read0();                      // the store has left the slice before the CTA retires
}

// Schur work units: one warp reduces <= kUnitItems products G_a * G_b^T of one (cam a, cam b) key,
// one item per lane and round.  The 64 blocks of a round (144 B each, scattered in HBM) are fetched
// COOPERATIVELY: consecutive lanes copy consecutive 16-byte pieces of a block with cp.async (LDGSTS),
// so one instruction touches ~5 lines instead of 32, into a kSchurStages-deep ring in shared memory
// (lane-per-block reads at 144 B stride are conflict-free).  Item indices are prefetched one round
// further ahead.  (One bulk async copy per block was measured slower: the TMA unit retires about one
// small copy per 15 cycles per SM.)
constexpr int kSchurStages = 2;
constexpr int kSchurWarps = 4;
constexpr int kSchurStageDoubles = 2 * 32 * 18;   // A blocks [32][18] then B blocks [32][18]
constexpr int kSchurSmem = kSchurWarps * kSchurStages * kSchurStageDoubles * 8;

__global__ void __launch_bounds__(kSchurWarps * 32, 3)
ba_schur_units(const LMState* __restrict__ st, const double* __restrict__ G,
               const int* __restrict__ item_a, const int* __restrict__ item_b,
               const int* __restrict__ unit_start, const int* __restrict__ unit_len, int n_units,
               double* __restrict__ partial) {
  extern __shared__ __align__(128) double sch_smem[];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int u = blockIdx.x * kSchurWarps + wid;
  // (static work lists: the unit's bounds are fetched before the wait, see ba_schur_finalize)
  const int s0 = u < n_units ? __ldg(unit_start + u) : 0, len_raw = u < n_units ? __ldg(unit_len + u) : 0;
  const int len = len_raw & 0x3fffffff;
  const bool diag = (len_raw >> 30) & 1;   // self-paired items (set by the host): each block is fetched once
  pdl_wait();
  pdl_trigger();
  if (st->done || u >= n_units) return;
  double* ring = sch_smem + (size_t)wid * kSchurStages * kSchurStageDoubles;
  const int n_it = (len + 31) >> 5;
  auto load_idx = [&](int it, int& ia, int& ib) {
    const int i = it * 32 + lane;
    ia = -1; ib = -1;
    if (it < n_it && i < len) { ia = item_a[s0 + i]; ib = item_b[s0 + i]; }
  };
  auto issue = [&](int it, int ia, int ib) {   // round `it` -> stage it % kSchurStages
    if (it < n_it) {
      double* stg = ring + (it % kSchurStages) * kSchurStageDoubles;
#pragma unroll
      for (int j = 0; j < 18; ++j) {
        if (j >= 9 && diag) break;
        const int q = j * 32 + lane;                 // 16-byte piece of the stage
        const int b = (q * 7282) >> 16;              // q / 9: block 0..63 (A blocks, then B blocks)
        const int r = q - 9 * b;
        const int idx = __shfl_sync(0xffffffffu, j < 9 ? ia : ib, b & 31);
        if (idx >= 0) cp_async16(stg + 2 * q, G + (size_t)idx * 18 + 2 * r);
      }
    }
    cp_async_commit();
  };
  int ia_n, ib_n, ia_m, ib_m;   // indices of rounds it + kSchurStages and it + kSchurStages + 1
  {
    int ia[kSchurStages], ib[kSchurStages];
#pragma unroll
    for (int s = 0; s < kSchurStages; ++s) load_idx(s, ia[s], ib[s]);
    load_idx(kSchurStages, ia_n, ib_n);
    load_idx(kSchurStages + 1, ia_m, ib_m);
#pragma unroll
    for (int s = 0; s < kSchurStages; ++s) issue(s, ia[s], ib[s]);
  }
  double acc[36];
#pragma unroll
  for (int i = 0; i < 36; ++i) acc[i] = 0.0;
#pragma unroll 1
  for (int it = 0; it < n_it; ++it) {
    const int i = it * 32 + lane;
    const double* stg = ring + (it % kSchurStages) * kSchurStageDoubles;
    cp_async_wait<kSchurStages - 1>();
    __syncwarp();                                    // every lane's pieces of this round have landed
    if (i < len) {
      double ga[18], gb[18];
      load18(stg + lane * 18, ga);
      load18(stg + (diag ? 0 : 32 * 18) + lane * 18, gb);
#pragma unroll
      for (int r = 0; r < 6; ++r) {
#pragma unroll
        for (int c = 0; c < 6; ++c)
          acc[6 * r + c] = fma(ga[3 * r], gb[3 * c], fma(ga[3 * r + 1], gb[3 * c + 1],
                           fma(ga[3 * r + 2], gb[3 * c + 2], acc[6 * r + c])));
      }
    }
    // refill this stage with round it + kSchurStages (indices arrived a round ago)
    const int ia_u = ia_n, ib_u = ib_n;
    ia_n = ia_m; ib_n = ib_m;
    load_idx(it + kSchurStages + 2, ia_m, ib_m);     // two rounds ahead of its use
    __syncwarp();                                    // all lanes are done reading the stage
    issue(it + kSchurStages, ia_u, ib_u);
  }
  // 36 warp sums: 32 of them by the transposed butterfly (31 shuffles instead of 160), 4 plain
  {
    double tail[4] = {acc[32], acc[33], acc[34], acc[35]};
    double (&head)[32] = reinterpret_cast<double (&)[32]>(acc);
    warp_transpose_sum32(head, lane);                // lane L now holds the total of entry L in head[0]
#pragma unroll
    for (int i = 0; i < 4; ++i) tail[i] = warp_sum(tail[i]);
    partial[(size_t)u * 36 + lane] = head[0];
    if (lane < 4) partial[(size_t)u * 36 + 32 + lane] = lane == 0 ? tail[0] : lane == 1 ? tail[1] : lane == 2 ? tail[2] : tail[3];
  }
}

// Schur complement by RESIDENT CHUNKS (the default; the sorted-unit kernel above is kept as a variant).  A CTA
// owns kSchurChunk camera-major observations of ONE free camera a.  Load phase: for each of them
// Y_o = Hpl_o (Hll_j + lambda I)^-1 (6x3; the 3x3 inverse is formed on the fly, so neither the point inverses nor
// G = Hpl W are ever written to HBM) goes to shared memory, and the chunk's share of sum Hpl Hinv bl (bhat) is
// reduced.  Item phase: an item pairs a resident observation with one partner observation o' of the same point
// (camera column b >= a's, its own included); its contribution to block (a, b) is Y_o Hpl_o'^T.  Items are sorted by
// b; a warp takes one (chunk, b) unit at a time, one item per lane and round; only the PARTNER blocks are gathered
// (cooperative cp.async, consecutive lanes copy consecutive 16-byte pieces): half the L2 traffic of gathering both
// factors, and no G pass.  36 warp sums per unit by the transposed butterfly -> ck_partial[unit], units numbered
// in (a, b, chunk) order so that ba_schur_finalize sums a block's partials in a fixed order.
constexpr int kSchurChunk = 512;
constexpr int kChunkThreads = 256;
constexpr int kChunkWarps = kChunkThreads / 32;
constexpr int kChunkStageDoubles = 32 * 18;
constexpr int kChunkSmem = (kSchurChunk * 18 + kChunkWarps * kChunkStageDoubles) * 8;

struct ChunkArgs {
  const int* ck_cam; const int* ck_start; const int* ck_end; const int* ck_ustart; const int* cu_unit;
  const int* cu_istart; const unsigned short* it_local; const int* it_partner;
  const double* Hpl; const double* Hll; const double* bl;
  double* ck_bh; double* partial;
};

__global__ void __launch_bounds__(kChunkThreads, 2) ba_schur_chunks(BAView V, ChunkArgs A) {
  extern __shared__ __align__(128) double ch_smem[];
  __shared__ double red[kChunkWarps * 6];
  double* Ys = ch_smem;                                            // [kSchurChunk][18]
  const int tid = threadIdx.x, wid = tid >> 5, lane = tid & 31, ck = blockIdx.x;
  double* stage = ch_smem + kSchurChunk * 18 + wid * kChunkStageDoubles;   // this warp's 32 gathered partner blocks
  const int o0 = __ldg(A.ck_start + ck), o1 = __ldg(A.ck_end + ck);
  const int u0 = __ldg(A.ck_ustart + ck), u1 = __ldg(A.ck_ustart + ck + 1);
  pdl_wait();
  pdl_trigger();
  if (V.st->done) return;
  const double lambda = V.st->lambda;
  // ---- load phase
  double bh[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
  for (int rep = 0; rep < kSchurChunk / kChunkThreads; ++rep) {
    const int loc = rep * kChunkThreads + tid, o = o0 + loc;
    double y[18];
#pragma unroll
    for (int i = 0; i < 18; ++i) y[i] = 0.0;
    if (o < o1) {
      const int pcol = V.pt_col[V.cm_pt[o]];
      if (pcol >= 0) {
        double h[6], hi[6], b[3], H[18];
#pragma unroll
        for (int j = 0; j < 6; ++j) h[j] = A.Hll[6 * (size_t)pcol + j];
#pragma unroll
        for (int j = 0; j < 3; ++j) b[j] = A.bl[3 * (size_t)pcol + j];
        load18(A.Hpl + (size_t)o * 18, H);
        h[0] += lambda; h[3] += lambda; h[5] += lambda;
        sym3_inverse(h, hi);
        double z[3];
        sym3_mul(hi, b, z);
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          const double h0 = H[3 * r], h1 = H[3 * r + 1], h2 = H[3 * r + 2];
          y[3 * r] = h0 * hi[0] + h1 * hi[1] + h2 * hi[2];
          y[3 * r + 1] = h0 * hi[1] + h1 * hi[3] + h2 * hi[4];
          y[3 * r + 2] = h0 * hi[2] + h1 * hi[4] + h2 * hi[5];
          bh[r] += h0 * z[0] + h1 * z[1] + h2 * z[2];
        }
      }
    }
    store18(Ys + (size_t)loc * 18, y);
  }
  block_reduce<6, kChunkThreads>(bh, red);                         // (ends with a block barrier: Ys is visible)
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 6; ++i) A.ck_bh[6 * (size_t)ck + i] = bh[i];
  }
  // ---- item phase
  for (int ul = u0 + wid; ul < u1; ul += kChunkWarps) {
    const int unit = __ldg(A.cu_unit + ul);
    const int i0 = __ldg(A.cu_istart + ul), i1 = __ldg(A.cu_istart + ul + 1);
    double acc[36];
#pragma unroll
    for (int i = 0; i < 36; ++i) acc[i] = 0.0;
    int loc_n = -1, part_n = -1;
    if (i0 + lane < i1) { loc_n = A.it_local[i0 + lane]; part_n = A.it_partner[i0 + lane]; }
#pragma unroll 1
    for (int base = i0; base < i1; base += 32) {
      const int loc = loc_n, part = part_n;
      loc_n = -1; part_n = -1;
      if (base + 32 + lane < i1) { loc_n = A.it_local[base + 32 + lane]; part_n = A.it_partner[base + 32 + lane]; }
      __syncwarp();                                                // the stage has been read by every lane
#pragma unroll
      for (int j = 0; j < 9; ++j) {
        const int q = j * 32 + lane;                               // 16-byte piece of the stage
        const int bk = (q * 7282) >> 16;                           // q / 9: block 0..31
        const int r = q - 9 * bk;
        const int idx = __shfl_sync(0xffffffffu, part, bk);
        if (idx >= 0) cp_async16(stage + 2 * q, A.Hpl + (size_t)idx * 18 + 2 * r);
      }
      cp_async_commit();
      cp_async_wait<0>();
      __syncwarp();
      if (part >= 0) {
        double gb[18];
        load18(stage + lane * 18, gb);
        const double* yr = Ys + (size_t)loc * 18;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          const double y0 = yr[3 * r], y1 = yr[3 * r + 1], y2 = yr[3 * r + 2];
#pragma unroll
          for (int c = 0; c < 6; ++c)
            acc[6 * r + c] = fma(y0, gb[3 * c], fma(y1, gb[3 * c + 1], fma(y2, gb[3 * c + 2], acc[6 * r + c])));
        }
      }
    }
    double tail[4] = {acc[32], acc[33], acc[34], acc[35]};
    double (&head)[32] = reinterpret_cast<double (&)[32]>(acc);
    warp_transpose_sum32(head, lane);
#pragma unroll
    for (int i = 0; i < 4; ++i) tail[i] = warp_sum(tail[i]);
    A.partial[(size_t)unit * 36 + lane] = head[0];
    if (lane < 4) A.partial[(size_t)unit * 36 + 32 + lane] = lane == 0 ? tail[0] : lane == 1 ? tail[1] : lane == 2 ? tail[2] : tail[3];
  }
}

// S block (a,b) = [a==b](Hpp_a + lambda I) + scale-edge block - sum of its units; mirrored.
struct BhatArgs {   // bhat = (bp + bp_extra) - sum_j Hpl_j Hinv_j bl_j  (ordered chunk sum of ba_compute_G's partials)
  const double* camp; const double* bp; const double* bp_extra; double* bhat;
  unsigned* pcg_bar; int* pcg_flags;   // reset here for the solve that follows (no memset nodes in the step)
  const double* ck_bh; const int* cam_ck_start;   // chunked Schur variant: per-chunk partials [chunk][6], chunks of a camera
};
__device__ __forceinline__ double bhat_partial_sum(const BAView& V, const BhatArgs& Bh, int c, int i) {
  double s = 0.0;
  if (Bh.ck_bh != nullptr) {
    for (int k = Bh.cam_ck_start[c]; k < Bh.cam_ck_start[c + 1]; ++k) s += Bh.ck_bh[6 * (size_t)k + i];
  } else if (Bh.camp != nullptr) {
    const int n_obs = V.cm_start[c + 1] - V.cm_start[c];
    const int n_chunks = (n_obs + kCamChunk - 1) / kCamChunk;
    for (int k = 0; k < n_chunks; ++k) s += Bh.camp[((size_t)c * V.max_chunks + k) * kCamVals + i];
  }
  return s;
}
template <int PARTS>   // threads per entry: 4 when a block has many unit partials (small work units), else 1
__global__ void ba_schur_finalize(BAView V, int n_seg, int n, const int* __restrict__ seg_a,
                                  const int* __restrict__ seg_b,
                                  const int* __restrict__ seg_unit_start,
                                  const double* __restrict__ partial,
                                  const double* __restrict__ Hpp,
                                  const double* __restrict__ extra,
                                  const int* __restrict__ seg_extra_idx, double* __restrict__ S, BhatArgs Bh) {
  // the work lists are static structure: fetched BEFORE the wait, i.e. while the producing kernels still run
  // (programmatic dependent launch), so the dependent part starts with its indices in registers
  const int t = (blockIdx.x * blockDim.x + threadIdx.x) / PARTS, part = threadIdx.x % PARTS;
  const int seg = t / 36, e = t % 36;
  int a = 0, b = 0, ex = -1, u0 = 0, u1 = 0;
  if (seg < n_seg) {
    a = __ldg(seg_a + seg); b = __ldg(seg_b + seg); ex = __ldg(seg_extra_idx + seg);
    u0 = __ldg(seg_unit_start + seg); u1 = __ldg(seg_unit_start + seg + 1);
  }
  pdl_wait();
  pdl_trigger();
  if (blockIdx.x == 0 && threadIdx.x == 0) {   // PCG bookkeeping of this trial: grid-barrier counter, fail / status / iterations
    *Bh.pcg_bar = 0u;
    Bh.pcg_flags[0] = 0; Bh.pcg_flags[1] = 0; Bh.pcg_flags[2] = 0; Bh.pcg_flags[3] = 0;
  }
  if (V.st->done) return;
  const double lambda = V.st->lambda;
  const unsigned quad_mask = __ballot_sync(0xffffffffu, seg < n_seg);   // (the lanes that take part in the sums below)
  if (seg >= n_seg) {
    // the threads behind the S blocks build bhat (one per camera and row)
    const int tb = t - n_seg * 36;
    const int c = tb / 6, i = tb - 6 * c;
    if (Bh.bhat == nullptr || c >= V.C || part != 0) return;
    const int col = V.cam_col[c];
    if (col < 0) return;
    const double s = bhat_partial_sum(V, Bh, c, i);
    Bh.bhat[6 * col + i] = (Bh.bp[6 * col + i] + Bh.bp_extra[6 * col + i]) - s;
    return;
  }
  const int r = e / 6, c = e % 6;
  double v = ex >= 0 ? extra[(size_t)ex * 36 + e] : 0.0;
  if (a == b) v += Hpp[36 * a + e] + (r == c ? lambda : 0.0);
  double s = 0.0;
  for (int u = u0 + part; u < u1; u += 8 * PARTS) {   // eight loads in flight per thread, added in unit order
    double w[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) w[q] = (u + q * PARTS < u1) ? __ldcg(partial + (size_t)(u + q * PARTS) * 36 + e) : 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q)
      if (u + q * PARTS < u1) s += w[q];
  }
  if (PARTS == 4) {                     // (p0 + p1) + (p2 + p3): a fixed tree
    s += __shfl_xor_sync(quad_mask, s, 1);
    s += __shfl_xor_sync(quad_mask, s, 2);
    if (part != 0) return;
  }
  v -= s;
  S[(size_t)(6 * a + r) * n + 6 * b + c] = v;
  if (a != b) S[(size_t)(6 * b + c) * n + 6 * a + r] = v;
}

// Back-substitution, camera-major half: v_o = Hpl_o^T dp_cam(o), scattered to the point-major slot.
__global__ void __launch_bounds__(kCamThreads)
ba_backsub_cam(BAView V, const double* __restrict__ Hpl, const double* __restrict__ dp,
               double* __restrict__ tmp3) {
  const int c = blockIdx.y, chunk = blockIdx.x;   // (static structure before the wait, see ba_schur_finalize)
  const int start = __ldg(V.cm_start + c) + chunk * kCamChunk;
  const int end = min(__ldg(V.cm_start + c + 1), start + kCamChunk);
  const int col = __ldg(V.cam_col + c);
  const int o_mine = start + threadIdx.x;
  const int pcol_mine = o_mine < end ? __ldg(V.pt_col + __ldg(V.cm_pt + o_mine)) : -1;
  const int pm_mine = o_mine < end ? __ldg(V.cm_pm + o_mine) : 0;
  __shared__ __align__(128) double stage_all[(kCamThreads / 32) * 32 * 18];   // see ba_compute_G
  pdl_wait();
  pdl_trigger();
  if (V.st->done || start >= end || col < 0) return;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double* stage = stage_all + wid * (32 * 18);
  {
    const int base = start + wid * 32;
    const int n_blk = max(0, min(32, end - base));
    for (int i = lane; i < n_blk * 9; i += 32) cp_async16(stage + 2 * i, Hpl + (size_t)base * 18 + 2 * i);
    cp_async_commit();
  }
  double d[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) d[i] = dp[6 * col + i];
  cp_async_wait<0>();
  __syncwarp();
  for (int o = o_mine; o < end; o += kCamThreads) {
    if (pcol_mine < 0) continue;
    double h[18];
    load18(stage + lane * 18, h);
    double v[3] = {0, 0, 0};
#pragma unroll
    for (int r = 0; r < 6; ++r) {
#pragma unroll
      for (int j = 0; j < 3; ++j) v[j] += h[3 * r + j] * d[r];
    }
    // rows padded to one 32-byte sector: the scattered store covers the whole sector (no partial-sector fill)
    *reinterpret_cast<double4*>(tmp3 + (size_t)pm_mine * 4) = make_double4(v[0], v[1], v[2], 0.0);
  }
}

// SBACam::update on every free camera (t += d[0:3]; q <- normalize(q * (sqrt(1-|v|^2), v))).
__device__ __forceinline__ void update_cams_body(const BAView& V, const double* __restrict__ dp, int c) {
  if (c >= V.C) return;
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

// Back-substitution, point half: dl = Hinv (bl - sum_o v_o); trial point = point + dl.
// Also the per-CTA partial of sum dl (lambda dl + bl)  (computeScale, landmark part).
__global__ void __launch_bounds__(256)
ba_backsub_pt(BAView V, const double* __restrict__ tmp3, const double* __restrict__ Hll,
              const double* __restrict__ bl, double* __restrict__ dl_out,
              double* __restrict__ scale_partial, const double* __restrict__ dp, int n_pt_blocks,
              const double* __restrict__ bp, const double* __restrict__ bp_extra) {
  __shared__ double red[8];
  // static structure before the wait (see ba_schur_finalize)
  const int p_mine = blockIdx.x * blockDim.x + threadIdx.x;
  const bool pt_role = (int)blockIdx.x < n_pt_blocks && p_mine < V.P;
  const int col_mine = pt_role ? __ldg(V.pt_col + p_mine) : -1;
  const int ps0 = pt_role ? __ldg(V.pt_start + p_mine) : 0, ps1 = pt_role ? __ldg(V.pt_start + p_mine + 1) : 0;
  pdl_wait();
  pdl_trigger();
  if (V.st->done) return;
  if ((int)blockIdx.x >= n_pt_blocks) {   // the blocks behind the points update the cameras (SBACam::update)
    const int c = ((int)blockIdx.x - n_pt_blocks) * 256 + (int)threadIdx.x;
    update_cams_body(V, dp, c);
    // ... and leave the camera part of computeScale, sum dp (lambda dp + bp + bp_extra), as this block's partial: the
    // controller reads partials only (the speculative linearisation that ends a trial overwrites bp before it runs)
    double sc[1] = {0.0};
    const int col = c < V.C ? V.cam_col[c] : -1;
    if (col >= 0) {
      const double lambda = V.st->lambda;
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        const double d = dp[6 * col + i];
        sc[0] += d * (lambda * d + (bp[6 * col + i] + bp_extra[6 * col + i]));
      }
    }
    block_reduce<1, 256>(sc, red);
    if (threadIdx.x == 0) scale_partial[blockIdx.x] = sc[0];
    return;
  }
  const double lambda = V.st->lambda;
  const double* __restrict__ pts = V.pts[V.st->cur];
  double* __restrict__ pts_trial = V.pts[V.st->cur ^ 1];
  const int p = p_mine;
  double sc[1] = {0.0};
  if (p < V.P) {
    const int col = col_mine;
    double X[3] = {pts[4 * (size_t)p], pts[4 * (size_t)p + 1], pts[4 * (size_t)p + 2]};
    if (col >= 0) {
      double r[3] = {bl[3 * (size_t)col], bl[3 * (size_t)col + 1], bl[3 * (size_t)col + 2]};
      const double b0 = r[0], b1 = r[1], b2 = r[2];
      // (rows of observations by fixed cameras are never written by ba_backsub_cam and were zeroed at build time:
      // subtracting them is exact, and the loop has no dependent index loads in front of its data loads)
      for (int o = ps0; o < ps1; ++o) {
        const double4 t = *reinterpret_cast<const double4*>(tmp3 + 4 * (size_t)o);
        r[0] -= t.x; r[1] -= t.y; r[2] -= t.z;
      }
      double h[6], hi[6], d[3];   // (Hll + lambda I)^-1 on the fly: the inverses are not kept in HBM
#pragma unroll
      for (int j = 0; j < 6; ++j) h[j] = Hll[6 * (size_t)col + j];
      h[0] += lambda; h[3] += lambda; h[5] += lambda;
      sym3_inverse(h, hi);
      sym3_mul(hi, r, d);
      sc[0] = d[0] * (lambda * d[0] + b0) + d[1] * (lambda * d[1] + b1) + d[2] * (lambda * d[2] + b2);
      X[0] += d[0]; X[1] += d[1]; X[2] += d[2];
      dl_out[3 * (size_t)col] = d[0]; dl_out[3 * (size_t)col + 1] = d[1]; dl_out[3 * (size_t)col + 2] = d[2];
    }
    *reinterpret_cast<double4*>(pts_trial + 4 * (size_t)p) = make_double4(X[0], X[1], X[2], 0.0);
  }
  block_reduce<1, 256>(sc, red);
  if (threadIdx.x == 0) scale_partial[blockIdx.x] = sc[0];
}

}  // namespace vslam
