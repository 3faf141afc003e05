// Bundle adjustment engine: problem structure (host, built once), LM loop (host scalar logic,
// device reductions), C ABI.  Replaces the g2o SparseOptimizer / BlockSolverSE3 /
// OptimizationAlgorithmLevenberg stack driven by reference src/v2/LocalBA.py:20-229.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <limits>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "ba_kernels.cuh"
#include "ba_pcg.cuh"
#include "tuning.h"
#include "ba_window.cuh"
#include "ba_direct.cuh"

namespace vslam {

// (Re)initialises the controller for optimize(max_iterations); keeps `cur`.
__global__ void ba_lm_reset(LMState* st, int max_iterations, int manual) {
  st->dirty = 1; st->done = 0; st->manual = manual; st->first = 1;
  st->iters = 0; st->trials = 0; st->q = 0; st->max_iterations = max_iterations;
  st->last_pcg_status = 0; st->last_pcg_iters = 0;
  st->lambda = 0.0; st->ni = 2.0; st->cur_cost = 0.0; st->initial_cost = 0.0; st->max_diag = 0.0;
  st->pcg_total = 0.0; st->rejected = 0.0; st->breakdowns = 0.0;
  st->last_cost = 0.0; st->last_scale = 0.0; st->last_rho = 0.0;
}
// Stage API pokes: op 0 = force a rebuild, 1 = set lambda, 2 = accept the trial state.
__global__ void ba_lm_poke(LMState* st, int op, double value) {
  if (op == 0) { st->dirty = 1; st->done = 0; st->manual = 1; }
  else if (op == 1) { st->lambda = value; st->done = 0; st->manual = 1; }
  else if (op == 2) { st->cur ^= 1; st->dirty = 1; }
  else if (op == 3) { st->cur = 0; }
}
// Copies the current (which = 0) or trial (1) state out.
__global__ void ba_copy_state(const LMState* __restrict__ st, int which, const double* __restrict__ a0,
                              const double* __restrict__ a1, double* __restrict__ dst, size_t n) {
  const double* src = ((st->cur ^ which) & 1) ? a1 : a0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}

__global__ void ba_add2(const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ o, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = a[i] + b[i];
}

// Public point arrays are [P][3]; the state buffers are padded to [P][4] (see load_point).
__global__ void ba_pad_points(const double* __restrict__ src, double* __restrict__ dst, int P) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < P) *reinterpret_cast<double4*>(dst + 4 * (size_t)p) = make_double4(src[3 * (size_t)p], src[3 * (size_t)p + 1], src[3 * (size_t)p + 2], 0.0);
}
__global__ void ba_unpad_points(const LMState* __restrict__ st, int which, const double* __restrict__ a0,
                                const double* __restrict__ a1, double* __restrict__ dst, int P) {
  const double* src = ((st->cur ^ which) & 1) ? a1 : a0;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < P) { dst[3 * (size_t)p] = src[4 * (size_t)p]; dst[3 * (size_t)p + 1] = src[4 * (size_t)p + 1]; dst[3 * (size_t)p + 2] = src[4 * (size_t)p + 2]; }
}

__global__ void ba_gather_uv(int N, const int* __restrict__ src, const double2* __restrict__ uv,
                             double2* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) out[i] = (src[i] >= 0) ? uv[src[i]] : make_double2(0.0, 0.0);
}

// Device buffer with capacity: a problem object is re-used for a new structure (vslam_ba_set_structure) without
// going back to cudaMalloc while the new sizes fit; growth doubles.  Contents are NOT kept across a growth (every
// structural array is re-uploaded and every state array is re-set after a structure change).
template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0, cap = 0;
  cudaError_t alloc(size_t count) {
    n = count;
    const size_t need = std::max<size_t>(count, 1);
    if (need <= cap) return cudaSuccess;
    if (p) { cudaFree(p); p = nullptr; }
    cap = std::max(need, 2 * cap);
    cudaError_t e = cudaMalloc(&p, cap * sizeof(T));
    if (e != cudaSuccess) { cap = 0; p = nullptr; }
    return e;
  }
  cudaError_t upload(const std::vector<T>& h) {
    cudaError_t e = alloc(h.size());
    if (e != cudaSuccess || h.empty()) return e;
    return cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
  }
  ~DevBuf() { if (p) cudaFree(p); }
};

struct BA {
  int C = 0, P = 0, N = 0, E = 0, Cf = 0, Pf = 0, n = 0, max_chunks = 1;
  int n_items = 0, n_units = 0, n_seg = 0, n_extra = 0, n_sp = 0;
  double4 K = make_double4(0, 0, 0, 0);
  // structure
  DevBuf<int> cam_col, pt_col, cm_start, cm_pt, cm_pm, cm_src, pt_start, pm_cam, pm_src;
  DevBuf<int> item_a, item_b, unit_start, unit_len, seg_a, seg_b, seg_unit_start, seg_extra_idx;
  DevBuf<int> sa, sb, diag_extra, inc_start, inc_edge, extra_a, extra_b;
  // chunked Schur work lists (ba_schur_chunks): per (free camera, chunk of its observations) the partner blocks
  DevBuf<int> ck_cam, ck_start, ck_end, ck_ustart, cu_unit, cu_istart, it_partner, cam_ck_start, seg_cunit_start;
  DevBuf<unsigned short> it_local;
  DevBuf<double> ck_bh, ck_partial;
  int n_chunks = 0, n_cunits = 0;
  // 1 = sorted units over G (default: 323 us at C4), 0 = resident chunks (measured slower: 455 us; with 8-lane unit
  // groups 712 us, with a double-buffered gather at one CTA per SM 1064 us -- latency-bound on its index loads)
  int schur_variant = 1;
  // fused linearisation work lists
  DevBuf<LinWork> lin_work, cam_work;
  DevBuf<unsigned> cam_ticket, lin_ticket;
  DevBuf<int> pw_first, item_row, ell_pt, ell_cam, ell_src;
  DevBuf<double2> ell_uv;
  int n_pt_warps = 0, n_ell_rows = 0, cams_in_smem = 0, lin_ctas = 0;
  size_t lin_smem = 0;
  DevBuf<double> pt_maxd;
  DevBuf<long long> trace;       // allocated by vslam_ba_trace_linearize only
  int lin_nA = 0, lin_nB = 0;
  DevBuf<double> edge_w, edge_es, edge_u;
  DevBuf<double2> cm_uv;
  // state: [0] current, [1] trial
  DevBuf<double> cams[2], pts[2], scale_meas;
  DevBuf<CamRT> camrt[2];
  DevBuf<LMState> lm;            // device-resident LM controller state
  LMState* lm_host = nullptr;    // pinned mirror, valid after ba_sync_state()
  cudaGraphExec_t step_graph = nullptr;   // one LM trial slot (rebuild-if-dirty + trial + controller)
  bool graph_tried = false;
  int optimizes_since_build = 0;          // the step is captured into a graph from the SECOND optimize() of a structure
  cudaGraphExec_t multi_graph = nullptr;  // `multi_slots` LM trial slots in ONE graph (batched optimisation: one launch per problem)
  int multi_slots = 0;
  cudaStream_t graph_stream = nullptr;
  // system
  DevBuf<double> Hpl, Gbuf, Hpp, bp, bhat, Hll, bl, Hinv, Wfac, camp, cam_cost, S, dp, dl, tmp3;
  DevBuf<double> unit_partial, seg_extra_dense, extra, extra_diag, bp_extra, scale_cost, scale_partial, res, p_glob, w_loc, part;
  DevBuf<unsigned> bar;
  DevBuf<int> pcg_flags;  // [0] fail flag, [1] status, [2] iters
  // PCG launch configuration
  int pcg_G = 1, pcg_rows = 6, pcg_slab = 1, pcg_cluster = 0;
  size_t pcg_smem = 0;
  double pcg_tol = 1e-10;
  int pcg_max_iter = 0;
  int linear_solver = 0;   // 0 auto = PCG (measured faster: 22.6 + 7 us against 31 us at n = 54), 1 PCG, 2 in-CTA LDL^T
  bool use_direct() const { return n > 0 && n <= kDirectMax && (linear_solver == 2 || linear_solver == 3); }
  bool use_gj() const { return use_direct() && linear_solver == 2 && n <= kGJMax; }   // 3 forces the LDL^T kernel

  BAView view() const {
    BAView V;
    V.C = C; V.P = P; V.N = N; V.Cf = Cf; V.Pf = Pf; V.E = E; V.max_chunks = max_chunks; V.K = K;
    V.st = lm.p;
    for (int i = 0; i < 2; ++i) { V.cams[i] = cams[i].p; V.pts[i] = pts[i].p; V.camrt[i] = camrt[i].p; }
    V.cam_col = cam_col.p; V.pt_col = pt_col.p; V.cm_start = cm_start.p; V.cm_pt = cm_pt.p;
    V.cm_pm = cm_pm.p; V.cm_uv = cm_uv.p; V.pt_start = pt_start.p; V.pm_cam = pm_cam.p;
    return V;
  }
  LinArgs lin_args() const {
    LinArgs A;
    A.work = lin_work.p; A.cam_work = cam_work.p;
    A.n_pt_ctas = lin_nB; A.n_pt_warps = n_pt_warps; A.n_tickets = C + lin_nB + (E > 0 ? 1 : 0);
    A.cams_in_smem = cams_in_smem; A.has_scale = E > 0 ? 1 : 0;
    A.pw_first = pw_first.p; A.item_row = item_row.p; A.ell_pt = ell_pt.p; A.ell_cam = ell_cam.p;
    A.ell_uv = ell_uv.p;
    A.cam_ticket = cam_ticket.p; A.lin_ticket = lin_ticket.p;
    A.Hpl = Hpl.p; A.camp = camp.p; A.Hpp = Hpp.p; A.bp = bp.p; A.cam_cost = cam_cost.p;
    A.Hll = Hll.p; A.bl = bl.p; A.pt_maxd = pt_maxd.p;
    A.extra_diag = extra_diag.p; A.scale_cost = scale_cost.p;
    A.st_rw = lm.p; A.trace = trace.p; A.force = 0; A.spec = 0;
    A.trial.n_sp = (P ? n_sp : 0) + (C + 255) / 256;   // point blocks + camera blocks of ba_backsub_pt
    A.trial.n = n; A.trial.scale_partial = scale_partial.p; A.trial.dp = dp.p;
    A.trial.bp = bp.p; A.trial.bp_extra = bp_extra.p; A.trial.pcg_status = pcg_flags.p + 1;
    A.trial.pcg_iters = pcg_flags.p + 2;
    return A;
  }
  ScaleView scale_view() const {
    ScaleView V;
    V.st = lm.p; V.cams2[0] = cams[0].p; V.cams2[1] = cams[1].p;
    V.E = E; V.n_extra = n_extra; V.Cf = Cf; V.sa = sa.p; V.sb = sb.p; V.meas = scale_meas.p;
    V.cam_col = cam_col.p; V.inc_start = inc_start.p; V.inc_edge = inc_edge.p; V.extra_a = extra_a.p;
    V.extra_b = extra_b.p; V.edge_w = edge_w.p; V.edge_es = edge_es.p; V.edge_u = edge_u.p;
    V.extra = extra.p; V.bp_extra = bp_extra.p; V.scale_cost = scale_cost.p; V.extra_diag = extra_diag.p;
    return V;
  }
  ~BA() {
    if (multi_graph) cudaGraphExecDestroy(multi_graph);
    if (step_graph) cudaGraphExecDestroy(step_graph);
    if (lm_host) cudaFreeHost(lm_host);
  }
};

// Launch with the programmatic-stream-serialization attribute (see pdl_wait in ba_kernels.cuh); captured
// into the step graph it becomes a programmatic edge.  VSLAM_BA_PDL=0 falls back to plain launches.
static bool use_pdl() {
  static const bool on = [] { const char* e = getenv("VSLAM_BA_PDL"); return !(e && atoi(e) == 0); }();
  return on;
}
template <typename... KArgs, typename... Args>
static void launch_step_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                               Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  // inside a stream capture the kernels stay on plain graph edges: measured faster than programmatic
  // edges (C3: 13.8k vs 12.7k LM it/s); the attribute pays on direct launches (trial 72 -> 62 us)
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cap);
  cfg.attrs = attr; cfg.numAttrs = (use_pdl() && cap == cudaStreamCaptureStatusNone) ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);   // errors surface in check_launch
}

// The dynamic shared memory limit of a kernel is a per-function attribute of the process, not of a launch: every user of
// ba_pcg_kernel (solvers of different sizes, the stand-alone entry points) raises it to the device maximum, so that
// none lowers it under another's feet.
static cudaError_t pcg_kernel_allow_max_smem() {
  int max_smem = 0, dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  e = cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return e;
  cudaFuncAttributes fa{};
  e = cudaFuncGetAttributes(&fa, ba_pcg_kernel);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(ba_pcg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              max_smem - (int)fa.sharedSizeBytes);   // (static + dynamic share the opt-in limit)
}
// The two instances of the cluster PCG kernel: 1 = <2, 48, 8> (n <= 384), 2 = <1, 64, 16> (n <= 480, non-portable cluster size)
using PcgClusterFn = void (*)(PcgArgs);
static PcgClusterFn pcg_cluster_fn(int kind) {
  return kind == 2 ? ba_pcg_cluster_kernel<1, kPcgCluster16MaxCols, kPcgCluster16MaxCtas>
                   : ba_pcg_cluster_kernel<2, kPcgClusterMaxCols, kPcgClusterMaxCtas>;
}
static int pcg_cluster_max_ctas(int kind) { return kind == 2 ? kPcgCluster16MaxCtas : kPcgClusterMaxCtas; }
static cudaError_t pcg_cluster_kernel_allow_max_smem(int kind) {
  int max_smem = 0, dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  e = cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return e;
  cudaFuncAttributes fa{};
  e = cudaFuncGetAttributes(&fa, pcg_cluster_fn(kind));
  if (e != cudaSuccess) return e;
  if (kind == 2) {
    e = cudaFuncSetAttribute(pcg_cluster_fn(kind), cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
  }
  return cudaFuncSetAttribute(pcg_cluster_fn(kind), cudaFuncAttributeMaxDynamicSharedMemorySize,
                              max_smem - (int)fa.sharedSizeBytes);
}
// Plan of the cluster PCG for a system of nb camera blocks (n = 6 nb): which instance, camera blocks per CTA, CTAs, shared
// memory -- or kind 0 when n is out of its range, the switch is off or the device does not place such a cluster.
struct PcgClusterPlan { int kind = 0, blocks = 0, ctas = 0; size_t smem = 0; };
static PcgClusterPlan pcg_cluster_plan_uncached(int nb, size_t budget);
static PcgClusterPlan pcg_cluster_plan(int nb, size_t budget) {
  // (per device and size: the occupancy query and the attribute calls are host work a serving loop should pay once)
  static std::mutex mu;
  static std::unordered_map<long long, PcgClusterPlan> cache;
  int dev = 0;
  cudaGetDevice(&dev);
  const long long key = ((long long)dev << 32) | (unsigned)nb;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
  }
  const PcgClusterPlan P = pcg_cluster_plan_uncached(nb, budget);
  std::lock_guard<std::mutex> lock(mu);
  cache[key] = P;
  return P;
}
static PcgClusterPlan pcg_cluster_plan_uncached(int nb, size_t budget) {
  PcgClusterPlan P;
  const int n = 6 * nb;
  static const int cluster_env = [] { const char* e = getenv("VSLAM_BA_PCG_CLUSTER"); return e ? atoi(e) : -1; }();
  // (measured: up to 40 free cameras, where eight CTAs need one row per lane, the two are level or <2, 48, 8> ahead;
  //  beyond, sixteen CTAs of one row per lane win: 44 cameras 10.5k -> 11.3k LM it/s, 59 cameras 9.9k -> 10.4k)
  const int first_kind = nb > 40 ? 2 : 1;
  for (int pass = 0; pass < 2 && !P.kind; ++pass) {
    const int kind = pass == 0 ? first_kind : 3 - first_kind;
    if (cluster_env == 0 || (cluster_env > 0 && cluster_env != kind)) continue;
    const int max_ctas = pcg_cluster_max_ctas(kind);
    const int max_rows = kind == 2 ? kPcgCluster16MaxRows : kPcgClusterMaxRows;
    const int max_n = kind == 2 ? kPcgCluster16MaxN : kPcgClusterMaxN;
    if (n <= kPcgMidMax || n > max_n) continue;
    const int t = (nb + max_ctas - 1) / max_ctas;
    const size_t need = pcg_cluster_smem_doubles(n, 6 * t, max_ctas) * sizeof(double);
    if (6 * t > max_rows || need > budget || pcg_cluster_kernel_allow_max_smem(kind) != cudaSuccess) { cudaGetLastError(); continue; }
    // the device must be able to place such a cluster (partitioned or partly busy devices may not): else the
    // cooperative kernel takes the system
    cudaLaunchConfig_t cfg = {};
    const int Gc = (nb + t - 1) / t;
    cfg.gridDim = dim3(Gc); cfg.blockDim = dim3(kPcgThreads); cfg.dynamicSmemBytes = need;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)Gc; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int n_clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&n_clusters, pcg_cluster_fn(kind), &cfg) == cudaSuccess && n_clusters >= 1) {
      P.kind = kind; P.blocks = t; P.ctas = Gc; P.smem = need;
    } else {
      cudaGetLastError();
    }
  }
  return P;
}
static cudaError_t launch_pcg_cluster(int kind, int ctas, size_t smem, const PcgArgs& A, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(kPcgThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)ctas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cap);
  cfg.attrs = attr; cfg.numAttrs = (use_pdl() && cap == cudaStreamCaptureStatusNone) ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, pcg_cluster_fn(kind), A);
}
static size_t pcg_smem_budget() {
  int max_smem = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return (size_t)std::max(max_smem - 2048, 16384);
}
// n <= 64: the pipelined single-CTA kernel unless VSLAM_BA_PCG_SMALL_PIPELINED=0
static PcgClusterFn pcg_small_fn() {
  static const bool pipelined = [] { const char* e = getenv("VSLAM_BA_PCG_SMALL_PIPELINED"); return !(e && atoi(e) == 0); }();
  return pipelined ? ba_pcg_small_pipelined_kernel : ba_pcg_small_kernel;
}
static size_t pcg_small_smem(int n) { return sizeof(double) * ((size_t)n * n + 6 * (size_t)n + 16 * 64); }   // S, M^-1, r/u copies, partials

static void counting_sort(const std::vector<int>& in, const int* key, int n_keys, std::vector<int>& out) {
  std::vector<int> cnt(n_keys + 1, 0);
  for (int i : in) cnt[key[i] + 1]++;
  for (int k = 0; k < n_keys; ++k) cnt[k + 1] += cnt[k];
  out.resize(in.size());
  for (int i : in) out[cnt[key[i]]++] = i;
}

#define BA_TRY(expr)                                                   \
  do {                                                                 \
    cudaError_t _e = (expr);                                           \
    if (_e != cudaSuccess) { set_cuda_error(_e, #expr); return VSLAM_ERR_CUDA; } \
  } while (0)

static int ba_build(BA& B, int C, int P, int N, int E, const int32_t* obs_cam, const int32_t* obs_pt,
                    const uint8_t* cam_fixed, const uint8_t* pt_fixed, const int32_t* sa,
                    const int32_t* sb) {
  if (B.step_graph) { cudaGraphExecDestroy(B.step_graph); B.step_graph = nullptr; }
    if (B.multi_graph) { cudaGraphExecDestroy(B.multi_graph); B.multi_graph = nullptr; B.multi_slots = 0; }
  B.graph_tried = false; B.optimizes_since_build = 0;
  B.K = make_double4(0, 0, 0, 0);
  B.C = C; B.P = P; B.N = N; B.E = E;
  std::vector<int> cam_col(C), pt_col(P);
  int Cf = 0, Pf = 0;
  for (int c = 0; c < C; ++c) cam_col[c] = (cam_fixed && cam_fixed[c]) ? -1 : Cf++;
  for (int p = 0; p < P; ++p) pt_col[p] = (pt_fixed && pt_fixed[p]) ? -1 : Pf++;
  B.Cf = Cf; B.Pf = Pf; B.n = 6 * Cf;
  for (int i = 0; i < N; ++i)
    if (obs_cam[i] < 0 || obs_cam[i] >= C || obs_pt[i] < 0 || obs_pt[i] >= P) return VSLAM_ERR_BAD_ARG;
  for (int k = 0; k < E; ++k)
    if (sa[k] < 0 || sa[k] >= C || sb[k] < 0 || sb[k] >= C || sa[k] == sb[k]) return VSLAM_ERR_BAD_ARG;
  // two stable counting sorts per ordering
  std::vector<int> ident(N), by_pt, by_cam, cm_src, pm_src;
  for (int i = 0; i < N; ++i) ident[i] = i;
  counting_sort(ident, obs_pt, P, by_pt);
  counting_sort(by_pt, obs_cam, C, cm_src);   // (cam, pt)
  counting_sort(ident, obs_cam, C, by_cam);
  counting_sort(by_cam, obs_pt, P, pm_src);   // (pt, cam)
  std::vector<int> pm_pos(N);
  for (int i = 0; i < N; ++i) pm_pos[pm_src[i]] = i;
  std::vector<int> cm_pos(N);
  for (int i = 0; i < N; ++i) cm_pos[cm_src[i]] = i;
  std::vector<int> cm_start(C + 1, 0), pt_start(P + 1, 0), cm_pt(N), cm_pm(N), pm_cam(N);
  for (int i = 0; i < N; ++i) { cm_start[obs_cam[i] + 1]++; pt_start[obs_pt[i] + 1]++; }
  for (int c = 0; c < C; ++c) cm_start[c + 1] += cm_start[c];
  for (int p = 0; p < P; ++p) pt_start[p + 1] += pt_start[p];
  int max_obs = 0;
  for (int c = 0; c < C; ++c) max_obs = std::max(max_obs, cm_start[c + 1] - cm_start[c]);
  B.max_chunks = std::max(1, (max_obs + kCamChunk - 1) / kCamChunk);
  for (int i = 0; i < N; ++i) {
    cm_pt[i] = obs_pt[cm_src[i]];
    cm_pm[i] = pm_pos[cm_src[i]];
    pm_cam[i] = obs_cam[pm_src[i]];
  }
  // Schur items, counted then filled per (col_a <= col_b) key
  if ((long long)Cf * Cf > (1ll << 26)) return VSLAM_ERR_BAD_ARG;  // dense reduced system limit
  std::vector<long long> cnt((size_t)Cf * Cf + 1, 0);
  auto for_items = [&](auto&& fn) {
    for (int p = 0; p < P; ++p) {
      if (pt_col[p] < 0) continue;
      for (int i = pt_start[p]; i < pt_start[p + 1]; ++i) {
        const int ci = cam_col[pm_cam[i]];
        if (ci < 0) continue;
        for (int j = i; j < pt_start[p + 1]; ++j) {
          const int cj = cam_col[pm_cam[j]];
          if (cj < 0) continue;
          fn((size_t)ci * Cf + cj, cm_pos[pm_src[i]], cm_pos[pm_src[j]]);
          if (i != j && ci == cj) fn((size_t)ci * Cf + cj, cm_pos[pm_src[j]], cm_pos[pm_src[i]]);
        }
      }
    }
  };
  for_items([&](size_t key, int, int) { cnt[key + 1]++; });
  std::vector<char> present((size_t)Cf * Cf, 0);
  for (size_t k = 0; k < (size_t)Cf * Cf; ++k) present[k] = cnt[k + 1] > 0;
  for (int c = 0; c < Cf; ++c) present[(size_t)c * Cf + c] = 1;
  for (int k = 0; k < E; ++k) {
    const int ca = cam_col[sa[k]], cb = cam_col[sb[k]];
    if (ca >= 0 && cb >= 0) present[(size_t)std::min(ca, cb) * Cf + std::max(ca, cb)] = 1;
  }
  for (size_t k = 0; k < (size_t)Cf * Cf; ++k) cnt[k + 1] += cnt[k];
  const long long n_items = cnt[(size_t)Cf * Cf];
  if (n_items > (1ll << 31) - 1) return VSLAM_ERR_BAD_ARG;
  std::vector<int> item_a(n_items), item_b(n_items);
  {
    std::vector<long long> fill(cnt.begin(), cnt.end() - 1);
    for_items([&](size_t key, int a, int b) { long long o = fill[key]++; item_a[o] = a; item_b[o] = b; });
  }
  std::vector<int> seg_a, seg_b, seg_unit_start, unit_start, unit_len, seg_of_key((size_t)Cf * Cf, -1);
  // items per warp work unit: kUnitItems, or more when the problem still offers every resident warp several units
  // (a unit pays ~600 instructions of prologue + 36-value reduction; C4's blocks hold ~300 items, so 256-item units
  // split most of them into a full unit and a one-round remainder)
  static const int unit_env = [] { const char* e = getenv("VSLAM_BA_UNIT_ITEMS"); return e ? atoi(e) : 0; }();
  // ... and fewer when it is small: a local window (C3: 105k items) wants about one unit per resident warp
  // (64-item units: Schur stage 30.7 -> 22.6 us, measured; 32-item units pay for their reductions again)
  const long long warp_slots = (long long)std::max(1, sm_count()) * 12;
  long long unit_items = (n_items / warp_slots + 31) / 32 * 32;
  unit_items = std::max<long long>(kUnitItemsSmall, std::min<long long>(unit_items, kUnitItemsLarge));
  if (unit_env >= 32) unit_items = unit_env / 32 * 32;
  for (int a = 0; a < Cf; ++a)
    for (int b = a; b < Cf; ++b) {
      const size_t key = (size_t)a * Cf + b;
      if (!present[key]) continue;
      seg_of_key[key] = (int)seg_a.size();
      seg_a.push_back(a); seg_b.push_back(b);
      seg_unit_start.push_back((int)unit_start.size());
      for (long long s = cnt[key]; s < cnt[key + 1]; s += unit_items) {
        const int len = (int)std::min<long long>(unit_items, cnt[key + 1] - s);
        // bit 30: every item of the unit pairs an observation with itself (a diagonal block without duplicate
        // observations): the kernel then fetches each block once
        bool self = a == b;
        for (long long i = s; self && i < s + len; ++i) self = item_a[i] == item_b[i];
        unit_start.push_back((int)s);
        unit_len.push_back(len | (self ? (1 << 30) : 0));
      }
    }
  seg_unit_start.push_back((int)unit_start.size());
  B.n_items = (int)n_items; B.n_units = (int)unit_start.size(); B.n_seg = (int)seg_a.size();
  // ---- chunked Schur lists: CTA = (free camera a, kSchurChunk of its camera-major observations).  The chunk's
  // Y_o = Hpl_o (Hll + lambda I)^-1 stay in shared memory; an item pairs a resident observation with a partner
  // observation o' of the same point whose camera column is >= a's (its own included); items are sorted by the
  // partner's column, one unit per (chunk, column) -> partial 6x6 block; units are numbered in (a, b, chunk) order
  // so that the blocks' partials are consecutive (seg_cunit_start).
  std::vector<int> ck_cam, ck_start, ck_end, ck_ustart, cu_unit, cu_istart, it_partner, cam_ck_start(C + 1, 0), seg_cunit_start;
  std::vector<unsigned short> it_local;
  {
    std::vector<int> chunk_first(C, 0);
    for (int c = 0; c < C; ++c) {
      cam_ck_start[c] = (int)ck_cam.size();
      if (cam_col[c] < 0) continue;
      for (int s = cm_start[c]; s < cm_start[c + 1]; s += kSchurChunk) {
        ck_cam.push_back(c); ck_start.push_back(s); ck_end.push_back(std::min(cm_start[c + 1], s + kSchurChunk));
      }
    }
    cam_ck_start[C] = (int)ck_cam.size();
    const int n_ck = (int)ck_cam.size();
    // pass 1: which (chunk, partner column) pairs have items -> unit numbering in (a, b, chunk) order
    std::vector<std::vector<int>> ck_cols(n_ck);            // sorted partner columns with >= 1 item
    std::vector<int> mark(Cf, -1);
    auto for_chunk_items = [&](int ck, auto&& fn) {
      const int ca = cam_col[ck_cam[ck]];
      for (int o = ck_start[ck]; o < ck_end[ck]; ++o) {
        const int p = cm_pt[o];
        if (pt_col[p] < 0) continue;
        for (int q = pt_start[p]; q < pt_start[p + 1]; ++q) {
          const int cb = cam_col[pm_cam[q]];
          if (cb < ca) continue;                            // (fixed cameras have column -1)
          fn(o - ck_start[ck], cm_pos[pm_src[q]], cb);
        }
      }
    };
    for (int ck = 0; ck < n_ck; ++ck) {
      for_chunk_items(ck, [&](int, int, int cb) { if (mark[cb] != ck) { mark[cb] = ck; ck_cols[ck].push_back(cb); } });
      std::sort(ck_cols[ck].begin(), ck_cols[ck].end());
    }
    // units of block (a, b): the chunks of camera a that have items with column b, in chunk order
    std::vector<int> seg_cunits(B.n_seg, 0);
    for (int ck = 0; ck < n_ck; ++ck) {
      const int ca = cam_col[ck_cam[ck]];
      for (int cb : ck_cols[ck]) seg_cunits[seg_of_key[(size_t)ca * Cf + cb]]++;
    }
    seg_cunit_start.assign(B.n_seg + 1, 0);
    for (int s = 0; s < B.n_seg; ++s) seg_cunit_start[s + 1] = seg_cunit_start[s] + seg_cunits[s];
    std::vector<int> seg_fill(seg_cunit_start.begin(), seg_cunit_start.end() - 1);
    // pass 2: items per chunk sorted by partner column (counting sort over the chunk's columns)
    std::vector<int> col_slot(Cf, -1), cnt_c, fill_c;
    ck_ustart.assign(n_ck + 1, 0);
    for (int ck = 0; ck < n_ck; ++ck) {
      const int ca = cam_col[ck_cam[ck]];
      const int nc = (int)ck_cols[ck].size();
      for (int i = 0; i < nc; ++i) col_slot[ck_cols[ck][i]] = i;
      cnt_c.assign(nc + 1, 0);
      for_chunk_items(ck, [&](int, int, int cb) { cnt_c[col_slot[cb] + 1]++; });
      for (int i = 0; i < nc; ++i) cnt_c[i + 1] += cnt_c[i];
      const size_t base = it_partner.size();
      it_partner.resize(base + cnt_c[nc]); it_local.resize(base + cnt_c[nc]);
      fill_c.assign(cnt_c.begin(), cnt_c.end() - 1);
      for_chunk_items(ck, [&](int loc, int partner, int cb) {
        const size_t at = base + fill_c[col_slot[cb]]++;
        it_local[at] = (unsigned short)loc; it_partner[at] = partner;
      });
      for (int i = 0; i < nc; ++i) {
        const int seg = seg_of_key[(size_t)ca * Cf + ck_cols[ck][i]];
        cu_unit.push_back(seg_fill[seg]++);
        cu_istart.push_back((int)(base + cnt_c[i]));
      }
      ck_ustart[ck + 1] = (int)cu_unit.size();
    }
    cu_istart.push_back((int)it_partner.size());
    if (it_partner.size() > (size_t)((1ll << 31) - 1)) return VSLAM_ERR_BAD_ARG;
    B.n_chunks = n_ck; B.n_cunits = (int)cu_unit.size();
  }
  // extra (scale-edge) blocks: one per touched segment; incident-edge CSR per free camera
  std::vector<int> seg_extra_idx(B.n_seg, -1), diag_extra(Cf, -1), extra_a, extra_b;
  int n_extra = 0;
  auto touch = [&](int a, int b) {
    const int lo = std::min(a, b), hi = std::max(a, b);
    int seg = seg_of_key[(size_t)lo * Cf + hi];
    if (seg_extra_idx[seg] < 0) { seg_extra_idx[seg] = n_extra++; extra_a.push_back(lo); extra_b.push_back(hi); }
    return seg_extra_idx[seg];
  };
  std::vector<int> inc_start(Cf + 1, 0), inc_edge;
  for (int k = 0; k < E; ++k) {
    const int ca = cam_col[sa[k]], cb = cam_col[sb[k]];
    if (ca >= 0) { diag_extra[ca] = touch(ca, ca); inc_start[ca + 1]++; }
    if (cb >= 0) { diag_extra[cb] = touch(cb, cb); inc_start[cb + 1]++; }
    if (ca >= 0 && cb >= 0) touch(ca, cb);
  }
  for (int c = 0; c < Cf; ++c) inc_start[c + 1] += inc_start[c];
  inc_edge.resize(inc_start[Cf]);
  {
    std::vector<int> fill(inc_start.begin(), inc_start.end() - 1);
    for (int k = 0; k < E; ++k) {
      const int ca = cam_col[sa[k]], cb = cam_col[sb[k]];
      if (ca >= 0) inc_edge[fill[ca]++] = k;
      if (cb >= 0) inc_edge[fill[cb]++] = k;
    }
  }
  B.n_extra = n_extra;
  B.n_sp = (P + 255) / 256;
  // fused linearisation: camera chunks (every camera owns at least one, possibly empty), point blocks
  // and the scale-edge CTA, interleaved in proportion so that all roles are resident at the same time
  // small problems get smaller chunks (fewer observations per thread, more CTAs in flight): the grid
  // should offer about four CTAs per SM before a thread takes a second observation
  const int slots = 4 * std::max(1, sm_count());
  const int obs_iters = std::min(kLinObs, std::max(1, (N + kLinThreads * slots - 1) / (kLinThreads * slots)));
  const int lin_chunk = kLinThreads * obs_iters;
  const int pt_rows_target = std::min(kPtRowsPerWarp, std::max(2, (N / 32 + 2 * slots * kLinWarps - 1) / (2 * slots * kLinWarps)));
  std::vector<LinWork> cam_work, lin_work;
  {
    int row = 0;
    for (int c = 0; c < C; ++c) {
      const int n_obs_c = cm_start[c + 1] - cm_start[c];
      const int nch = std::max(1, (n_obs_c + lin_chunk - 1) / lin_chunk);
      for (int k = 0; k < nch; ++k) {
        LinWork w{};
        w.kind = 0; w.cam = c; w.start = cm_start[c] + k * lin_chunk;
        w.end = std::min(cm_start[c + 1], w.start + lin_chunk);
        w.row = row + k; w.n_chunks = nch; w.first_row = row; w.col = cam_col[c];
        cam_work.push_back(w);
      }
      row += nch;
    }
  }
  B.lin_nA = (int)cam_work.size();
  // point role, ELL layout: free points with observations sorted by degree (descending, stable), 32
  // per item (one per lane); row j of an item holds the j-th observation of each of its points
  std::vector<int> pw_first, item_row, ell_pt, ell_cam, ell_src;
  {
    std::vector<int> fp;
    for (int p = 0; p < P; ++p)
      if (pt_col[p] >= 0 && pt_start[p + 1] > pt_start[p]) fp.push_back(p);
    // ... then by camera list: the 32 points of an item then mostly see the SAME camera in a row, and
    // equal shared-memory addresses are a broadcast instead of a bank conflict in the camera table
    std::stable_sort(fp.begin(), fp.end(), [&](int a, int b) {
      const int da = pt_start[a + 1] - pt_start[a], db = pt_start[b + 1] - pt_start[b];
      if (da != db) return da > db;
      return std::lexicographical_compare(pm_cam.begin() + pt_start[a], pm_cam.begin() + pt_start[a + 1],
                                          pm_cam.begin() + pt_start[b], pm_cam.begin() + pt_start[b + 1]);
    });
    const int n_it = ((int)fp.size() + 31) / 32;
    item_row.assign(n_it + 1, 0);
    ell_pt.assign((size_t)n_it * 32, -1);
    for (int it = 0; it < n_it; ++it) {
      const int p0 = fp[(size_t)it * 32];
      item_row[it + 1] = item_row[it] + (pt_start[p0 + 1] - pt_start[p0]);
    }
    ell_cam.assign((size_t)item_row[n_it] * 32, -1);
    ell_src.assign((size_t)item_row[n_it] * 32, -1);
    for (size_t q = 0; q < fp.size(); ++q) {
      const int p = fp[q], it = (int)(q / 32), l = (int)(q % 32);
      ell_pt[(size_t)it * 32 + l] = p;
      for (int j = 0; j < pt_start[p + 1] - pt_start[p]; ++j) {
        ell_cam[((size_t)item_row[it] + j) * 32 + l] = pm_cam[pt_start[p] + j];
        ell_src[((size_t)item_row[it] + j) * 32 + l] = pm_src[pt_start[p] + j];
      }
    }
    // warps take runs of items of about kPtRowsPerWarp rows
    pw_first.push_back(0);
    int rows = 0;
    for (int it = 0; it < n_it; ++it) {
      rows += item_row[it + 1] - item_row[it];
      if (rows >= pt_rows_target || it + 1 == n_it) { pw_first.push_back(it + 1); rows = 0; }
    }
    B.n_pt_warps = (int)pw_first.size() - 1;
    B.n_ell_rows = item_row[n_it];
  }
  B.lin_nB = (B.n_pt_warps + kLinWarps - 1) / kLinWarps;
  B.cams_in_smem = (C <= kCamSmemMax) ? 1 : 0;
  B.lin_smem = std::max((size_t)kLinCamSmem,
                        B.cams_in_smem ? (size_t)C * kCamTabStride * sizeof(double) : 0);
  BA_TRY(cudaFuncSetAttribute(ba_schur_units, cudaFuncAttributeMaxDynamicSharedMemorySize, kSchurSmem));
  BA_TRY(cudaFuncSetAttribute(ba_schur_chunks, cudaFuncAttributeMaxDynamicSharedMemorySize, kChunkSmem));
  BA_TRY(cudaFuncSetAttribute(ba_linearize_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)(kCamSmemMax * kCamTabStride * sizeof(double))));
  {
    if (E > 0) { LinWork w{}; w.kind = 2; lin_work.push_back(w); }
    const long long tot = (long long)B.lin_nA + B.lin_nB;
    int a = 0, b = 0;
    for (long long i = 0; i < tot; ++i) {
      // point block when the running share of point blocks falls behind
      const bool want_b = (long long)(b + 1) * tot <= (i + 1) * (long long)B.lin_nB;
      if ((want_b && b < B.lin_nB) || a >= B.lin_nA) { LinWork w{}; w.kind = 1; w.cam = b++; lin_work.push_back(w); }
      else lin_work.push_back(cam_work[a++]);
    }
    B.lin_ctas = (int)lin_work.size();
  }
  // uploads
  BA_TRY(B.cam_col.upload(cam_col)); BA_TRY(B.pt_col.upload(pt_col));
  BA_TRY(B.cm_start.upload(cm_start)); BA_TRY(B.cm_pt.upload(cm_pt)); BA_TRY(B.cm_pm.upload(cm_pm));
  BA_TRY(B.cm_src.upload(cm_src)); BA_TRY(B.pt_start.upload(pt_start)); BA_TRY(B.pm_cam.upload(pm_cam));
  BA_TRY(B.pm_src.upload(pm_src)); BA_TRY(B.item_a.upload(item_a)); BA_TRY(B.item_b.upload(item_b));
  BA_TRY(B.unit_start.upload(unit_start)); BA_TRY(B.unit_len.upload(unit_len));
  BA_TRY(B.ck_cam.upload(ck_cam)); BA_TRY(B.ck_start.upload(ck_start)); BA_TRY(B.ck_end.upload(ck_end));
  BA_TRY(B.ck_ustart.upload(ck_ustart)); BA_TRY(B.cu_unit.upload(cu_unit)); BA_TRY(B.cu_istart.upload(cu_istart));
  BA_TRY(B.it_partner.upload(it_partner)); BA_TRY(B.it_local.upload(it_local)); BA_TRY(B.cam_ck_start.upload(cam_ck_start));
  BA_TRY(B.seg_cunit_start.upload(seg_cunit_start));
  BA_TRY(B.ck_bh.alloc((size_t)std::max(B.n_chunks, 1) * 6)); BA_TRY(B.ck_partial.alloc((size_t)std::max(B.n_cunits, 1) * 36));
  BA_TRY(B.seg_a.upload(seg_a)); BA_TRY(B.seg_b.upload(seg_b));
  BA_TRY(B.seg_unit_start.upload(seg_unit_start)); BA_TRY(B.seg_extra_idx.upload(seg_extra_idx));
  BA_TRY(B.sa.upload(std::vector<int>(sa, sa + E))); BA_TRY(B.sb.upload(std::vector<int>(sb, sb + E)));
  BA_TRY(B.inc_start.upload(inc_start)); BA_TRY(B.inc_edge.upload(inc_edge));
  BA_TRY(B.extra_a.upload(extra_a)); BA_TRY(B.extra_b.upload(extra_b));
  BA_TRY(B.edge_w.alloc(E)); BA_TRY(B.edge_es.alloc(E)); BA_TRY(B.edge_u.alloc((size_t)E * 3));
  BA_TRY(B.diag_extra.upload(diag_extra));
  BA_TRY(B.lin_work.upload(lin_work)); BA_TRY(B.cam_work.upload(cam_work));
  BA_TRY(B.pw_first.upload(pw_first)); BA_TRY(B.item_row.upload(item_row)); BA_TRY(B.ell_pt.upload(ell_pt));
  BA_TRY(B.ell_cam.upload(ell_cam)); BA_TRY(B.ell_src.upload(ell_src)); BA_TRY(B.ell_uv.alloc(ell_cam.size()));
  BA_TRY(B.cam_ticket.alloc(C)); BA_TRY(B.lin_ticket.alloc(1));
  BA_TRY(B.pt_maxd.alloc(std::max(B.lin_nB, 1)));
  BA_TRY(cudaMemset(B.cam_ticket.p, 0, (size_t)C * sizeof(unsigned)));
  BA_TRY(cudaMemset(B.lin_ticket.p, 0, sizeof(unsigned)));
  BA_TRY(B.cm_uv.alloc(N)); BA_TRY(B.scale_meas.alloc(E));
  for (int s = 0; s < 2; ++s) {
    BA_TRY(B.cams[s].alloc((size_t)C * 7)); BA_TRY(B.pts[s].alloc((size_t)P * 4)); BA_TRY(B.camrt[s].alloc(C));   // points padded to 32 bytes
  }
  BA_TRY(B.Hpl.alloc((size_t)N * 18)); BA_TRY(B.Gbuf.alloc((size_t)N * 18));
  BA_TRY(B.Hpp.alloc((size_t)Cf * 36)); BA_TRY(B.bp.alloc(B.n)); BA_TRY(B.bhat.alloc(B.n));
  BA_TRY(B.Hll.alloc((size_t)Pf * 6)); BA_TRY(B.bl.alloc((size_t)Pf * 3)); BA_TRY(B.Hinv.alloc((size_t)Pf * 6)); BA_TRY(B.Wfac.alloc((size_t)Pf * 6));
  BA_TRY(B.camp.alloc(std::max((size_t)C * B.max_chunks, (size_t)B.lin_nA) * kCamVals)); BA_TRY(B.cam_cost.alloc(C));
  BA_TRY(B.S.alloc((size_t)B.n * B.n)); BA_TRY(B.dp.alloc(B.n)); BA_TRY(B.dl.alloc((size_t)Pf * 3));
  BA_TRY(B.tmp3.alloc((size_t)N * 4));   // [N][4]: one 32-byte sector per observation
  BA_TRY(cudaMemset(B.tmp3.p, 0, std::max<size_t>((size_t)N * 4, 1) * sizeof(double)));   // rows of fixed cameras / points stay 0
  BA_TRY(B.unit_partial.alloc((size_t)B.n_units * 36));
  BA_TRY(B.extra.alloc((size_t)std::max(n_extra, 1) * 36)); BA_TRY(B.bp_extra.alloc(std::max(B.n, 1))); BA_TRY(B.extra_diag.alloc(std::max(3 * Cf, 1)));
  BA_TRY(B.scale_cost.alloc(1)); BA_TRY(B.scale_partial.alloc((size_t)B.n_sp + (C + 255) / 256)); BA_TRY(B.res.alloc(8));
  BA_TRY(B.p_glob.alloc(2 * (size_t)std::max(B.n, 1))); BA_TRY(B.w_loc.alloc(std::max(B.n, 1)));   // (two exchange buffers)
  BA_TRY(cudaMemset(B.Hll.p, 0, std::max<size_t>((size_t)Pf * 6, 1) * sizeof(double)));   // free points without
  BA_TRY(cudaMemset(B.bl.p, 0, std::max<size_t>((size_t)Pf * 3, 1) * sizeof(double)));    // observations stay 0
  BA_TRY(cudaMemset(B.S.p, 0, std::max<size_t>((size_t)B.n * B.n, 1) * sizeof(double)));
  BA_TRY(cudaMemset(B.extra.p, 0, (size_t)std::max(n_extra, 1) * 36 * sizeof(double)));
  BA_TRY(cudaMemset(B.bp_extra.p, 0, (size_t)std::max(B.n, 1) * sizeof(double)));
  BA_TRY(cudaMemset(B.extra_diag.p, 0, (size_t)std::max(3 * Cf, 1) * sizeof(double)));
  BA_TRY(cudaMemset(B.scale_cost.p, 0, sizeof(double)));
  BA_TRY(cudaMemset(B.camp.p, 0, B.camp.n * sizeof(double)));
  BA_TRY(B.lm.alloc(1));
  BA_TRY(cudaMemset(B.lm.p, 0, sizeof(LMState)));
  if (!B.lm_host) BA_TRY(cudaMallocHost(&B.lm_host, sizeof(LMState)));
  std::memset(B.lm_host, 0, sizeof(LMState));
  // PCG configuration: as many CTAs as keep their slab of S in shared memory, at most one per SM
  {
    const int nb = Cf;
    const int sms = std::max(1, sm_count());
    int max_smem = 0, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    const size_t budget = (size_t)std::max(max_smem - 2048, 16384);
    auto smem_for = [&](int rows, bool slab) {
      return sizeof(double) * ((size_t)B.n + 10 * (size_t)rows + 6 * (size_t)rows + 8 +
                               (slab ? (size_t)rows * B.n : 0));
    };
    int bpc = std::max(1, (nb + sms - 1) / sms);  // blocks per CTA with one CTA per SM
    // As many CTAs as there are camera blocks, up to 128: with one grid barrier per iteration the local product is what
    // is left to shrink, and the barrier costs about the same for 30 as for 120 arrivals.  (Measured, optimize(10):
    // 29 free cameras 4.3k -> 6.0k LM it/s against two CTAs of 150 rows, 59 cameras 3.7k -> 4.7k, 119 cameras
    // 2.9k -> 3.3k, C4 (199 cameras, 100 CTAs instead of 67) 1.90k -> 1.93k.)
    int best = std::max(bpc, (nb + 127) / 128);
    while (best > bpc && smem_for(6 * best, true) > budget) --best;
    if (B.n <= kPcgMidMax) best = std::max(nb, 1);   // one CTA: ba_pcg_small_kernel / ba_pcg_mid_kernel keep S in registers
    if (const char* e = std::getenv("VSLAM_BA_PCG_BLOCKS")) {   // (experiments: camera blocks per PCG CTA)
      const int t = std::atoi(e);
      if (t >= bpc && t <= nb && smem_for(6 * t, true) <= budget) best = t;
    }
    // 22 .. 80 free cameras: one thread-block cluster (ba_pcg_cluster_kernel), exchange through DSMEM
    {
      const PcgClusterPlan P = pcg_cluster_plan(nb, budget);
      B.pcg_cluster = P.kind;
      if (P.kind) best = P.blocks;
    }
    B.pcg_slab = smem_for(6 * best, true) <= budget ? 1 : 0;
    B.pcg_rows = 6 * std::max(best, 1);
    B.pcg_G = std::max(1, (nb + best - 1) / std::max(best, 1));
    B.pcg_smem = B.pcg_cluster ? pcg_cluster_smem_doubles(B.n, B.pcg_rows, pcg_cluster_max_ctas(B.pcg_cluster)) * sizeof(double) : smem_for(B.pcg_rows, B.pcg_slab);
    if (B.pcg_smem > budget) return VSLAM_ERR_BAD_ARG;
    BA_TRY(pcg_kernel_allow_max_smem());
    BA_TRY(cudaFuncSetAttribute(ba_solve_direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)direct_smem(kDirectMax)));
  }
  BA_TRY(B.part.alloc((size_t)6 * B.pcg_G)); BA_TRY(B.bar.alloc(1)); BA_TRY(B.pcg_flags.alloc(4));
  BA_TRY(cudaMemset(B.bar.p, 0, sizeof(unsigned))); BA_TRY(cudaMemset(B.pcg_flags.p, 0, 4 * sizeof(int)));
  B.pcg_max_iter = std::max(200, 20 * B.n);
  return VSLAM_OK;
}

static int ba_set_state(BA& B, const double* cams, const double* pts, const double* uv,
                        const double* scale_meas, const double* K, cudaStream_t st) {
  const double4 Knew = make_double4(K[0], K[1], K[2], K[3]);
  if (Knew.x != B.K.x || Knew.y != B.K.y || Knew.z != B.K.z || Knew.w != B.K.w) {
    if (B.step_graph) { cudaGraphExecDestroy(B.step_graph); B.step_graph = nullptr; }
    if (B.multi_graph) { cudaGraphExecDestroy(B.multi_graph); B.multi_graph = nullptr; B.multi_slots = 0; }  // K is baked into the nodes
    B.graph_tried = false;
  }
  B.K = Knew;
  ba_lm_reset<<<1, 1, 0, st>>>(B.lm.p, 0, 1);
  ba_lm_poke<<<1, 1, 0, st>>>(B.lm.p, 3, 0.0);
  BA_TRY(cudaMemcpyAsync(B.cams[0].p, cams, (size_t)B.C * 7 * sizeof(double), cudaMemcpyDeviceToDevice, st));
  if (B.P) ba_pad_points<<<(B.P + 255) / 256, 256, 0, st>>>(pts, B.pts[0].p, B.P);
  if (B.E) BA_TRY(cudaMemcpyAsync(B.scale_meas.p, scale_meas, (size_t)B.E * sizeof(double), cudaMemcpyDeviceToDevice, st));
  if (B.N) {
    const int blocks = (B.N + 255) / 256;
    ba_gather_uv<<<blocks, 256, 0, st>>>(B.N, B.cm_src.p, reinterpret_cast<const double2*>(uv), B.cm_uv.p);
    const int nx = B.n_ell_rows * 32;
    if (nx) ba_gather_uv<<<(nx + 255) / 256, 256, 0, st>>>(nx, B.ell_src.p, reinterpret_cast<const double2*>(uv), B.ell_uv.p);
  }
  if (B.C) ba_prepare_cams<<<(B.C + 127) / 128, 128, 0, st>>>(B.cams[0].p, B.camrt[0].p, B.C);
  return check_launch("ba_set_state");
}

static int ba_sync_state(BA& B, cudaStream_t st) {
  BA_TRY(cudaMemcpyAsync(B.lm_host, B.lm.p, sizeof(LMState), cudaMemcpyDeviceToHost, st));
  BA_TRY(cudaStreamSynchronize(st));
  return VSLAM_OK;
}

// buildSystem() on the current state (every kernel returns at once unless LMState::dirty).
static int enqueue_linearize(BA& B, cudaStream_t st) {
  const BAView V = B.view();
  dim3 grid(B.max_chunks, B.C);
  launch_step_kernel(ba_linearize_kernel<false>, B.lin_ctas, kLinThreads, B.lin_smem, st, V, B.lin_args(), B.scale_view());
  return check_launch("ba_linearize");
}

// Schur complement of the damped system: Hinv = W W^T, G = Hpl W, bhat, S = Hpp - sum G G^T
// (sorted-segment reduction).  A small system is then solved by the in-CTA direct kernel from S, bhat in HBM when
// that solver is selected.  (Measured and dropped: letting the last unit of a block finish S and the last chunk of
// a camera finish bhat, with W formed on the fly -- two launches instead of four.  The launches overlap under
// programmatic dependent launch anyway: C3 gained 1 us per trial, C4 lost 23 us to the ticket / fence tails of its
// 31k units.)
static int enqueue_schur(BA& B, cudaStream_t st) {
  const BAView V = B.view();
  const LMState* lm = B.lm.p;
  dim3 grid(B.max_chunks, B.C);
  const bool chunked = B.schur_variant == 0;
  if (B.Pf && chunked) {
    if (B.n_chunks) {
      ChunkArgs A{B.ck_cam.p, B.ck_start.p, B.ck_end.p, B.ck_ustart.p, B.cu_unit.p, B.cu_istart.p, B.it_local.p,
                  B.it_partner.p, B.Hpl.p, B.Hll.p, B.bl.p, B.ck_bh.p, B.ck_partial.p};
      launch_step_kernel(ba_schur_chunks, B.n_chunks, kChunkThreads, kChunkSmem, st, V, A);
    }
  } else if (B.Pf) {
    // small problems form W inside the G pass (same statements, one launch fewer: 2.3 us of a 54 us LM slot at C3);
    // large ones keep the per-point launch (C4: on-the-fly factors measured 99 against 83 + 9 us)
    const bool fuse_w = B.N <= 200000;
    if (!fuse_w)
      launch_step_kernel(ba_point_invert, (B.Pf + 127) / 128, 128, 0, st, lm, B.Pf, B.Hll.p, B.Hinv.p, B.Wfac.p);
    launch_step_kernel(ba_compute_G, grid, kCamThreads, 0, st, V, B.Hpl.p, B.Wfac.p, B.bl.p, B.Gbuf.p, B.camp.p,
                       fuse_w ? (const double*)B.Hll.p : (const double*)nullptr);
    if (B.n_units)
      launch_step_kernel(ba_schur_units, (B.n_units + kSchurWarps - 1) / kSchurWarps, kSchurWarps * 32, kSchurSmem, st,
                         lm, B.Gbuf.p, B.item_a.p, B.item_b.p, B.unit_start.p, B.unit_len.p, B.n_units,
                         B.unit_partial.p);
  }
  if (B.n) {
    BhatArgs Bh{(B.Pf && !chunked) ? B.camp.p : nullptr, B.bp.p, B.bp_extra.p, B.bhat.p, B.bar.p, B.pcg_flags.p,
                (B.Pf && chunked) ? B.ck_bh.p : nullptr, B.cam_ck_start.p};
    const int* seg_us = chunked ? B.seg_cunit_start.p : B.seg_unit_start.p;
    const double* partials = chunked ? B.ck_partial.p : B.unit_partial.p;
    // four threads per entry when a block has many unit partials to add (a local window with 64-item units)
    const long long n_part = chunked ? B.n_cunits : B.n_units;
    if (n_part > 8ll * B.n_seg)
      launch_step_kernel(ba_schur_finalize<4>, ((B.n_seg * 36 + B.C * 6) * 4 + 127) / 128, 128, 0, st,
                         V, B.n_seg, B.n, B.seg_a.p, B.seg_b.p, seg_us, partials, B.Hpp.p,
                         B.extra.p, B.seg_extra_idx.p, B.S.p, Bh);
    else
      launch_step_kernel(ba_schur_finalize<1>, (B.n_seg * 36 + B.C * 6 + 127) / 128, 128, 0, st,
                         V, B.n_seg, B.n, B.seg_a.p, B.seg_b.p, seg_us, partials, B.Hpp.p,
                         B.extra.p, B.seg_extra_idx.p, B.S.p, Bh);
    if (B.use_direct()) {
      DirectArgs D{};
      D.n = B.n; D.ld = direct_ld(B.n); D.done = &B.lm.p->done; D.S = B.S.p; D.b = B.bhat.p;
      D.x = B.dp.p; D.status = B.pcg_flags.p + 1; D.iters = B.pcg_flags.p + 2;
      if (B.use_gj()) launch_step_kernel(ba_solve_gj_kernel, 1, kGJThreads, 0, st, D);
      else launch_step_kernel(ba_solve_direct_kernel, 1, kDirectThreads, direct_smem(B.n), st, D);
    }
  }
  return check_launch("ba_schur");
}

// One LM trial at LMState::lambda: solve (H + lambda I) d = b through Schur + PCG, write the trial
// state, evaluate it, then the controller decides (accept / reject / terminate) on the device.
static int enqueue_trial(BA& B, cudaStream_t st, bool speculative = false) {
  const BAView V = B.view();
  dim3 grid(B.max_chunks, B.C);
  {
    const int rc = enqueue_schur(B, st);
    if (rc) return rc;
  }
  if (B.n && !B.use_direct()) {
    PcgArgs A{};   // (barrier counter and flags were reset by ba_schur_finalize)
    A.done = &B.lm.p->done;
    A.S = B.S.p; A.b = B.bhat.p; A.x = B.dp.p; A.n = B.n; A.rows = B.pcg_rows; A.G = B.pcg_G;
    A.slab_in_smem = B.pcg_slab; A.max_iter = B.pcg_max_iter; A.tol = B.pcg_tol;
    A.p_glob = B.p_glob.p; A.w_loc = B.w_loc.p; A.part = B.part.p; A.bar = B.bar.p; A.fail_flag = B.pcg_flags.p;
    A.status = B.pcg_flags.p + 1; A.iters = B.pcg_flags.p + 2;
    static const bool pipelined = [] { const char* e = getenv("VSLAM_BA_PCG_PIPELINED"); return !(e && atoi(e) == 0); }();
    A.pipelined = (pipelined && B.pcg_G > 1) ? 1 : 0;
    void* args[] = {&A};
    if (B.pcg_cluster) {
      BA_TRY(launch_pcg_cluster(B.pcg_cluster, B.pcg_G, B.pcg_smem, A, st));
    } else if (B.pcg_G > 1) {
      BA_TRY(cudaLaunchCooperativeKernel((void*)ba_pcg_kernel, dim3(B.pcg_G), dim3(kPcgThreads), args,
                                         B.pcg_smem, st));
    } else if (B.n <= kPcgSmallMax) {
      launch_step_kernel(pcg_small_fn(), 1, kPcgSmallThreads, pcg_small_smem(B.n), st, A);
    } else if (B.n <= kPcgMidMax) {
      launch_step_kernel(ba_pcg_mid_kernel, 1, kPcgMidThreads, kPcgMidSmemDoubles * sizeof(double), st, A);
    } else {
      launch_step_kernel(ba_pcg_kernel, 1, kPcgThreads, B.pcg_smem, st, A);
    }
  }
  if (B.Pf) launch_step_kernel(ba_backsub_cam, grid, kCamThreads, 0, st, V, B.Hpl.p, B.dp.p, B.tmp3.p);
  {
    const int n_pt_blocks = B.P ? B.n_sp : 0;
    launch_step_kernel(ba_backsub_pt, n_pt_blocks + (B.C + 255) / 256, 256, 0, st, V, B.tmp3.p, B.Hll.p, B.bl.p,
                       B.dl.p, B.scale_partial.p, B.dp.p, n_pt_blocks, B.bp.p, B.bp_extra.p);
  }
  // trial cost; its last CTA is the LM controller (accept / reject / terminate).  Inside the LM loop the cost pass is
  // a SPECULATIVE full linearisation of the trial state (see spec_controller): an accepted trial leaves the next
  // iteration's system in place, so an iteration costs one evaluation of the observations' residuals instead of two
  // (C3: the 10 us cost pass disappears from a 62 us LM slot; C4: 26 of 555 us)
  if (speculative) {
    LinArgs LA = B.lin_args();
    LA.spec = 1;
    launch_step_kernel(ba_linearize_kernel<false>, B.lin_ctas, kLinThreads, B.lin_smem, st, V, LA, B.scale_view());
  } else {
    launch_step_kernel(ba_linearize_kernel<true>, B.lin_nA, kLinThreads, kLinCostSmem, st, V, B.lin_args(), B.scale_view());
  }
  return check_launch("ba_trial");
}

static int enqueue_step(BA& B, cudaStream_t st) {
  static const bool spec = [] { const char* e = getenv("VSLAM_BA_SPECULATIVE"); return !(e && atoi(e) == 0); }();
  int rc = enqueue_linearize(B, st);
  if (rc) return rc;
  return enqueue_trial(B, st, spec);
}

// Captures one LM trial slot into a CUDA graph (all arguments are fixed device pointers; lambda and
// the buffer roles live in LMState).  On any capture problem the step is launched kernel by kernel.
static void try_capture_step(BA& B) {
  // a structure that is optimised once (the reference builds a fresh optimiser per call, main.py:213,322) is
  // launched kernel by kernel (programmatic dependent launches); capture + instantiation (~1 ms) pays from the
  // second optimize() of the same structure on
  static const bool no_graph = [] { const char* e = getenv("VSLAM_BA_NO_GRAPH"); return e && atoi(e) != 0; }();
  if (no_graph || B.graph_tried || B.optimizes_since_build < 1) return;
  B.graph_tried = true;
  cudaStream_t cs = nullptr;
  if (cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); return; }
  cudaGraph_t graph = nullptr;
  bool ok = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
  if (ok) {
    const int rc = enqueue_step(B, cs);
    const cudaError_t e = cudaStreamEndCapture(cs, &graph);
    ok = (rc == VSLAM_OK) && (e == cudaSuccess) && graph != nullptr;
  }
  if (ok) ok = cudaGraphInstantiate(&B.step_graph, graph, 0) == cudaSuccess;
  if (graph) cudaGraphDestroy(graph);
  if (!ok) { B.step_graph = nullptr; cudaGetLastError(); }
  cudaStreamDestroy(cs);
}

// `count` trial slots as ONE graph launch (captured once per structure and count); false if capture is not possible.
static bool launch_steps_as_one_graph(BA& B, int count, cudaStream_t st) {
  if (B.optimizes_since_build < 1 || count < 2) return false;
  if (B.multi_graph && B.multi_slots != count) { cudaGraphExecDestroy(B.multi_graph); B.multi_graph = nullptr; }
  if (!B.multi_graph) {
    cudaStream_t cs = nullptr;
    if (cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); return false; }
    cudaGraph_t graph = nullptr;
    bool ok = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
    if (ok) {
      int rc = VSLAM_OK;
      for (int i = 0; i < count && rc == VSLAM_OK; ++i) rc = enqueue_step(B, cs);
      const cudaError_t e = cudaStreamEndCapture(cs, &graph);
      ok = (rc == VSLAM_OK) && (e == cudaSuccess) && graph != nullptr;
    }
    if (ok) ok = cudaGraphInstantiate(&B.multi_graph, graph, 0) == cudaSuccess;
    if (graph) cudaGraphDestroy(graph);
    cudaStreamDestroy(cs);
    if (!ok) { B.multi_graph = nullptr; cudaGetLastError(); return false; }
    B.multi_slots = count;
  }
  return cudaGraphLaunch(B.multi_graph, st) == cudaSuccess;
}

static int launch_steps(BA& B, int count, cudaStream_t st) {
  try_capture_step(B);
  for (int i = 0; i < count; ++i) {
    if (B.step_graph) BA_TRY(cudaGraphLaunch(B.step_graph, st));
    else { int rc = enqueue_step(B, st); if (rc) return rc; }
  }
  return VSLAM_OK;
}

}  // namespace vslam

using namespace vslam;

struct vslam_ba { BA impl; };

// OptimizationAlgorithmLevenberg::solve driven by SparseOptimizer::optimize(max_iterations)
// (reference LocalBA.py:39-42).  The whole loop runs on the device: `max_iterations` trial slots are
// enqueued back to back (one CUDA graph launch each), the state is read once, and more slots follow
// only if a trial was rejected.  stats: [0] initial cost, [1] final cost, [2] outer iterations,
// [3] trials, [4] final lambda, [5] total PCG iterations, [6] rejected trials, [7] PCG breakdowns.
static int ba_optimize(BA& B, int max_iterations, double* stats, cudaStream_t st) {
  if (B.C == 0) return VSLAM_ERR_BAD_ARG;
  struct Count { BA& b; ~Count() { b.optimizes_since_build++; } } count{B};
  ba_lm_reset<<<1, 1, 0, st>>>(B.lm.p, max_iterations, max_iterations == 0 ? 1 : 0);
  if (max_iterations == 0) {   // report the cost of the state as given
    int rc = enqueue_linearize(B, st);
    if (rc) return rc;
    rc = ba_sync_state(B, st);
    if (rc) return rc;
    if (stats) {
      for (int i = 0; i < 8; ++i) stats[i] = 0.0;
      stats[0] = stats[1] = B.lm_host->cur_cost;
    }
    return VSLAM_OK;
  }
  int budget = max_iterations * 10 + 1;   // at most 10 trials per outer iteration
  int batch = max_iterations;
  static const bool one_graph = [] { const char* e = getenv("VSLAM_BA_ONE_GRAPH"); return !(e && atoi(e) == 0); }();
  while (true) {
    int rc = VSLAM_OK;
    // the first batch (max_iterations slots) goes out as ONE graph launch; follow-up slots after a rejected trial
    // one by one (their count varies, and a graph per count would be re-captured every time)
    const bool first_batch = batch == max_iterations && budget == max_iterations * 10 + 1;
    if (!(one_graph && first_batch && launch_steps_as_one_graph(B, batch, st))) rc = launch_steps(B, batch, st);
    if (rc) return rc;
    rc = ba_sync_state(B, st);
    if (rc) return rc;
    budget -= batch;
    if (B.lm_host->done || budget <= 0) break;
    batch = std::max(1, std::min(budget, max_iterations - B.lm_host->iters));
  }
  const LMState& L = *B.lm_host;
  if (stats) {
    stats[0] = L.initial_cost; stats[1] = L.cur_cost; stats[2] = L.iters; stats[3] = L.trials;
    stats[4] = L.lambda; stats[5] = L.pcg_total; stats[6] = L.rejected; stats[7] = L.breakdowns;
  }
  return VSLAM_OK;
}

// Independent BA problems (one local window per sequence) optimised concurrently: every problem's LM
// slots are enqueued on one of a small pool of streams without waiting, the controller states are read once all
// are in flight, and only problems that are not done yet (a rejected trial) get further slots.  Small windows are
// latency-bound on their own (C3: ~70 us per LM iteration at a few per cent of the machine); side by side they
// fill it.
static int ba_optimize_batch(BA** Bs, int count, int max_iterations, double* stats) {
  static cudaStream_t pool[32];
  static int pool_n = 0;
  const int want = std::min(count, 32);
  while (pool_n < want) {
    BA_TRY(cudaStreamCreateWithFlags(&pool[pool_n], cudaStreamNonBlocking));
    ++pool_n;
  }
  std::vector<int> budget(count, max_iterations * 10 + 1), batch(count, max_iterations);
  std::vector<char> done(count, 0);
  BA_TRY(cudaDeviceSynchronize());          // the callers' set_state calls ran on other streams
  for (int i = 0; i < count; ++i) {
    if (Bs[i]->C == 0 || max_iterations <= 0) return VSLAM_ERR_BAD_ARG;
    ba_lm_reset<<<1, 1, 0, pool[i % pool_n]>>>(Bs[i]->lm.p, max_iterations, 0);
  }
  int open = count;
  while (open > 0) {
    for (int i = 0; i < count; ++i) {
      if (done[i]) continue;
      if (launch_steps_as_one_graph(*Bs[i], batch[i], pool[i % pool_n])) continue;
      const int rc = launch_steps(*Bs[i], batch[i], pool[i % pool_n]);
      if (rc) return rc;
    }
    for (int i = 0; i < count; ++i)
      if (!done[i]) BA_TRY(cudaMemcpyAsync(Bs[i]->lm_host, Bs[i]->lm.p, sizeof(LMState), cudaMemcpyDeviceToHost, pool[i % pool_n]));
    for (int s = 0; s < pool_n; ++s) BA_TRY(cudaStreamSynchronize(pool[s]));
    for (int i = 0; i < count; ++i) {
      if (done[i]) continue;
      budget[i] -= batch[i];
      if (Bs[i]->lm_host->done || budget[i] <= 0) { done[i] = 1; --open; Bs[i]->optimizes_since_build++; }
      else batch[i] = std::max(1, std::min(budget[i], max_iterations - Bs[i]->lm_host->iters));
    }
  }
  if (stats) {
    for (int i = 0; i < count; ++i) {
      const LMState& L = *Bs[i]->lm_host;
      double* s = stats + 8 * (size_t)i;
      s[0] = L.initial_cost; s[1] = L.cur_cost; s[2] = L.iters; s[3] = L.trials;
      s[4] = L.lambda; s[5] = L.pcg_total; s[6] = L.rejected; s[7] = L.breakdowns;
    }
  }
  return VSLAM_OK;
}

extern "C" int vslam_ba_create(vslam_ba** out, int n_cams, int n_pts, int n_obs, int n_scale,
                               const int32_t* obs_cam, const int32_t* obs_pt, const uint8_t* cam_fixed,
                               const uint8_t* pt_fixed, const int32_t* scale_a, const int32_t* scale_b) {
  if (!out || n_cams <= 0 || n_pts < 0 || n_obs < 0 || n_scale < 0) return VSLAM_ERR_BAD_ARG;
  if ((n_obs && (!obs_cam || !obs_pt)) || (n_scale && (!scale_a || !scale_b))) return VSLAM_ERR_BAD_ARG;
  if (sm_count() <= 0) return VSLAM_ERR_NO_DEVICE;
  vslam_ba* h = new vslam_ba();
  int rc = ba_build(h->impl, n_cams, n_pts, n_obs, n_scale, obs_cam, obs_pt, cam_fixed, pt_fixed,
                    scale_a, scale_b);
  if (rc) { delete h; return rc; }
  *out = h;
  return VSLAM_OK;
}

extern "C" int vslam_ba_set_structure(vslam_ba* h, int n_cams, int n_pts, int n_obs, int n_scale,
                                      const int32_t* obs_cam, const int32_t* obs_pt, const uint8_t* cam_fixed,
                                      const uint8_t* pt_fixed, const int32_t* scale_a, const int32_t* scale_b) {
  if (!h || n_cams <= 0 || n_pts < 0 || n_obs < 0 || n_scale < 0) return VSLAM_ERR_BAD_ARG;
  if ((n_obs && (!obs_cam || !obs_pt)) || (n_scale && (!scale_a || !scale_b))) return VSLAM_ERR_BAD_ARG;
  cudaError_t e = cudaDeviceSynchronize();   // nothing of the previous structure may still be in flight
  if (e != cudaSuccess) { set_cuda_error(e, "vslam_ba_set_structure"); return VSLAM_ERR_CUDA; }
  return ba_build(h->impl, n_cams, n_pts, n_obs, n_scale, obs_cam, obs_pt, cam_fixed, pt_fixed, scale_a, scale_b);
}

extern "C" int vslam_ba_destroy(vslam_ba* h) {
  delete h;
  return VSLAM_OK;
}

extern "C" int vslam_ba_set_state(vslam_ba* h, const double* cams, const double* pts, const double* obs_uv,
                                  const double* scale_meas, const double* K, void* stream) {
  if (!h || !cams || !K || (h->impl.P && !pts) || (h->impl.N && !obs_uv) || (h->impl.E && !scale_meas))
    return VSLAM_ERR_BAD_ARG;
  return ba_set_state(h->impl, cams, pts, obs_uv, scale_meas, K, static_cast<cudaStream_t>(stream));
}

extern "C" int vslam_ba_get_state(vslam_ba* h, double* cams, double* pts, void* stream) {
  if (!h) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (cams) ba_copy_state<<<64, 256, 0, st>>>(B.lm.p, 0, B.cams[0].p, B.cams[1].p, cams, (size_t)B.C * 7);
  if (pts && B.P) ba_unpad_points<<<(B.P + 255) / 256, 256, 0, st>>>(B.lm.p, 0, B.pts[0].p, B.pts[1].p, pts, B.P);
  return check_launch("ba_copy_state");
}

extern "C" int vslam_ba_set_solver(vslam_ba* h, double pcg_tol, int pcg_max_iter) {
  if (!h || !(pcg_tol > 0)) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  const int mi = pcg_max_iter > 0 ? pcg_max_iter : B.pcg_max_iter;
  if (pcg_tol != B.pcg_tol || mi != B.pcg_max_iter) {   // both are baked into the captured step
    if (B.step_graph) { cudaGraphExecDestroy(B.step_graph); B.step_graph = nullptr; }
    if (B.multi_graph) { cudaGraphExecDestroy(B.multi_graph); B.multi_graph = nullptr; B.multi_slots = 0; }
    B.graph_tried = false;
  }
  B.pcg_tol = pcg_tol;
  B.pcg_max_iter = mi;
  return VSLAM_OK;
}

extern "C" int vslam_ba_set_schur_variant(vslam_ba* h, int variant) {
  if (!h || variant < 0 || variant > 1) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  if (variant != B.schur_variant) {
    if (B.step_graph) { cudaGraphExecDestroy(B.step_graph); B.step_graph = nullptr; }
    B.graph_tried = false;
  }
  B.schur_variant = variant;
  return VSLAM_OK;
}

extern "C" int vslam_ba_set_linear_solver(vslam_ba* h, int kind) {
  if (!h || kind < 0 || kind > 3) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  if ((kind == 2 || kind == 3) && B.n > kDirectMax) return VSLAM_ERR_BAD_ARG;
  if (kind != B.linear_solver) {   // the choice is baked into the captured step
    if (B.step_graph) { cudaGraphExecDestroy(B.step_graph); B.step_graph = nullptr; }
    if (B.multi_graph) { cudaGraphExecDestroy(B.multi_graph); B.multi_graph = nullptr; B.multi_slots = 0; }
    B.graph_tried = false;
  }
  B.linear_solver = kind;
  return VSLAM_OK;
}

extern "C" int vslam_ba_linearize(vslam_ba* h, double* out2, void* stream) {
  if (!h) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  if (B.C == 0) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ba_lm_poke<<<1, 1, 0, st>>>(B.lm.p, 0, 0.0);
  int rc = enqueue_linearize(B, st);
  if (rc) return rc;
  rc = ba_sync_state(B, st);
  if (rc) return rc;
  if (out2) { out2[0] = B.lm_host->cur_cost; out2[1] = B.lm_host->max_diag; }
  return VSLAM_OK;
}

// Times one stage with CUDA events on `stream` (the stream the kernels run on): stage 0 = the fused
// linearisation kernel alone (buildSystem), 1 = Schur complement assembly (point inverse + G + units +
// finalize), 2 = one whole LM trial, 3 = the fused linearisation kernel launched `reps` times back to
// back between ONE pair of events (ms_out[0] = average per launch; its working set at C4, 194 MB, is
// larger than the L2, and nothing but the kernel runs inside the timed region).  For stages 0-2 every
// launch is bracketed on its own (ms_out[reps]) after flush_bytes of flush_buf are overwritten (L2
// flush; skipped when null).  Synchronises.
extern "C" int vslam_ba_time_stage(vslam_ba* h, int stage, int reps, void* flush_buf, int64_t flush_bytes,
                                   float* ms_out, void* stream) {
  if (!h || reps < 1 || !ms_out || stage < 0 || stage > 3) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  if (B.C == 0) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaEvent_t e0, e1;
  BA_TRY(cudaEventCreate(&e0)); BA_TRY(cudaEventCreate(&e1));
  int rc = VSLAM_OK;
  if (stage == 3) {
    ba_lm_poke<<<1, 1, 0, st>>>(B.lm.p, 0, 0.0);
    LinArgs A = B.lin_args();
    A.force = 1;
    const BAView V = B.view();
    const ScaleView SV = B.scale_view();
    ba_linearize_kernel<false><<<B.lin_ctas, kLinThreads, B.lin_smem, st>>>(V, A, SV);   // warm-up
    cudaEventRecord(e0, st);
    for (int r = 0; r < reps; ++r) ba_linearize_kernel<false><<<B.lin_ctas, kLinThreads, B.lin_smem, st>>>(V, A, SV);
    cudaEventRecord(e1, st);
    rc = check_launch("ba_linearize");
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_cuda_error(e, "vslam_ba_time_stage"); rc = VSLAM_ERR_CUDA; }
    else { cudaEventElapsedTime(&ms_out[0], e0, e1); ms_out[0] /= (float)reps; }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return rc;
  }
  for (int r = 0; r < reps && rc == VSLAM_OK; ++r) {
    ba_lm_poke<<<1, 1, 0, st>>>(B.lm.p, 0, 0.0);
    if (stage != 0) rc = enqueue_linearize(B, st);
    if (flush_buf && flush_bytes > 0) cudaMemsetAsync(flush_buf, r & 0xff, (size_t)flush_bytes, st);
    cudaEventRecord(e0, st);
    if (rc == VSLAM_OK) {
      if (stage == 0) rc = enqueue_linearize(B, st);
      else if (stage == 1) rc = enqueue_schur(B, st);
      else rc = enqueue_trial(B, st);
    }
    cudaEventRecord(e1, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_cuda_error(e, "vslam_ba_time_stage"); rc = VSLAM_ERR_CUDA; break; }
    cudaEventElapsedTime(&ms_out[r], e0, e1);
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return rc;
}

// Profiling: runs one linearisation with per-CTA globaltimer stamps.  out (host) receives, per CTA of
// the fused grid, {kind, start_ns, end_ns} relative to the earliest start; returns the CTA count in *n.
extern "C" int vslam_ba_trace_linearize(vslam_ba* h, int64_t* out, int64_t capacity, int32_t* n, void* stream) {
  if (!h || !out || !n) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  *n = B.lin_ctas;
  if (capacity < (int64_t)3 * B.lin_ctas) return VSLAM_ERR_BAD_ARG;
  if (!B.trace.p) BA_TRY(B.trace.alloc((size_t)2 * B.lin_ctas + 8));
  BA_TRY(cudaMemsetAsync(B.trace.p, 0, ((size_t)2 * B.lin_ctas + 8) * sizeof(long long), st));
  ba_lm_poke<<<1, 1, 0, st>>>(B.lm.p, 0, 0.0);
  ba_stamp<<<1, 1, 0, st>>>(B.trace.p + 2 * B.lin_ctas + 1);
  int rc = enqueue_linearize(B, st);
  if (rc) return rc;
  ba_stamp<<<1, 1, 0, st>>>(B.trace.p + 2 * B.lin_ctas + 2);
  std::vector<long long> t((size_t)2 * B.lin_ctas + 8);
  std::vector<LinWork> w(B.lin_ctas);
  BA_TRY(cudaMemcpyAsync(t.data(), B.trace.p, t.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
  BA_TRY(cudaMemcpyAsync(w.data(), B.lin_work.p, w.size() * sizeof(LinWork), cudaMemcpyDeviceToHost, st));
  BA_TRY(cudaStreamSynchronize(st));
  cudaFree(B.trace.p); B.trace.p = nullptr; B.trace.cap = 0;
  long long t0 = std::numeric_limits<long long>::max();
  for (int i = 0; i < B.lin_ctas; ++i) if (t[2 * i]) t0 = std::min(t0, t[2 * i]);
  for (int i = 0; i < B.lin_ctas; ++i) {
    out[3 * i] = w[i].kind; out[3 * i + 1] = t[2 * i] - t0; out[3 * i + 2] = t[2 * i + 1] - t0;
  }
  if (capacity >= (int64_t)3 * B.lin_ctas + 7) {
    for (int k = 3; k < 7; ++k) out[3 * B.lin_ctas + k] = t[2 * B.lin_ctas + k] - t0;   // stamp kernel before, controller end, stamp kernel after
    out[3 * B.lin_ctas] = t[2 * B.lin_ctas + 1] - t0; out[3 * B.lin_ctas + 1] = t[2 * B.lin_ctas] - t0;
    out[3 * B.lin_ctas + 2] = t[2 * B.lin_ctas + 2] - t0;
  }
  return VSLAM_OK;
}

extern "C" int vslam_ba_trial(vslam_ba* h, double lambda, double* out4, void* stream) {
  if (!h) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ba_lm_poke<<<1, 1, 0, st>>>(B.lm.p, 1, lambda);
  int rc = enqueue_linearize(B, st);   // no-op unless the state changed since the last rebuild
  if (rc) return rc;
  rc = enqueue_trial(B, st);
  if (rc) return rc;
  rc = ba_sync_state(B, st);
  if (rc) return rc;
  if (out4) {
    out4[0] = B.lm_host->last_cost; out4[1] = B.lm_host->last_scale;
    out4[2] = (double)B.lm_host->last_pcg_status; out4[3] = (double)B.lm_host->last_pcg_iters;
  }
  return VSLAM_OK;
}

extern "C" int vslam_ba_accept(vslam_ba* h) {
  if (!h) return VSLAM_ERR_BAD_ARG;
  ba_lm_poke<<<1, 1, 0, 0>>>(h->impl.lm.p, 2, 0.0);   // legacy default stream: ordered after prior work
  return check_launch("ba_lm_poke");
}

extern "C" int vslam_ba_optimize(vslam_ba* h, int max_iterations, double* stats8, void* stream) {
  if (!h || max_iterations < 0) return VSLAM_ERR_BAD_ARG;
  return ba_optimize(h->impl, max_iterations, stats8, static_cast<cudaStream_t>(stream));
}

extern "C" int vslam_ba_optimize_batch(vslam_ba** hs, int count, int max_iterations, double* stats8) {
  if (!hs || count < 0 || max_iterations <= 0) return VSLAM_ERR_BAD_ARG;
  std::vector<BA*> Bs(count);
  for (int i = 0; i < count; ++i) {
    if (!hs[i]) return VSLAM_ERR_BAD_ARG;
    Bs[i] = &hs[i]->impl;
  }
  return count ? ba_optimize_batch(Bs.data(), count, max_iterations, stats8) : VSLAM_OK;
}

extern "C" int vslam_ba_dims(vslam_ba* h, int32_t* out8) {
  if (!h || !out8) return VSLAM_ERR_BAD_ARG;
  const BA& B = h->impl;
  out8[0] = B.Cf; out8[1] = B.Pf; out8[2] = B.n; out8[3] = B.n_items; out8[4] = B.n_units;
  out8[5] = B.n_seg; out8[6] = B.pcg_G; out8[7] = B.pcg_rows;
  return VSLAM_OK;
}

extern "C" int vslam_ba_read(vslam_ba* h, int what, double* dst, int64_t capacity, void* stream) {
  if (!h || !dst) return VSLAM_ERR_BAD_ARG;
  BA& B = h->impl;
  const double* src = nullptr;
  size_t n = 0;
  switch (what) {
    case 0: src = B.Hpp.p; n = (size_t)B.Cf * 36; break;
    case 1:   // b = projection part + scale-edge part
      n = B.n;
      if ((int64_t)n > capacity) return VSLAM_ERR_BAD_ARG;
      if (n) ba_add2<<<(B.n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(B.bp.p, B.bp_extra.p, dst, B.n);
      return check_launch("ba_add2");
    case 2: src = B.Hll.p; n = (size_t)B.Pf * 6; break;
    case 3: src = B.bl.p; n = (size_t)B.Pf * 3; break;
    case 4: src = B.S.p; n = (size_t)B.n * B.n; break;
    case 5: src = B.bhat.p; n = B.n; break;
    case 6: src = B.dp.p; n = B.n; break;
    case 7: src = B.dl.p; n = (size_t)B.Pf * 3; break;
    case 8:
      n = (size_t)B.C * 7;
      if ((int64_t)n > capacity) return VSLAM_ERR_BAD_ARG;
      ba_copy_state<<<64, 256, 0, static_cast<cudaStream_t>(stream)>>>(B.lm.p, 1, B.cams[0].p, B.cams[1].p, dst, n);
      return check_launch("ba_copy_state");
    case 9:
      n = (size_t)B.P * 3;
      if ((int64_t)n > capacity) return VSLAM_ERR_BAD_ARG;
      ba_unpad_points<<<(B.P + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(B.lm.p, 1, B.pts[0].p, B.pts[1].p, dst, B.P);
      return check_launch("ba_unpad_points");
    default: return VSLAM_ERR_BAD_ARG;
  }
  if ((int64_t)n > capacity) return VSLAM_ERR_BAD_ARG;
  if (n) BA_TRY(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  return VSLAM_OK;
}

// Stand-alone small-system solve of a caller-supplied dense symmetric system (tests: fixture F1): the register-
// resident Gauss-Jordan kernel for n <= 63, the shared-memory LDL^T kernel above that (negative n forces LDL^T).
extern "C" int vslam_direct_solve(const double* S, const double* b, double* x, int n_signed, int32_t* status_host,
                                  void* stream) {
  const int n = n_signed < 0 ? -n_signed : n_signed;
  const bool force_ldl = n_signed < 0;
  if (!S || !b || !x || n <= 0 || n % 6 || n > kDirectMax || !status_host) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DevBuf<int> flags;
  BA_TRY(flags.alloc(4));
  BA_TRY(cudaMemsetAsync(flags.p, 0, 4 * sizeof(int), st));
  DirectArgs D{};
  D.n = n; D.ld = direct_ld(n); D.x = x; D.status = flags.p + 1; D.iters = flags.p + 2; D.S = S; D.b = b;
  int rc;
  if (n <= kGJMax && !force_ldl) {
    ba_solve_gj_kernel<<<1, kGJThreads, 0, st>>>(D);
    rc = check_launch("ba_solve_gj_kernel");
  } else {
    BA_TRY(cudaFuncSetAttribute(ba_solve_direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)direct_smem(kDirectMax)));
    ba_solve_direct_kernel<<<1, kDirectThreads, direct_smem(n), st>>>(D);
    rc = check_launch("ba_solve_direct_kernel");
  }
  if (rc) return rc;
  int host[4];
  BA_TRY(cudaMemcpyAsync(host, flags.p, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
  BA_TRY(cudaStreamSynchronize(st));
  *status_host = host[1];
  return VSLAM_OK;
}

// Stand-alone PCG on a caller-supplied dense SPD system (tests: fixture F1 = reference debug.txt).
extern "C" int vslam_pcg_solve(const double* S, const double* b, double* x, int n, double tol, int max_iter,
                               int32_t* status_iters_host, void* stream) {
  if (!S || !b || !x || n <= 0 || n % 6 || !status_iters_host) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DevBuf<double> p_glob, w_loc, part;
  DevBuf<unsigned> bar;
  DevBuf<int> flags;
  BA_TRY(p_glob.alloc(n)); BA_TRY(w_loc.alloc(n)); BA_TRY(part.alloc(4)); BA_TRY(bar.alloc(1)); BA_TRY(flags.alloc(4));
  BA_TRY(cudaMemsetAsync(bar.p, 0, sizeof(unsigned), st));
  BA_TRY(cudaMemsetAsync(flags.p, 0, 4 * sizeof(int), st));
  PcgArgs A{};
  A.S = S; A.b = b; A.x = x; A.n = n; A.rows = n; A.G = 1; A.max_iter = max_iter; A.tol = tol;
  A.p_glob = p_glob.p; A.w_loc = w_loc.p; A.part = part.p; A.bar = bar.p; A.fail_flag = flags.p; A.status = flags.p + 1;
  A.iters = flags.p + 2;
  size_t smem_slab = sizeof(double) * ((size_t)n + 12 * (size_t)n + 8 + (size_t)n * n);
  size_t smem_noslab = sizeof(double) * ((size_t)n + 12 * (size_t)n + 8);
  A.slab_in_smem = smem_slab <= 200 * 1024;
  size_t smem = A.slab_in_smem ? smem_slab : smem_noslab;
  BA_TRY(pcg_kernel_allow_max_smem());
  if (n <= kPcgSmallMax) pcg_small_fn()<<<1, kPcgSmallThreads, pcg_small_smem(n), st>>>(A);
  else if (n <= kPcgMidMax && n % 6 == 0) ba_pcg_mid_kernel<<<1, kPcgMidThreads, kPcgMidSmemDoubles * sizeof(double), st>>>(A);
  else {
    // (n is a multiple of 6, checked above) the thread-block cluster kernel where it applies, as inside the BA engine
    const PcgClusterPlan P = pcg_cluster_plan(n / 6, pcg_smem_budget());
    if (P.kind) {
      A.rows = 6 * P.blocks; A.G = P.ctas;
      BA_TRY(launch_pcg_cluster(P.kind, P.ctas, P.smem, A, st));
    } else {
      ba_pcg_kernel<<<1, kPcgThreads, smem, st>>>(A);
    }
  }
  int rc = check_launch("ba_pcg_kernel");
  if (rc) return rc;
  int host[4];
  BA_TRY(cudaMemcpyAsync(host, flags.p, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
  BA_TRY(cudaStreamSynchronize(st));
  status_iters_host[0] = host[1];
  status_iters_host[1] = host[2];
  return VSLAM_OK;
}

// ---- device-resident tracking window (ba_window.cuh) ---------------------------------------------------------
struct vslam_window { Window impl; };

extern "C" int vslam_window_create(vslam_window** out, int max_frames, int slot, int max_points) {
  if (!out || max_frames < 2 || slot < 1 || max_points < 1) return VSLAM_ERR_BAD_ARG;
  int sms = sm_count();
  if (sms <= 0) return VSLAM_ERR_NO_DEVICE;
  if (max_frames > sms) return VSLAM_ERR_BAD_ARG;        // one co-resident CTA per frame
  vslam_window* h = new vslam_window();
  Window& W = h->impl;
  W.max_frames = max_frames; W.slot = slot; W.max_pts = max_points;
  const size_t F = max_frames, S = slot;
  cudaError_t e = cudaSuccess;
  auto take = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
  take((void**)&W.pts4, (size_t)max_points * 4 * sizeof(double));
  take((void**)&W.cams, F * 7 * sizeof(double));
  take((void**)&W.fixed, F);
  take((void**)&W.obs_pt, F * S * sizeof(int));
  take((void**)&W.obs_uv, F * S * sizeof(double2));
  take((void**)&W.n_obs, F * sizeof(int));
  take((void**)&W.part, 3 * F * kWinPart * sizeof(double));
  take((void**)&W.bar, sizeof(unsigned));
  take((void**)&W.stats, 8 * sizeof(double));
  if (e == cudaSuccess) e = cudaMallocHost(&W.stats_host, 8 * sizeof(double));
  if (e == cudaSuccess) e = cudaMemset(W.n_obs, 0, F * sizeof(int));
  if (e != cudaSuccess) { set_cuda_error(e, "vslam_window_create"); delete h; return VSLAM_ERR_CUDA; }
  *out = h;
  return VSLAM_OK;
}

extern "C" int vslam_window_destroy(vslam_window* h) {
  delete h;
  return VSLAM_OK;
}

extern "C" int vslam_window_reset(vslam_window* h, const double* pts, const float* kf_uv, int n_pts,
                                  const double* kf_cam7, const double* K, void* stream) {
  if (!h || !pts || !kf_uv || !kf_cam7 || !K || n_pts < 1) return VSLAM_ERR_BAD_ARG;
  Window& W = h->impl;
  if (n_pts > W.max_pts) return VSLAM_ERR_BAD_ARG;
  W.K = make_double4(K[0], K[1], K[2], K[3]);
  W.n_pts = n_pts; W.n_frames = 1;
  window_reset_kernel<<<(std::max(n_pts, 7) + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      pts, kf_uv, kf_cam7, n_pts, W.slot, W.pts4, W.cams, W.fixed, W.obs_pt, W.obs_uv, W.n_obs);
  return check_launch("window_reset_kernel");
}

extern "C" int vslam_window_append(vslam_window* h, const double* cam7, const int32_t* keep_q, const int32_t* keep_t,
                                   const int32_t* n_keep, const float* kp_cur, void* stream) {
  if (!h || !cam7 || !keep_q || !keep_t || !n_keep || !kp_cur) return VSLAM_ERR_BAD_ARG;
  Window& W = h->impl;
  if (W.n_frames < 1 || W.n_frames >= W.max_frames) return VSLAM_ERR_BAD_ARG;   // reset first; window full
  const int f = W.n_frames++;
  window_append_kernel<<<(std::max(W.slot, 7) + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      f, W.slot, cam7, keep_q, keep_t, n_keep, kp_cur, W.n_pts, W.cams, W.fixed, W.obs_pt, W.obs_uv, W.n_obs);
  return check_launch("window_append_kernel");
}

extern "C" int vslam_window_motion_ba(vslam_window* h, int max_iterations, double* stats8, void* stream) {
  if (!h || max_iterations < 0) return VSLAM_ERR_BAD_ARG;
  Window& W = h->impl;
  if (W.n_frames < 1) return VSLAM_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  WinView V;
  V.n_frames = W.n_frames; V.slot = W.slot; V.max_iterations = max_iterations; V.K = W.K;
  V.pts4 = W.pts4; V.cams = W.cams; V.fixed = W.fixed; V.obs_pt = W.obs_pt; V.obs_uv = W.obs_uv; V.n_obs = W.n_obs;
  V.part = W.part; V.bar = W.bar; V.stats = W.stats;
  BA_TRY(cudaMemsetAsync(W.bar, 0, sizeof(unsigned), st));
  void* args[] = {&V};
  BA_TRY(cudaLaunchCooperativeKernel((void*)window_motion_ba_kernel, dim3(W.n_frames), dim3(kWinThreads), args, 0, st));
  if (stats8) {
    BA_TRY(cudaMemcpyAsync(W.stats_host, W.stats, 8 * sizeof(double), cudaMemcpyDeviceToHost, st));
    BA_TRY(cudaStreamSynchronize(st));
    for (int i = 0; i < 8; ++i) stats8[i] = W.stats_host[i];
  }
  return check_launch("window_motion_ba_kernel");
}

extern "C" int vslam_window_read_cams(vslam_window* h, double* dst, int first, int count, void* stream) {
  if (!h || !dst || first < 0 || count < 0 || first + count > h->impl.n_frames) return VSLAM_ERR_BAD_ARG;
  if (count)
    BA_TRY(cudaMemcpyAsync(dst, h->impl.cams + 7 * (size_t)first, (size_t)count * 7 * sizeof(double),
                           cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  return VSLAM_OK;
}

extern "C" int vslam_window_info(vslam_window* h, int32_t* out4, double** cams_dev, double** stats_dev) {
  if (!h) return VSLAM_ERR_BAD_ARG;
  Window& W = h->impl;
  if (out4) { out4[0] = W.n_frames; out4[1] = W.n_pts; out4[2] = W.max_frames; out4[3] = W.slot; }
  if (cams_dev) *cams_dev = W.cams;
  if (stats_dev) *stats_dev = W.stats;
  return VSLAM_OK;
}

// ---- vslam_pcg_solve without any allocation -------------------------------------------------------------------
// Added (no counterpart in the reference): callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it) get the stand-alone PCG solve with
// its scratch (grid-barrier counter, flags, partial sums, the published search direction and the local S u rows) in
// a file-scope device array.  Every GPU that loads the library holds kPcgStaticBytes (24.25 KiB) until exit.  All calls
// share the array and it keeps the last call's contents: the header (barrier counter, flags, partials) is cleared
// on the caller's stream before every launch; the vectors are fully written by the kernel before it reads them.
// Calls that overlap on the GPU (two streams, two host threads) race on it -- serialise them.
constexpr int kPcgStaticMaxN = 1536;                                   // unknowns (a multiple of 6): 256 free cameras
constexpr size_t kPcgStaticHeader = 256;                               // bar (u32) @0, flags (4 x i32) @16, part (4 x f64) @64
constexpr size_t kPcgStaticBytes = kPcgStaticHeader + 2 * (size_t)kPcgStaticMaxN * sizeof(double);
static __device__ __align__(256) unsigned char g_pcg_static_scratch[kPcgStaticBytes];

// In host code the array's name is only a stand-in for the symbol; its device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int pcg_static_scratch(unsigned char** out) {
  constexpr int kMaxDevices = 64;
  static unsigned char* addr[kMaxDevices] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_cuda_error(e, "cudaGetDevice"); return VSLAM_ERR_CUDA; }
  if (dev < 0 || dev >= kMaxDevices) return VSLAM_ERR_BAD_ARG;
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_pcg_static_scratch);
    if (e != cudaSuccess || !p) { set_cuda_error(e, "cudaGetSymbolAddress(g_pcg_static_scratch)"); return VSLAM_ERR_CUDA; }
    addr[dev] = static_cast<unsigned char*>(p);
  }
  *out = addr[dev];
  return VSLAM_OK;
}

extern "C" int vslam_pcg_solve_static_max_n(void) { return kPcgStaticMaxN; }

extern "C" int vslam_pcg_solve_static(const double* S, const double* b, double* x, int n, double tol, int max_iter,
                                      int32_t* status_iters_dev, void* stream) {
  if (!S || !b || !x || n <= 0 || n % 6 || !status_iters_dev || !(tol > 0) || max_iter < 0) return VSLAM_ERR_BAD_ARG;
  if (n > kPcgStaticMaxN) {
    snprintf_last_error("vslam_pcg_solve_static: n = %d exceeds the static scratch capacity of %d unknowns", n, kPcgStaticMaxN);
    return VSLAM_ERR_TOO_LARGE;
  }
  unsigned char* base = nullptr;
  int rc = pcg_static_scratch(&base);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  BA_TRY(cudaMemsetAsync(base, 0, kPcgStaticHeader, st));             // what must start at zero: barrier counter, flags, partials
  PcgArgs A{};
  A.S = S; A.b = b; A.x = x; A.n = n; A.rows = n; A.G = 1; A.max_iter = max_iter; A.tol = tol;
  A.bar = reinterpret_cast<unsigned*>(base);
  A.fail_flag = reinterpret_cast<int*>(base + 16);
  A.part = reinterpret_cast<double*>(base + 64);
  A.p_glob = reinterpret_cast<double*>(base + kPcgStaticHeader);
  A.w_loc = A.p_glob + kPcgStaticMaxN;
  A.status = status_iters_dev; A.iters = status_iters_dev + 1;         // outputs go straight to the caller's device buffer
  if (n <= kPcgSmallMax) {
    pcg_small_fn()<<<1, kPcgSmallThreads, pcg_small_smem(n), st>>>(A);
    return check_launch("ba_pcg_small_kernel");
  }
  if (n <= kPcgMidMax && n % 6 == 0) {
    ba_pcg_mid_kernel<<<1, kPcgMidThreads, kPcgMidSmemDoubles * sizeof(double), st>>>(A);
    return check_launch("ba_pcg_mid_kernel");
  }
  {
    const PcgClusterPlan P = pcg_cluster_plan(n / 6, pcg_smem_budget());
    if (P.kind) {
      A.rows = 6 * P.blocks; A.G = P.ctas;
      BA_TRY(launch_pcg_cluster(P.kind, P.ctas, P.smem, A, st));
      return check_launch("ba_pcg_cluster_kernel");
    }
  }
  const size_t smem_slab = sizeof(double) * ((size_t)n + 12 * (size_t)n + 8 + (size_t)n * n);
  const size_t smem_noslab = sizeof(double) * ((size_t)n + 12 * (size_t)n + 8);
  A.slab_in_smem = smem_slab <= 200 * 1024;
  const size_t smem = A.slab_in_smem ? smem_slab : smem_noslab;
  BA_TRY(pcg_kernel_allow_max_smem());
  ba_pcg_kernel<<<1, kPcgThreads, smem, st>>>(A);
  return check_launch("ba_pcg_kernel");
}
