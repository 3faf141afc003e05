// Direct solve of a SMALL reduced camera system (n = 6 * free cameras <= kDirectMax, e.g. the 9 free key frames of
// BASELINE configs[2]) in ONE CTA, with the assembly of the system fused in front of it: what g2o's
// LinearSolverCholmodSE3 does for such a window (reference src/v2/LocalBA.py:23) -- a Cholesky factorisation that
// reports a non-positive pivot as a failed solve (the case behind the reference's debug.txt) -- instead of ~20
// PCG iterations behind a separate assembly launch.  PCG stays the solver of large systems (north_star; C4).
//
//   prologue   S = [a == b](Hpp + lambda I) + scale-edge blocks - sum of the Schur unit partials, bhat = bp + bp_extra
//              - sum of the per-chunk G y partials: the statements of ba_schur_finalize, written into shared memory
//              (and to HBM for the stage API / tests)
//   factor     blocked right-looking Cholesky, 6x6 block columns: every thread factors the diagonal block redundantly
//              in registers (no broadcast round trip), one thread per row below solves its 6-wide row against it,
//              then all threads update the trailing triangle; the right-hand side rides along as row n, so the
//              forward substitution L y = b costs nothing extra
//   backward   L^T x = y by warp 0, one block at a time (six warp reductions + a 6x6 triangular solve)
// Fixed work assignment, no atomics: bit-reproducible.
#pragma once
#include "ba_kernels.cuh"
#include "ba_pcg.cuh"

namespace vslam {

constexpr int kDirectMax = 96;
constexpr int kDirectThreads = 256;

static inline size_t direct_smem(int n) { return sizeof(double) * (size_t)(n + 1) * (size_t)(n + 2); }

struct DirectArgs {
  int n_seg, n, ld;
  const int* seg_a; const int* seg_b; const int* seg_unit_start; const double* partial; const double* Hpp;
  const double* extra; const int* seg_extra_idx;
  double* S;            // HBM copy of the assembled system (row-major n x n)
  BhatArgs Bh;
  double* x; int* status; int* iters;
  const double* dense_S; const double* dense_b;   // stand-alone use (vslam_direct_solve): the system as given
};

__global__ void __launch_bounds__(kDirectThreads) ba_solve_direct_kernel(BAView V, DirectArgs A) {
  extern __shared__ __align__(16) double Ls[];      // (n + 1) rows x ld: lower triangle + right-hand side row
  __shared__ int s_fail;
  const int n = A.n, ld = A.ld, tid = threadIdx.x;
  pdl_wait();
  pdl_trigger();
  if (tid == 0) {
    s_fail = 0;
    if (A.Bh.pcg_bar) *A.Bh.pcg_bar = 0u;
    if (A.Bh.pcg_flags) { A.Bh.pcg_flags[0] = 0; A.Bh.pcg_flags[1] = 0; A.Bh.pcg_flags[2] = 0; A.Bh.pcg_flags[3] = 0; }
  }
  const bool standalone = A.dense_S != nullptr;
  if (!standalone && V.st->done) return;
  const double lambda = standalone ? 0.0 : V.st->lambda;
  for (int i = tid; i < (n + 1) * ld; i += kDirectThreads) Ls[i] = 0.0;
  __syncthreads();
  if (standalone) {
    for (int i = tid; i < n * n; i += kDirectThreads) Ls[(i / n) * ld + (i % n)] = A.dense_S[i];
    for (int i = tid; i < n; i += kDirectThreads) Ls[n * ld + i] = A.dense_b[i];
  }
  // ---- assembly (ba_schur_finalize's arithmetic)
  for (int t = tid; t < (standalone ? 0 : A.n_seg * 36); t += kDirectThreads) {
    const int seg = t / 36, e = t - 36 * seg;
    const int a = A.seg_a[seg], b = A.seg_b[seg], r = e / 6, c = e - 6 * r;
    const int ex = A.seg_extra_idx[seg];
    double v = ex >= 0 ? A.extra[(size_t)ex * 36 + e] : 0.0;
    if (a == b) v += A.Hpp[36 * a + e] + (r == c ? lambda : 0.0);
    double s = 0.0;
    for (int u = A.seg_unit_start[seg]; u < A.seg_unit_start[seg + 1]; ++u) s += A.partial[(size_t)u * 36 + e];
    v -= s;
    Ls[(6 * a + r) * ld + 6 * b + c] = v;
    if (a != b) Ls[(6 * b + c) * ld + 6 * a + r] = v;
    A.S[(size_t)(6 * a + r) * n + 6 * b + c] = v;
    if (a != b) A.S[(size_t)(6 * b + c) * n + 6 * a + r] = v;
  }
  for (int t = tid; t < (standalone ? 0 : V.C * 6); t += kDirectThreads) {
    const int c = t / 6, i = t - 6 * c;
    const int col = V.cam_col[c];
    if (col < 0) continue;
    double s = 0.0;
    if (A.Bh.camp != nullptr) {
      const int n_obs = V.cm_start[c + 1] - V.cm_start[c];
      const int n_chunks = (n_obs + kCamChunk - 1) / kCamChunk;
      for (int k = 0; k < n_chunks; ++k) s += A.Bh.camp[((size_t)c * V.max_chunks + k) * kCamVals + i];
    }
    const double v = (A.Bh.bp[6 * col + i] + A.Bh.bp_extra[6 * col + i]) - s;
    Ls[n * ld + 6 * col + i] = v;
    A.Bh.bhat[6 * col + i] = v;
  }
  __syncthreads();
  // ---- blocked Cholesky; row n (the right-hand side) is carried as one more row of the lower triangle
  const int nb = n / 6;
  for (int k = 0; k < nb; ++k) {
    const int kb = 6 * k;
    // diagonal block, factored by every thread in registers: D = C C^T, ri = 1 / diag(C)
    double Cf[6][6], ri[6];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j <= i; ++j) Cf[i][j] = Ls[(kb + i) * ld + kb + j];
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      double d = Cf[j][j];
#pragma unroll
      for (int q = 0; q < j; ++q) d = fma(-Cf[j][q], Cf[j][q], d);
      ok = ok && (d > 1e-290) && (d < 1e290);
      ri[j] = fast_rsqrt(ok ? d : 1.0);
      Cf[j][j] = d * ri[j];
#pragma unroll
      for (int i = j + 1; i < 6; ++i) {
        double t = Cf[i][j];
#pragma unroll
        for (int q = 0; q < j; ++q) t = fma(-Cf[i][q], Cf[j][q], t);
        Cf[i][j] = t * ri[j];
      }
    }
    if (!ok) { if (tid == 0) s_fail = 1; }
    __syncthreads();                                  // everybody has read the diagonal block
    if (s_fail) break;
    // rows kb .. n: the block's own rows take the factor, the rows below solve l C^T = a
    if (tid < 6) {                                    // (static indices only: the factor stays in registers)
      double* row = Ls + (kb + tid) * ld + kb;
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        double v = 0.0;
#pragma unroll
        for (int i = j; i < 6; ++i) v = (i == tid) ? Cf[i][j] : v;
        row[j] = v;
      }
    }
    for (int r = kb + 6 + tid; r <= n; r += kDirectThreads) {
      double* row = Ls + r * ld + kb;
      double l[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        double t = row[j];
#pragma unroll
        for (int q = 0; q < j; ++q) t = fma(-l[q], Cf[j][q], t);
        l[j] = t * ri[j];
      }
#pragma unroll
      for (int j = 0; j < 6; ++j) row[j] = l[j];
    }
    __syncthreads();
    // trailing update of the lower triangle (rows i > kb + 5, columns kb + 6 .. min(i, n - 1))
    const int m0 = kb + 6, m = n + 1 - m0;            // rows m0 .. n
    const int total = m * (m + 1) / 2 - 1;            // triangle without its last entry (row n, column n)
    for (int e = tid; e < total; e += kDirectThreads) {
      int i = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
      while ((i + 1) * (i + 2) / 2 <= e) ++i;
      while (i * (i + 1) / 2 > e) --i;
      const int j = e - i * (i + 1) / 2;
      const double* ri_ = Ls + (m0 + i) * ld + kb;
      const double* rj_ = Ls + (m0 + j) * ld + kb;
      double s = Ls[(m0 + i) * ld + m0 + j];
#pragma unroll
      for (int q = 0; q < 6; ++q) s = fma(-ri_[q], rj_[q], s);
      Ls[(m0 + i) * ld + m0 + j] = s;
    }
    __syncthreads();
  }
  const bool fail = s_fail != 0;
  // ---- backward substitution L^T x = y (y = row n), one block at a time, warp 0
  if (tid < 32 && !fail) {
    double* y = Ls + n * ld;
    for (int k = nb - 1; k >= 0; --k) {
      const int kb = 6 * k;
      double acc[6] = {0, 0, 0, 0, 0, 0};
      for (int i = kb + 6 + tid; i < n; i += 32) {
        const double xi = y[i];
#pragma unroll
        for (int c = 0; c < 6; ++c) acc[c] = fma(Ls[i * ld + kb + c], xi, acc[c]);
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) acc[c] = warp_sum(acc[c]);
      // x_k = C^-T (y_k - acc), C = the block's factor (lower): solve from the last row up; every lane does it
      double xk[6];
#pragma unroll
      for (int c = 5; c >= 0; --c) {
        double t = y[kb + c] - acc[c];
#pragma unroll
        for (int q = c + 1; q < 6; ++q) t = fma(-Ls[(kb + q) * ld + kb + c], xk[q], t);
        xk[c] = t / Ls[(kb + c) * ld + kb + c];
      }
      __syncwarp();
      if (tid == 0) {
#pragma unroll
        for (int c = 0; c < 6; ++c) y[kb + c] = xk[c];
      }
      __syncwarp();
    }
  }
  __syncthreads();
  for (int i = tid; i < n; i += kDirectThreads) A.x[i] = fail ? 0.0 : Ls[n * ld + i];
  if (tid == 0) { *A.status = fail ? 1 : 0; *A.iters = 0; }
}

}  // namespace vslam
