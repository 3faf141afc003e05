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
constexpr int kDirectThreads = 512;

// (n + 1) rows x ld (odd: column walks are conflict-free) of the system + right-hand side row, and the (row, column)
// table of the trailing triangle's entries
static inline int direct_ld(int n) { return (n + 2) | 1; }
static inline size_t direct_smem(int n) {
  const size_t m = (size_t)n + 1;
  return sizeof(double) * m * (size_t)direct_ld(n) + sizeof(uint16_t) * (m * (m + 1) / 2 + 2);
}

struct DirectArgs {
  int n_seg, n, ld;
  const int* seg_a; const int* seg_b; const int* seg_unit_start; const double* partial; const double* Hpp;
  const double* extra; const int* seg_extra_idx;
  double* S;            // HBM copy of the assembled system (row-major n x n)
  BhatArgs Bh;
  double* x; int* status; int* iters;
  const double* dense_S; const double* dense_b;   // stand-alone use (vslam_direct_solve): the system as given
};

// Right-looking LDL^T, one pivot per step: the work per pivot is a rank-1 update of the trailing triangle -- one
// FMA per entry, all entries independent -- so the sequential depth is the 54 pivots themselves and not the
// arithmetic inside a block factorisation (a blocked variant with redundant 6x6 inverses measured 27 us at n = 54:
// ~170 dependent FP64 instructions per block step in every warp).  The lower triangle holds the unscaled columns,
// the UPPER triangle receives L^T (l_ik = a_ik / d_k at [k][i]), so a step needs no second barrier; the reciprocal of
// the NEXT pivot is formed by every thread from three entries while the update of the current step is in flight.
// The right-hand side is row n: its scaled entries v = D^-1 L^-1 b land at [k][n], and the backward pass
// x_k = v_k - sum_{i > k} l_ik x_i runs column by column in warp 0 (no divisions, no reductions).
__global__ void __launch_bounds__(kDirectThreads) ba_solve_direct_kernel(BAView V, DirectArgs A) {
  extern __shared__ __align__(16) double Ls[];
  const int n = A.n, ld = A.ld, tid = threadIdx.x;
  uint16_t* tri = reinterpret_cast<uint16_t*>(Ls + (size_t)(n + 1) * ld);   // entry e of the triangle -> (row << 8 | col)
  pdl_wait();
  pdl_trigger();
  if (tid == 0) {
    if (A.Bh.pcg_bar) *A.Bh.pcg_bar = 0u;
    if (A.Bh.pcg_flags) { A.Bh.pcg_flags[0] = 0; A.Bh.pcg_flags[1] = 0; A.Bh.pcg_flags[2] = 0; A.Bh.pcg_flags[3] = 0; }
  }
  const bool standalone = A.dense_S != nullptr;
  if (!standalone && V.st->done) return;
  const double lambda = standalone ? 0.0 : V.st->lambda;
  const double* __restrict__ partial = A.partial;
  const double* __restrict__ Hpp = A.Hpp;
  const double* __restrict__ extra = A.extra;
  // ---- assembly (ba_schur_finalize's arithmetic); every entry of the lower triangle that belongs to a present
  // block is written exactly once, absent blocks are zeroed first
  for (int i = tid; i < (n + 1) * ld; i += kDirectThreads) Ls[i] = 0.0;
  {
    const int m = n + 1, total = m * (m + 1) / 2;
    for (int e = tid; e < total; e += kDirectThreads) {
      int i = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
      while ((i + 1) * (i + 2) / 2 <= e) ++i;
      while (i * (i + 1) / 2 > e) --i;
      tri[e] = (uint16_t)((i << 8) | (e - i * (i + 1) / 2));
    }
  }
  __syncthreads();
  if (standalone) {
    for (int i = tid; i < n * n; i += kDirectThreads) Ls[(i / n) * ld + (i % n)] = A.dense_S[i];
    for (int i = tid; i < n; i += kDirectThreads) Ls[n * ld + i] = A.dense_b[i];
  } else {
    for (int t = tid; t < A.n_seg * 36; t += kDirectThreads) {
      const int seg = t / 36, e = t - 36 * seg;
      const int a = A.seg_a[seg], b = A.seg_b[seg], r = e / 6, c = e - 6 * r;
      const int ex = A.seg_extra_idx[seg];
      const int u0 = A.seg_unit_start[seg], u1 = A.seg_unit_start[seg + 1];
      double v = ex >= 0 ? extra[(size_t)ex * 36 + e] : 0.0;
      if (a == b) v += Hpp[36 * a + e] + (r == c ? lambda : 0.0);
      double s = 0.0;
      for (int u = u0; u < u1; ++u) s += partial[(size_t)u * 36 + e];
      v -= s;
      Ls[(6 * a + r) * ld + 6 * b + c] = v;
      if (a != b) Ls[(6 * b + c) * ld + 6 * a + r] = v;
    }
    for (int t = tid; t < V.C * 6; t += kDirectThreads) {
      const int c = t / 6, i = t - 6 * c;
      const int col = V.cam_col[c];
      if (col < 0) continue;
      const double s = bhat_partial_sum(V, A.Bh, c, i);
      Ls[n * ld + 6 * col + i] = (A.Bh.bp[6 * col + i] + A.Bh.bp_extra[6 * col + i]) - s;
    }
  }
  __syncthreads();
  if (!standalone) {      // HBM copies for the stage API (vslam_ba_read) -- off the critical path
    for (int i = tid; i < n * n; i += kDirectThreads) A.S[i] = Ls[(i / n) * ld + (i % n)];
    for (int i = tid; i < n; i += kDirectThreads) A.Bh.bhat[i] = Ls[n * ld + i];
    __syncthreads();      // (the upper triangle is overwritten below)
  }
  bool ok = true;
  double inv;
  {
    const double d0 = Ls[0];
    ok = (d0 > 1e-290) && (d0 < 1e290);
    inv = fast_rcp(ok ? d0 : 1.0);
  }
  for (int k = 0; k < n; ++k) {
    const double* colk = Ls + k;                      // a_ik = colk[i * ld]
    double inv_next = 1.0;
    if (k + 1 < n) {                                  // pivot k + 1 after this step's update, by every thread
      const double a10 = colk[(k + 1) * ld];
      const double dn = fma(-(a10 * inv), a10, Ls[(k + 1) * ld + k + 1]);
      ok = ok && (dn > 1e-290) && (dn < 1e290);
      inv_next = fast_rcp(ok ? dn : 1.0);
    }
    const int m0 = k + 1, m = n - k;                  // rows m0 .. n
    for (int r = tid; r < m; r += kDirectThreads) Ls[k * ld + m0 + r] = colk[(m0 + r) * ld] * inv;   // L^T row k
    const int total = m * (m + 1) / 2 - 1;            // trailing triangle without the entry (n, n)
    // (entry 0 = the next pivot: every thread has just formed its updated value in registers, and nobody reads
    // the stored one again, so it is left alone -- no thread writes what another one is reading in this step)
    for (int e = 1 + tid; e < total; e += kDirectThreads) {
      const int ij = tri[e], i = m0 + (ij >> 8), j = m0 + (ij & 255);
      Ls[i * ld + j] = fma(-(colk[i * ld] * inv), colk[j * ld], Ls[i * ld + j]);
    }
    inv = inv_next;
    __syncthreads();
  }
  const bool fail = !ok;
  if (tid < 32 && !fail) {
    for (int i = n - 1; i >= 1; --i) {                // x_i is final: fold it into the rows above
      const double xi = Ls[i * ld + n];
      for (int k = tid; k < i; k += 32) Ls[k * ld + n] = fma(-Ls[k * ld + i], xi, Ls[k * ld + n]);
      __syncwarp();
    }
  }
  __syncthreads();
  for (int i = tid; i < n; i += kDirectThreads) A.x[i] = fail ? 0.0 : Ls[i * ld + n];
  if (tid == 0) { *A.status = fail ? 1 : 0; *A.iters = 0; }
}

}  // namespace vslam
