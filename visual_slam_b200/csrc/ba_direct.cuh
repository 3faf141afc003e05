// Direct solve of a SMALL reduced camera system (n = 6 * free cameras <= kDirectMax, e.g. the 9 free key frames of
// BASELINE configs[2]) in ONE CTA: what g2o's LinearSolverCholmodSE3 does for such a window (reference
// src/v2/LocalBA.py:23) -- a factorisation that reports a non-positive pivot as a failed solve (the case behind the
// reference's debug.txt) -- instead of ~20 PCG iterations.  PCG stays the solver of large systems (north_star; C4).
// The system S, bhat is read from HBM, where the Schur kernels leave it (ba_schur_units / ba_compute_G finish it).
// Two kernels: register-resident Gauss-Jordan elimination for n <= 63 (below), and for 63 < n <= 96
//   factor     right-looking LDL^T in shared memory, one pivot per step; the right-hand side rides along as row n,
//              so the forward substitution costs nothing extra
//   backward   L^T x = y by warp 0
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
  int n, ld;
  const int* done;                 // optional (&LMState::done): the solve is skipped when the LM loop has terminated
  const double* S; const double* b;   // the system (row-major n x n, symmetric) and its right-hand side
  double* x; int* status; int* iters;
};

// Right-looking LDL^T, one pivot per step: the work per pivot is a rank-1 update of the trailing triangle -- one
// FMA per entry, all entries independent -- so the sequential depth is the 54 pivots themselves and not the
// arithmetic inside a block factorisation (a blocked variant with redundant 6x6 inverses measured 27 us at n = 54:
// ~170 dependent FP64 instructions per block step in every warp).  The lower triangle holds the unscaled columns,
// the UPPER triangle receives L^T (l_ik = a_ik / d_k at [k][i]), so a step needs no second barrier; the reciprocal of
// the NEXT pivot is formed by every thread from three entries while the update of the current step is in flight.
// The right-hand side is row n: its scaled entries v = D^-1 L^-1 b land at [k][n], and the backward pass
// x_k = v_k - sum_{i > k} l_ik x_i runs column by column in warp 0 (no divisions, no reductions).
__global__ void __launch_bounds__(kDirectThreads) ba_solve_direct_kernel(DirectArgs A) {
  extern __shared__ __align__(16) double Ls[];
  const int n = A.n, ld = A.ld, tid = threadIdx.x;
  uint16_t* tri = reinterpret_cast<uint16_t*>(Ls + (size_t)(n + 1) * ld);   // entry e of the triangle -> (row << 8 | col)
  pdl_wait();
  pdl_trigger();
  if (A.done != nullptr && *A.done) return;
  {
    const int m = n + 1, total = m * (m + 1) / 2;
    for (int e = tid; e < total; e += kDirectThreads) {
      int i = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
      while ((i + 1) * (i + 2) / 2 <= e) ++i;
      while (i * (i + 1) / 2 > e) --i;
      tri[e] = (uint16_t)((i << 8) | (e - i * (i + 1) / 2));
    }
  }
  for (int r = tid >> 5; r < n; r += kDirectThreads / 32)
    for (int c = tid & 31; c < n; c += 32) Ls[r * ld + c] = __ldcg(A.S + (size_t)r * n + c);
  for (int i = tid; i < n; i += kDirectThreads) Ls[n * ld + i] = __ldcg(A.b + i);
  __syncthreads();
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


// ---- Register-resident Gauss-Jordan elimination (n <= kGJMax): the faster of the two direct kernels.
// The LDL^T kernel above keeps the matrix in shared memory, so each of its n pivot steps is a shared-memory
// read-modify-write of the trailing triangle behind a table lookup, and a backward pass of n more dependent steps
// follows (33 us at n = 54).  Here the augmented system [b | S] lives in REGISTERS as 4 x 4 tiles: thread (tr, tc)
// of 16 x NT owns rows 4 tr .. 4 tr + 3 and augmented columns 4 tc .. 4 tc + 3 (augmented column 0 is the
// right-hand side, column j + 1 matrix column j).  Pivot step k: the owners of row k and of column k have published
// them in a double-buffered shared-memory strip (two 128-bit stores per thread), ONE named barrier, then every
// thread updates its tile from four 128-bit broadcast loads (rows above AND below the pivot are eliminated, so there
// is no backward pass).  The NEXT pivot is formed by every thread from three published entries, so no thread waits
// for a reciprocal to be broadcast.  The pivot loop is unrolled by four (the period of the local tile indices, which
// keeps every register index a compile-time constant).  A warp whose two tile columns lie left of the pivot has
// nothing left to do, arrives at the next barrier without waiting, and exits.  The pivots are those of Gaussian
// elimination without pivoting = the D of LDL^T: a pivot that is not positive reports status 1 exactly like the
// factorisation does (g2o: "Cholesky failure").  Entries of finished columns are not maintained (they may hold
// anything, nobody reads them).  Fixed work assignment, no atomics: bit-reproducible.
// Measured on B200 at n = 54 (clock64 stamps): 3.3k cycles for the load, 27k cycles = ~500 per pivot for the loop,
// 17 us for the launch.  The pivot step is NOT bound by its arithmetic (8-cycle DFMA, 17-cycle MUFU.RCP64H) but by
// the publish -> barrier -> read round trip: tools/probes/bar_sts.cu measures 150 cycles for two stores per warp
// and 290 for five at 256 threads (a block barrier drains every store in flight, ~30 cycles each), 95 at 64
// threads; a strip layout (one row per four threads, 14 stores per step), a fully unrolled pivot loop (80 KB of
// straight-line code) and an earlier look-ahead all measured 440-550 cycles per pivot.  PCG at the same size needs
// 21 iterations of ~1 us, so the two are level (22 us) and PCG stays the default (north_star).
constexpr int kGJMax = 63;
constexpr int kGJThreads = 256;

__device__ __forceinline__ void named_bar_sync(int count) { asm volatile("bar.sync 1, %0;" ::"r"(count) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(int count) { asm volatile("bar.arrive 1, %0;" ::"r"(count) : "memory"); }

// warps (of two tile columns each) that still take part in pivot step kk: warp 0 (right-hand side) and every warp
// whose second tile column is not left of the pivot's
__host__ __device__ constexpr int gj_active_warps(int kk, int nw) {
  const int kc = kk >> 2;
  int first = kc >> 1;                 // first warp w >= 1 with 2 w + 1 >= kc
  if (first < 1) first = 1;
  return 1 + (nw > first ? nw - first : 0);
}

__global__ void __launch_bounds__(kGJThreads) ba_solve_gj_kernel(DirectArgs A) {
  __shared__ __align__(16) double strips[4 * 64 + 2];
  const int n = A.n, tid = threadIdx.x;
  const int NT = (n + 4) >> 2;                       // tile columns: 4 NT >= n + 1
  const int NW = (NT + 1) >> 1;                      // eliminating warps (two tile columns each)
  double* prow = strips;                             // [2][64]  pivot row, augmented column index
  double* pcol = prow + 2 * 64;                      // [2][64]  pivot column, row index
  double* pdia = pcol + 2 * 64;                      // [2]      diagonal entry of the pivot after the next one
  pdl_wait();
  pdl_trigger();
  if (A.done != nullptr && *A.done) return;
  const int wid = tid >> 5;
  if (wid >= NW) return;
  const int tr = tid & 15, tc = tid >> 4;            // warp w holds tile columns 2 w, 2 w + 1
  double a[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int row = 4 * tr + r, jj = 4 * tc + c;
      double v = 0.0;
      if (row < n) {
        if (jj == 0) v = __ldcg(A.b + row);
        else if (jj <= n) v = __ldcg(A.S + (size_t)row * n + jj - 1);
      }
      a[r][c] = v;
    }
  }
  // publish pivot step kk = 1 (matrix row / column 0) into buffer 1, and the diagonal entry of the step after it
  if (tr == 0) {
    *reinterpret_cast<double2*>(prow + 64 + 4 * tc) = make_double2(a[0][0], a[0][1]);
    *reinterpret_cast<double2*>(prow + 64 + 4 * tc + 2) = make_double2(a[0][2], a[0][3]);
  }
  bool ok = true;
  double dinv[4] = {0.0, 0.0, 0.0, 0.0}, inv = 1.0;
  if (tc == 0) {
    *reinterpret_cast<double2*>(pcol + 64 + 4 * tr) = make_double2(a[0][1], a[1][1]);
    *reinterpret_cast<double2*>(pcol + 64 + 4 * tr + 2) = make_double2(a[2][1], a[3][1]);
    if (tr == 0) pdia[0] = a[1][2];                  // S[1][1], read by step kk = 1
  }
  {
    const double d0 = __ldcg(A.S);                   // the first pivot, by every thread
    ok = (d0 > 1e-290) && (d0 < 1e290);
    inv = fast_rcp(d0);
  }
#pragma unroll 1
  for (int q = 0; q < NT; ++q) {
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      const int kk = 4 * q + l;                      // augmented pivot column; matrix index k = kk - 1
      if (kk >= 1 && kk <= n) {
        const int kr = (l == 0) ? q - 1 : q;         // tile row of the pivot row k
        const int krl = (l + 3) & 3;                 // its local row (compile-time)
        const double* pr = prow + (l & 1) * 64;
        const double* pc = pcol + (l & 1) * 64;
        named_bar_sync(32 * gj_active_warps(kk > 1 ? kk - 1 : 1, NW));
        const double2 c01 = *reinterpret_cast<const double2*>(pc + 4 * tr);
        const double2 c23 = *reinterpret_cast<const double2*>(pc + 4 * tr + 2);
        const double2 r01 = *reinterpret_cast<const double2*>(pr + 4 * tc);
        const double2 r23 = *reinterpret_cast<const double2*>(pr + 4 * tc + 2);
        double f[4] = {c01.x * inv, c01.y * inv, c23.x * inv, c23.y * inv};
        if (tr == kr) { f[krl] = 0.0; dinv[krl] = inv; }
        const double rv[4] = {r01.x, r01.y, r23.x, r23.y};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
          for (int c = 0; c < 4; ++c) a[r][c] = fma(-f[r], rv[c], a[r][c]);
        }
        // row / column of the next pivot into the other buffer, and the diagonal entry of the one after
        const int kn = kk + 1, km = kk + 2;
        double inv_next = 1.0;
        if (kn <= n) {
          double* qr = prow + ((l + 1) & 1) * 64;
          double* qc = pcol + ((l + 1) & 1) * 64;
          const int nr = q;                          // pivot row k + 1 = kk as a row index: tile row q, local row l
          const int nc = (l == 3) ? q + 1 : q;       // pivot column kn: tile column, local column (l + 1) & 3
          if (tr == nr) {
            *reinterpret_cast<double2*>(qr + 4 * tc) = make_double2(a[l][0], a[l][1]);
            *reinterpret_cast<double2*>(qr + 4 * tc + 2) = make_double2(a[l][2], a[l][3]);
          }
          if (tc == nc) {
            *reinterpret_cast<double2*>(qc + 4 * tr) = make_double2(a[0][(l + 1) & 3], a[1][(l + 1) & 3]);
            *reinterpret_cast<double2*>(qc + 4 * tr + 2) = make_double2(a[2][(l + 1) & 3], a[3][(l + 1) & 3]);
          }
          if (km <= n) {                             // entry (row k + 2 = kn, column km)
            const int mc = (l >= 2) ? q + 1 : q;
            if (tr == nc && tc == mc) pdia[l & 1] = a[(l + 1) & 3][(l + 2) & 3];
          }
          // the next pivot, formed by every thread from the published entries exactly as its owner has just
          // updated it, so that no thread waits for a reciprocal to be broadcast
          const double dn = fma(-(pc[kk] * inv), pr[kn], pdia[(l + 1) & 1]);
          ok = ok && (dn > 1e-290) && (dn < 1e290);
          inv_next = fast_rcp(dn);
          // a warp that has nothing to do from the next step on still owes that step's barrier its arrival
          if (wid != 0 && 2 * wid + 1 < (kn >> 2)) { named_bar_arrive(32 * gj_active_warps(kk, NW)); return; }
        }
        inv = inv_next;
      }
    }
  }
  const bool fail = !ok;
  if (tc == 0) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
      if (4 * tr + r < n) A.x[4 * tr + r] = fail ? 0.0 : a[r][0] * dinv[r];
  }
  if (tid == 0) { *A.status = fail ? 1 : 0; *A.iters = 0; }
}

}  // namespace vslam
