// Preconditioned conjugate gradients on the dense reduced camera system S x = b (n = 6 * free cams),
// block-Jacobi (6x6) preconditioner.  Replaces g2o LinearSolverCholmodSE3 (reference
// src/v2/LocalBA.py:23) as BASELINE.json's north_star prescribes.
//
// One cooperative launch: CTA g owns a slab of `rows` rows of S, kept in shared memory for the whole
// solve, so an iteration moves only the search direction (n doubles) through L2.  Dot products are
// per-CTA partials summed in a fixed order by every CTA after a grid barrier -> bit-reproducible.
// A non-positive curvature p^T S p (or a non-SPD diagonal block) reports status 1: the LM loop
// treats that exactly like g2o treats a failed Cholesky (rejected step; cf. reference debug.txt).
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace vslam {

struct PcgArgs {
  const int* done;    // optional (&LMState::done): the solve is skipped when the LM loop has terminated
  const double* S; const double* b; double* x;
  int n, rows, G, slab_in_smem, max_iter;
  int pipelined;      // G > 1: one grid barrier per iteration (pipelined PCG); p_glob holds 2 n, part 6 G doubles
  double tol;
  double* p_glob;     // [n]  published u = M^-1 r
  double* w_loc;      // [n]  rows of S u, each CTA touches only its own rows
  double* part;       // [3 * G]
  unsigned* bar;      // grid barrier counter (zeroed before launch)
  int* fail_flag;     // zeroed before launch
  int* status;        // out: 0 converged, 1 breakdown (indefinite), 2 max_iter reached
  int* iters;         // out
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Arrival is a release reduction (orders the CTA's earlier writes, made visible to thread 0 by the block barrier,
// before the count) and the poll an acquire load: no separate membar.gl on either side (two full fences per
// barrier cost ~1 us of the ~2 us a barrier took; an iteration has two barriers).
__device__ __forceinline__ void red_release_add_u32(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& target, int G) {
  __syncthreads();
  if (G > 1 && threadIdx.x == 0) {
    target += (unsigned)G;
    red_release_add_u32(bar, 1u);
    while (ld_acquire_u32(bar) < target) {}
  }
  __syncthreads();
}

// In-place inverse of a 6x6 SPD matrix via Cholesky; returns false if a pivot is not positive (or not a
// finite normal number).  Columns are template instances so that every index is a constant and all three
// 6x6 arrays stay in registers (the looped form lived in local memory and cost ~12 us per call); one
// reciprocal square root per pivot instead of a sqrt and divisions.
template <int J>
__device__ __forceinline__ void spd6_chol_col(const double (&A)[36], double (&L)[36], double (&ri)[6], bool& ok) {
  double d = A[7 * J];
#pragma unroll
  for (int k = 0; k < J; ++k) d = fma(-L[6 * J + k], L[6 * J + k], d);
  ok = ok && (d > 1e-290) && (d < 1e290);
  ri[J] = fast_rsqrt(ok ? d : 1.0);
  L[7 * J] = d * ri[J];
#pragma unroll
  for (int i = J + 1; i < 6; ++i) {
    double t = A[6 * i + J];
#pragma unroll
    for (int k = 0; k < J; ++k) t = fma(-L[6 * i + k], L[6 * J + k], t);
    L[6 * i + J] = t * ri[J];
  }
}
template <int J>
__device__ __forceinline__ void spd6_linv_col(const double (&L)[36], const double (&ri)[6], double (&Li)[36]) {
  Li[7 * J] = ri[J];
#pragma unroll
  for (int i = J + 1; i < 6; ++i) {
    double t = 0.0;
#pragma unroll
    for (int k = J; k < i; ++k) t = fma(-L[6 * i + k], Li[6 * k + J], t);
    Li[6 * i + J] = t * ri[i];
  }
}
template <int I>
__device__ __forceinline__ void spd6_ltl_row(const double (&Li)[36], double (&A)[36]) {
#pragma unroll
  for (int j = I; j < 6; ++j) {
    double t = 0.0;
#pragma unroll
    for (int k = j; k < 6; ++k) t = fma(Li[6 * k + I], Li[6 * k + j], t);
    A[6 * I + j] = t;
    A[6 * j + I] = t;
  }
}
__device__ __forceinline__ bool spd6_inverse(double (&A)[36] /* row-major */) {
  double L[36], Li[36], ri[6];
  bool ok = true;
  spd6_chol_col<0>(A, L, ri, ok); spd6_chol_col<1>(A, L, ri, ok); spd6_chol_col<2>(A, L, ri, ok);
  spd6_chol_col<3>(A, L, ri, ok); spd6_chol_col<4>(A, L, ri, ok); spd6_chol_col<5>(A, L, ri, ok);
  if (!ok) return false;
  // inverse of L (lower), then A^-1 = L^-T L^-1
  spd6_linv_col<0>(L, ri, Li); spd6_linv_col<1>(L, ri, Li); spd6_linv_col<2>(L, ri, Li);
  spd6_linv_col<3>(L, ri, Li); spd6_linv_col<4>(L, ri, Li); spd6_linv_col<5>(L, ri, Li);
  spd6_ltl_row<0>(Li, A); spd6_ltl_row<1>(Li, A); spd6_ltl_row<2>(Li, A);
  spd6_ltl_row<3>(Li, A); spd6_ltl_row<4>(Li, A); spd6_ltl_row<5>(Li, A);
  return true;
}

constexpr int kPcgThreads = 256;

__device__ __forceinline__ double block_sum1(double v, double* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  for (int w = 0; w < kPcgThreads / 32; ++w) s += red[w];
  __syncthreads();
  return s;  // every thread holds the same, order-fixed sum
}

// Three block sums with one barrier pair; every thread ends with the same, order-fixed sums.
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c, double* red3) {
  warp_sum3(a, b, c);
  if ((threadIdx.x & 31) == 0) {
    const int w = threadIdx.x >> 5;
    red3[3 * w] = a; red3[3 * w + 1] = b; red3[3 * w + 2] = c;
  }
  __syncthreads();
  double sa = 0.0, sb = 0.0, sc = 0.0;
  for (int w = 0; w < kPcgThreads / 32; ++w) { sa += red3[3 * w]; sb += red3[3 * w + 1]; sc += red3[3 * w + 2]; }
  __syncthreads();
  a = sa; b = sb; c = sc;
}

// Sum of the G per-CTA partials of K quantities laid out as part[k * G + g]: warp 0 strides over g,
// butterfly-reduces, and publishes through shared memory.  Fixed shape -> deterministic, and only one
// warp per CTA touches L2.
template <int K>
__device__ __forceinline__ void gather_sums(const double* part, int G, double* sh /* [K] */, double (&out)[K]) {
  if (threadIdx.x < 32) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      double s = 0.0;
      for (int g = threadIdx.x; g < G; g += 32) s += __ldcg(part + k * G + g);
      s = warp_sum(s);
      if (threadIdx.x == 0) sh[k] = s;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; ++k) out[k] = sh[k];
  __syncthreads();
}

// Local rows of S * (vector in u_full).  A warp takes up to four rows (wid, wid + 8, ...) TOGETHER, so that an
// element of the vector is read from shared memory once for all of them: the product moves 16 bytes of shared
// memory per FMA otherwise, and at 18 rows x 1194 columns per CTA that alone is 2700 cycles of the one
// shared-memory pipe (measured 2.8 us per product, a third of a PCG iteration at C4).
__device__ __forceinline__ void pcg_local_matvec(const double* __restrict__ Srows, int n, int R, const double* u_full,
                                               double* y, int wid, int lane) {
  for (int r0 = wid; r0 < R; r0 += 4 * (kPcgThreads / 32)) {
    const double* row0p = Srows + (size_t)r0 * n;
    const size_t rs = (size_t)(kPcgThreads / 32) * n;           // stride between the warp's rows
    constexpr int W = kPcgThreads / 32;
    const bool h1 = r0 + W < R, h2 = r0 + 2 * W < R, h3 = r0 + 3 * W < R;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
    // n is a multiple of 6 and the slab starts on a 16-byte boundary: every row does, so a lane reads two
    // neighbouring columns with one 128-bit load
    const bool vec = ((reinterpret_cast<uintptr_t>(row0p) | reinterpret_cast<uintptr_t>(u_full)) & 15) == 0 && (n & 1) == 0;
    if (vec) {
      for (int j = 2 * lane; j < n; j += 64) {
        const double2 u = *reinterpret_cast<const double2*>(u_full + j);
        const double2 s0 = *reinterpret_cast<const double2*>(row0p + j);
        a0 = fma(s0.x, u.x, a0); b0 = fma(s0.y, u.y, b0);
        if (h1) { const double2 t = *reinterpret_cast<const double2*>(row0p + rs + j); a1 = fma(t.x, u.x, a1); b1 = fma(t.y, u.y, b1); }
        if (h2) { const double2 t = *reinterpret_cast<const double2*>(row0p + 2 * rs + j); a2 = fma(t.x, u.x, a2); b2 = fma(t.y, u.y, b2); }
        if (h3) { const double2 t = *reinterpret_cast<const double2*>(row0p + 3 * rs + j); a3 = fma(t.x, u.x, a3); b3 = fma(t.y, u.y, b3); }
      }
    } else {
      int j = lane;
      for (; j + 32 < n; j += 64) {
        const double u0 = u_full[j], u1 = u_full[j + 32];
        a0 = fma(row0p[j], u0, a0); b0 = fma(row0p[j + 32], u1, b0);
        if (h1) { a1 = fma(row0p[rs + j], u0, a1); b1 = fma(row0p[rs + j + 32], u1, b1); }
        if (h2) { a2 = fma(row0p[2 * rs + j], u0, a2); b2 = fma(row0p[2 * rs + j + 32], u1, b2); }
        if (h3) { a3 = fma(row0p[3 * rs + j], u0, a3); b3 = fma(row0p[3 * rs + j + 32], u1, b3); }
      }
      for (; j < n; j += 32) {
        const double u0 = u_full[j];
        a0 = fma(row0p[j], u0, a0);
        if (h1) a1 = fma(row0p[rs + j], u0, a1);
        if (h2) a2 = fma(row0p[2 * rs + j], u0, a2);
        if (h3) a3 = fma(row0p[3 * rs + j], u0, a3);
      }
    }
    a0 += b0; a1 += b1; a2 += b2; a3 += b3;
    warp_sum4(a0, a1, a2, a3);
    if (lane == 0) {
      y[r0] = a0;
      if (h1) y[r0 + W] = a1;
      if (h2) y[r0 + 2 * W] = a2;
      if (h3) y[r0 + 3 * W] = a3;
    }
  }
}

// Chronopoulos-Gear preconditioned CG: one matrix-vector product and ONE fused reduction per
// iteration, i.e. two grid barriers (publish u = M^-1 r; reduce (r,u), (Au,u), (r,r)).
__global__ void __launch_bounds__(kPcgThreads) ba_pcg_kernel(PcgArgs A) {
  extern __shared__ __align__(16) double sm[];
  pdl_wait();                                   // (no trigger: the cooperative grid keeps the SMs to itself)
  if (A.done != nullptr && *A.done) return;   // uniform over the grid: nobody writes it meanwhile
  const int n = A.n, G = A.G, g = blockIdx.x, tid = threadIdx.x;
  const int row0 = g * A.rows;
  const int R = max(0, min(A.rows, n - row0));
  double* u_full = sm;                 // [n]
  double* xl = u_full + n;             // [rows] each
  double* rl = xl + A.rows;
  double* ul = rl + A.rows;
  double* pl = ul + A.rows;
  double* sl = pl + A.rows;            // s = A p (recurrence)
  double* wl = sl + A.rows;            // w = S u, local rows
  double* minv = wl + A.rows;          // [rows/6][36]
  double* red = minv + 6 * A.rows;     // [8]
  double* slab = red + 8;              // [rows][n] when slab_in_smem
  // pipelined variant only (G > 1; the launch adds 4 * rows doubles behind the slab): m = M^-1 w, z = S q and
  // q = M^-1 s (recurrences), n = S m
  double* ml = slab + (A.slab_in_smem ? (size_t)A.rows * n : 0);
  double* zl = ml + A.rows;
  double* ql = zl + A.rows;
  double* nl = ql + A.rows;
  __shared__ int s_fail;
  __shared__ double s_sum[3];
  __shared__ double red3[3 * (kPcgThreads / 32)];
  unsigned target = 0;

  if (tid == 0) s_fail = 0;
  if (A.slab_in_smem) {
    // the slab arrives by 16-byte asynchronous copies, all of a thread's pieces in flight at once (a load -> store
    // loop keeps only a few requests per thread outstanding: ~15 us for the 172 KB of a C4 slab, per solve)
    const double* src = A.S + (size_t)row0 * n;
    const int cnt = R * n;
    if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(slab)) & 15) == 0 && (cnt & 1) == 0) {
      for (int i = tid; i < cnt / 2; i += kPcgThreads)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(slab + 2 * i)), "l"(src + 2 * i) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    } else {
      for (int i = tid; i < cnt; i += kPcgThreads) slab[i] = src[i];
    }
  }
  __syncthreads();
  const double* Srows = A.slab_in_smem ? slab : A.S + (size_t)row0 * n;
  // block-Jacobi preconditioner: invert the local diagonal 6x6 blocks
  for (int bi = tid; bi < R / 6; bi += kPcgThreads) {
    double blk[36];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) blk[6 * i + j] = Srows[(size_t)(6 * bi + i) * n + row0 + 6 * bi + j];
    }
    if (!spd6_inverse(blk)) s_fail = 1;
#pragma unroll
    for (int i = 0; i < 36; ++i) minv[36 * bi + i] = blk[i];
  }
  for (int i = tid; i < R; i += kPcgThreads) { xl[i] = 0.0; rl[i] = A.b[row0 + i]; pl[i] = 0.0; sl[i] = 0.0; }
  __syncthreads();
  if (tid == 0 && s_fail) atomicExch(A.fail_flag, 1);
  const int lane = tid & 31, wid = tid >> 5;
  int status = 2, it = 0;
  double gamma_old = 0.0, alpha_old = 0.0, thresh = 0.0;
  // local block-Jacobi application y = M^-1 v and the local rows of S * (vector in u_full)
  auto apply_minv = [&](const double* v, double* y) {
    for (int i = tid; i < R; i += kPcgThreads) {
      const int bi = i / 6, ii = i % 6;
      double z = 0.0;
      for (int j = 0; j < 6; ++j) z += minv[36 * bi + 6 * ii + j] * v[6 * bi + j];
      y[i] = z;
    }
  };
  auto local_matvec = [&](double* y) { pcg_local_matvec(Srows, n, R, u_full, y, wid, lane); };
  if (G > 1 && A.pipelined) {
    // ---- Pipelined PCG (Ghysels & Vanroose): the matrix-vector product of an iteration (n = S m, m = M^-1 w) does
    // not depend on that iteration's dot products, so the slab of m and the partial sums travel through ONE grid
    // barrier per iteration instead of two.  Same iterates as the two-barrier form in exact arithmetic (both are
    // preconditioned CG), three more recurrences (z = S q, q = M^-1 s, w = S u).  Measured at C4 (67 CTAs, 18 rows
    // of S each; %globaltimer stamps): an iteration is 0.9 us of local work before the barrier, 1.2 us barrier,
    // 1.4 us gather + scalars, 1.6 us product = 5.3 us, against 7.8 us with two barriers, a single-row product
    // (2.8 us: 16 bytes of shared memory per FMA) and three dependent partial-sum loads.
    apply_minv(rl, ul);
    // (the initial exchange uses buffer 1: iteration 0 publishes into buffer 0 while a slower CTA may still read this one)
    for (int i = tid; i < R; i += kPcgThreads) { A.p_glob[n + row0 + i] = ul[i]; zl[i] = 0.0; ql[i] = 0.0; }
    grid_barrier(A.bar, target, G);
    if (__ldcg(A.fail_flag)) {
      status = 1;
    } else {
      for (int i = tid; i < n; i += kPcgThreads) u_full[i] = __ldcg(A.p_glob + n + i);
      __syncthreads();
      local_matvec(wl);                                  // w = S u
      __syncthreads();
      double inv_gamma_old = 0.0, pap_old = 0.0;
      // With at most 32 local rows (the usual split: 18 at C4) the short serial part of an iteration -- recurrences,
      // m = M^-1 w, the three dot products, publication -- runs in warp 0 alone under __syncwarp, and the other warps
      // go straight to the grid barrier: four block-wide barriers and a shared-memory reduction fewer per iteration.
      const bool warp_tail = R <= 32;
      for (it = 0; it <= A.max_iter; ++it) {
        // the exchange buffers alternate with the iteration's parity: a CTA that has left the barrier may publish
        // the NEXT iteration's values while a slower one still reads this iteration's
        double* xch = A.p_glob + (size_t)(it & 1) * n;
        double* pch = A.part + (size_t)(it & 1) * 3 * G;
        if (warp_tail) {
          // (two warps, as in the cluster kernel: the dot products do not wait for M^-1 w and its publication)
          if (wid == 0) {
            asm volatile("bar.sync 1, 64;" ::: "memory");
            if (lane < R) {
              const int bi = lane / 6, ii = lane % 6;
              double z = 0.0;
#pragma unroll
              for (int j = 0; j < 6; ++j) z += minv[36 * bi + 6 * ii + j] * wl[6 * bi + j];
              ml[lane] = z;
              xch[row0 + lane] = z;
            }
          } else if (wid == 1) {
            asm volatile("bar.sync 1, 64;" ::: "memory");
            const double rv = lane < R ? rl[lane] : 0.0, uv = lane < R ? ul[lane] : 0.0, wv = lane < R ? wl[lane] : 0.0;
            double gam_p = rv * uv, del_p = wv * uv, rr_p = rv * rv;
            warp_sum3(gam_p, del_p, rr_p);
            if (lane == 0) { pch[g] = gam_p; pch[G + g] = del_p; pch[2 * G + g] = rr_p; }
          }
        } else {
          apply_minv(wl, ml);                            // m = M^-1 w
          double gam_p = 0.0, del_p = 0.0, rr_p = 0.0;
          for (int i = tid; i < R; i += kPcgThreads) {
            gam_p += rl[i] * ul[i];
            del_p += wl[i] * ul[i];
            rr_p += rl[i] * rl[i];
          }
          block_sum3(gam_p, del_p, rr_p, red3);          // (also orders ml before its publication below)
          for (int i = tid; i < R; i += kPcgThreads) xch[row0 + i] = ml[i];
          if (tid == 0) { pch[g] = gam_p; pch[G + g] = del_p; pch[2 * G + g] = rr_p; }
        }
        grid_barrier(A.bar, target, G);
        for (int i = tid; i < n; i += kPcgThreads) u_full[i] = __ldcg(xch + i);
        double sums[3];
        if (tid < 32) {                                    // all nine loads of a lane in flight, then the three sums
          double v[3][5];
#pragma unroll
          for (int k = 0; k < 3; ++k) {
#pragma unroll
            for (int q = 0; q < 5; ++q) {
              const int gg = tid + 32 * q;
              v[k][q] = gg < G ? __ldcg(pch + k * G + gg) : 0.0;
            }
          }
          if (G <= 160) {                                  // (one CTA per SM at most: always, on 148 SMs)
            double sk[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
              sk[k] = 0.0;
#pragma unroll
              for (int q = 0; q < 5; ++q)
                if (tid + 32 * q < G) sk[k] += v[k][q];
            }
            warp_sum3(sk[0], sk[1], sk[2]);                // (the three butterflies stage by stage, not one after the other)
            if (tid == 0) { s_sum[0] = sk[0]; s_sum[1] = sk[1]; s_sum[2] = sk[2]; }
          } else {                                         // (more CTAs than three per lane: the general form)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
              double sk = 0.0;
              for (int gg = tid; gg < G; gg += 32) sk += __ldcg(pch + k * G + gg);
              sk = warp_sum(sk);
              if (tid == 0) s_sum[k] = sk;
            }
          }
        }
        __syncthreads();                                   // (u_full is complete after it, too)
#pragma unroll
        for (int k = 0; k < 3; ++k) sums[k] = s_sum[k];    // (rewritten only after the next grid barrier)
        const double gamma = sums[0], delta = sums[1], rr = sums[2];
        if (it == 0) {
          if (!(rr > 0.0)) { status = 0; break; }        // b == 0: x = 0
          thresh = A.tol * A.tol * rr;
        }
        if (rr <= thresh && it > 0) { status = 0; break; }
        if (it == A.max_iter) break;
        // reciprocals instead of divisions on the scalar path that every thread walks between the barrier and the
        // product (p^T S p = delta - beta^2 * previous p^T S p, as in the single-CTA kernel)
        double beta = 0.0, pap = delta;
        if (it > 0) {
          beta = gamma * inv_gamma_old;
          pap = delta - beta * beta * pap_old;
        }
        if (!(pap > 1e-290) || !(pap < 1e290) || !(gamma > 1e-290) || !(gamma < 1e290)) { status = 1; break; }
        const double alpha = gamma * fast_rcp(pap);
        inv_gamma_old = fast_rcp(gamma); pap_old = pap;
        local_matvec(nl);                                // n = S m (u_full holds m)
        __syncthreads();
        if (!warp_tail || wid == 0) {
          for (int i = tid; i < R; i += kPcgThreads) {
            // (loads ahead of the stores: the arrays may alias as far as the compiler knows)
            const double nv = nl[i], zo = zl[i], mo = ml[i], qo = ql[i], wo = wl[i], so = sl[i], uo = ul[i], po = pl[i],
                         xo = xl[i], ro = rl[i];
            const double z = nv + beta * zo;
            const double q = mo + beta * qo;
            const double sv = wo + beta * so;
            const double pv = uo + beta * po;
            zl[i] = z; ql[i] = q; sl[i] = sv; pl[i] = pv;
            xl[i] = xo + alpha * pv;
            rl[i] = ro - alpha * sv;
            ul[i] = uo - alpha * q;
            wl[i] = wo - alpha * z;
          }
        }
        if (warp_tail) __syncwarp(); else __syncthreads();
      }
    }
  } else
  for (it = 0; it <= A.max_iter; ++it) {
    // u = M^-1 r (block local), published for the other CTAs
    for (int i = tid; i < R; i += kPcgThreads) {
      const int bi = i / 6, ii = i % 6;
      double z = 0.0;
      for (int j = 0; j < 6; ++j) z += minv[36 * bi + 6 * ii + j] * rl[6 * bi + j];
      ul[i] = z;
      if (G > 1) A.p_glob[row0 + i] = z; else u_full[i] = z;
    }
    if (G > 1) {
      grid_barrier(A.bar, target, G);
      if (it == 0 && __ldcg(A.fail_flag)) { status = 1; break; }
      for (int i = tid; i < n; i += kPcgThreads) u_full[i] = __ldcg(A.p_glob + i);
    } else if (it == 0 && s_fail) { status = 1; break; }
    __syncthreads();
    // w = S u for the local rows: one warp per row
    double gam_p = 0.0, del_p = 0.0, rr_p = 0.0;
    for (int r = wid; r < R; r += kPcgThreads / 32) {
      const double* row = Srows + (size_t)r * n;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;   // (see local_matvec)
      int j = lane;
      for (; j + 96 < n; j += 128) {
        a0 = fma(row[j], u_full[j], a0); a1 = fma(row[j + 32], u_full[j + 32], a1);
        a2 = fma(row[j + 64], u_full[j + 64], a2); a3 = fma(row[j + 96], u_full[j + 96], a3);
      }
      for (; j < n; j += 32) a0 = fma(row[j], u_full[j], a0);
      const double s = warp_sum((a0 + a1) + (a2 + a3));
      if (lane == 0) wl[r] = s;
    }
    __syncthreads();
    for (int i = tid; i < R; i += kPcgThreads) {
      const double w = wl[i];
      gam_p += rl[i] * ul[i];
      del_p += w * ul[i];
      rr_p += rl[i] * rl[i];
    }
    block_sum3(gam_p, del_p, rr_p, red3);
    double sums[3] = {gam_p, del_p, rr_p};
    if (G > 1) {
      if (tid == 0) { A.part[g] = gam_p; A.part[G + g] = del_p; A.part[2 * G + g] = rr_p; }
      grid_barrier(A.bar, target, G);
      gather_sums<3>(A.part, G, s_sum, sums);
    }
    const double gamma = sums[0], delta = sums[1], rr = sums[2];
    if (it == 0) {
      if (!(rr > 0.0)) { status = 0; break; }      // b == 0: x = 0
      thresh = A.tol * A.tol * rr;
    }
    if (rr <= thresh && it > 0) { status = 0; break; }
    if (it == A.max_iter) break;
    double beta = 0.0, pap = delta;
    if (it > 0) {
      beta = gamma / gamma_old;
      pap = delta - beta * gamma / alpha_old;        // = p^T S p in exact arithmetic
    }
    if (!(pap > 0.0) || !isfinite(pap) || !(gamma > 0.0)) { status = 1; break; }
    const double alpha = gamma / pap;
    for (int i = tid; i < R; i += kPcgThreads) {
      const double p = ul[i] + beta * pl[i];
      const double s = wl[i] + beta * sl[i];
      pl[i] = p; sl[i] = s;
      xl[i] += alpha * p;
      rl[i] -= alpha * s;
    }
    __syncthreads();
    gamma_old = gamma; alpha_old = alpha;
  }
  for (int i = tid; i < R; i += kPcgThreads) A.x[row0 + i] = (status == 1) ? 0.0 : xl[i];
  if (g == 0 && tid == 0) { *A.status = status; *A.iters = it; }
}

// Small systems (n <= kPcgSmallMax, e.g. the 9 free cameras of a local window): the same
// Chronopoulos-Gear iteration in ONE CTA of four warps with S in shared memory.  EVERY warp keeps the
// vectors (two rows per lane, in registers) and runs the scalar recurrences redundantly -- identical
// instructions on identical data, so all warps agree bit for bit and nobody waits for an owner; the
// matrix-vector product is split by COLUMN range over the four warps (S is symmetric, so a lane reads
// consecutive addresses) and the four partials are added in warp order -> deterministic.  ONE block
// barrier per iteration (partials double-buffered), reciprocals instead of divisions on the critical
// path (p^T S p = delta - beta^2 * previous p^T S p).  Setup: S arrives by cp.async while nine lanes
// invert the diagonal blocks straight from global memory.
constexpr int kPcgSmallMax = 64;
constexpr int kPcgSmallThreads = 128;
constexpr int kPcgSmallWarps = kPcgSmallThreads / 32;

__global__ void __launch_bounds__(kPcgSmallThreads) ba_pcg_small_kernel(PcgArgs A) {
  extern __shared__ __align__(16) double sm[];
  pdl_wait();
  pdl_trigger();
  if (A.done != nullptr && *A.done) return;
  const int n = A.n, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double* minv = sm + n * n;                        // [n/6][36]  (the first n * n doubles of the launch's shared memory are unused)
  double* rs = minv + 6 * n + wid * 64;             // [warps][64] each warp's own copy of r
  double* us = minv + 6 * n + (kPcgSmallWarps + wid) * 64;   // [warps][64] ... of u
  double* part = minv + 6 * n + 2 * kPcgSmallWarps * 64;     // [2][warps][64]
  // this lane's entries of S for the warp's column range go straight from global memory to registers, where they
  // stay for the whole solve (two rows x <= 16 columns; S is symmetric, so a warp reads consecutive addresses);
  // the loads are in flight while nine lanes invert the diagonal blocks.  (Reading S from shared memory in every
  // iteration cost 12.4 us per solve at n = 54 against 7.8 us from registers.)
  const int i0 = lane, i1 = lane + 32;
  const bool h0 = i0 < n, h1 = i1 < n;
  const int cols = ((n + 7) / 8) * 2;               // column range of this warp, <= 16 (multiples of 2)
  const int j0 = min(n, wid * cols), j1 = min(n, j0 + cols);
  double sr0[16], sr1[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    const int j = j0 + q;
    sr0[q] = (j < j1 && h0) ? __ldcg(A.S + (size_t)j * n + i0) : 0.0;
    sr1[q] = (j < j1 && h1) ? __ldcg(A.S + (size_t)j * n + i1) : 0.0;
  }
  int fail = 0;
  if (threadIdx.x < n / 6) {                        // n <= 64: at most ten blocks, all in warp 0
    const int bi = threadIdx.x;
    double blk[36];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) blk[6 * i + j] = __ldcg(A.S + (size_t)(6 * bi + i) * n + 6 * bi + j);
    }
    if (!spd6_inverse(blk)) fail = 1;
#pragma unroll
    for (int i = 0; i < 36; ++i) minv[36 * bi + i] = blk[i];
  }
  rs[lane] = 0.0; rs[lane + 32] = 0.0; us[lane] = 0.0; us[lane + 32] = 0.0;
  fail = __syncthreads_or(fail);
  // preconditioner rows of this lane, in registers
  double m0[6], m1[6];
  const int b0 = h0 ? i0 / 6 : 0, b1 = h1 ? i1 / 6 : 0;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    m0[k] = h0 ? minv[36 * b0 + 6 * (i0 - 6 * b0) + k] : 0.0;
    m1[k] = h1 ? minv[36 * b1 + 6 * (i1 - 6 * b1) + k] : 0.0;
  }
  double x0 = 0.0, x1 = 0.0, p0 = 0.0, p1 = 0.0, s0 = 0.0, s1 = 0.0;
  double r0 = h0 ? A.b[i0] : 0.0, r1 = h1 ? A.b[i1] : 0.0;
  int status = fail ? 1 : 2, it = 0, buf = 0;
  double inv_gamma_old = 0.0, pap_old = 0.0, thresh = 0.0;
  if (!fail) {
    for (it = 0;; ++it) {
      // u = M^-1 r through the warp's own shared-memory copy of r
      rs[i0] = r0; rs[i1] = r1;
      __syncwarp();
      const double* ra = rs + 6 * b0;
      const double* rb = rs + 6 * b1;
      const double u0 = (m0[0] * ra[0] + m0[1] * ra[1]) + (m0[2] * ra[2] + m0[3] * ra[3]) + (m0[4] * ra[4] + m0[5] * ra[5]);
      const double u1 = (m1[0] * rb[0] + m1[1] * rb[1]) + (m1[2] * rb[2] + m1[3] * rb[3]) + (m1[4] * rb[4] + m1[5] * rb[5]);
      us[i0] = u0; us[i1] = u1;
      __syncwarp();
      {   // partial w = S[:, j0:j1] u[j0:j1] for rows i0, i1
        double wa = 0.0, wb = 0.0, wc = 0.0, wd = 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) {   // (entries beyond the column range are zero, u beyond n is zero)
          const int j = min(j0 + 2 * q, 62);
          const double ua = us[j], ub = us[j + 1];
          wa = fma(sr0[2 * q], ua, wa); wb = fma(sr0[2 * q + 1], ub, wb);
          wc = fma(sr1[2 * q], ua, wc); wd = fma(sr1[2 * q + 1], ub, wd);
        }
        double* pw = part + (buf * kPcgSmallWarps + wid) * 64;
        pw[i0] = h0 ? wa + wb : 0.0;
        pw[i1] = h1 ? wc + wd : 0.0;
      }
      __syncthreads();
      const double* pr = part + buf * kPcgSmallWarps * 64;
      buf ^= 1;
      const double w0 = (pr[i0] + pr[64 + i0]) + (pr[128 + i0] + pr[192 + i0]);
      const double w1 = (pr[i1] + pr[64 + i1]) + (pr[128 + i1] + pr[192 + i1]);
      double gamma = r0 * u0 + r1 * u1, delta = w0 * u0 + w1 * u1, rr = r0 * r0 + r1 * r1;
      warp_sum3(gamma, delta, rr);
      if (it == 0) {
        if (!(rr > 0.0)) { status = 0; break; }      // b == 0: x = 0
        thresh = A.tol * A.tol * rr;
      }
      if (rr <= thresh && it > 0) { status = 0; break; }
      if (it == A.max_iter) break;
      double beta = 0.0, pap = delta;
      if (it > 0) {
        beta = gamma * inv_gamma_old;
        pap = delta - beta * beta * pap_old;           // = p^T S p in exact arithmetic
      }
      if (!(pap > 1e-290) || !(pap < 1e290) || !(gamma > 1e-290) || !(gamma < 1e290)) { status = 1; break; }
      const double alpha = gamma * fast_rcp(pap);
      inv_gamma_old = fast_rcp(gamma); pap_old = pap;
      p0 = u0 + beta * p0; p1 = u1 + beta * p1;
      s0 = w0 + beta * s0; s1 = w1 + beta * s1;
      x0 += alpha * p0; x1 += alpha * p1;
      r0 -= alpha * s0; r1 -= alpha * s1;
    }
  }
  if (wid == 0) {
    if (h0) A.x[i0] = (status == 1) ? 0.0 : x0;
    if (h1) A.x[i1] = (status == 1) ? 0.0 : x1;
    if (lane == 0) { *A.status = status; *A.iters = it; }
  }
}

// The same single-CTA layout running the PIPELINED recurrences (Ghysels-Vanroose, as in ba_pcg_kernel): the three dot
// products of an iteration do not depend on its matrix-vector product n = S (M^-1 w), so the five butterfly stages of
// their warp reduction -- a shuffle round trip each, ~350 cycles of an iteration's ~760 -- are written BETWEEN the
// phases of the product (shuffles issued, a phase of the product, then the additions): shuffles and barriers keep
// their program order, so the overlap has to be in the source.
__global__ void __launch_bounds__(kPcgSmallThreads) ba_pcg_small_pipelined_kernel(PcgArgs A) {
  extern __shared__ __align__(16) double sm[];
  pdl_wait();
  pdl_trigger();
  if (A.done != nullptr && *A.done) return;
  const int n = A.n, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double* minv = sm + n * n;                        // (same shared-memory layout as ba_pcg_small_kernel)
  double* rs = minv + 6 * n + wid * 64;             // [warps][64] each warp's own copy of w
  double* us = minv + 6 * n + (kPcgSmallWarps + wid) * 64;   // [warps][64] ... of m
  double* part = minv + 6 * n + 2 * kPcgSmallWarps * 64;     // [2][warps][64]
  const int i0 = lane, i1 = lane + 32;
  const bool h0 = i0 < n, h1 = i1 < n;
  const int cols = ((n + 7) / 8) * 2;
  const int j0 = min(n, wid * cols), j1 = min(n, j0 + cols);
  double sr0[16], sr1[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    const int j = j0 + q;
    sr0[q] = (j < j1 && h0) ? __ldcg(A.S + (size_t)j * n + i0) : 0.0;
    sr1[q] = (j < j1 && h1) ? __ldcg(A.S + (size_t)j * n + i1) : 0.0;
  }
  int fail = 0;
  if (threadIdx.x < n / 6) {
    const int bi = threadIdx.x;
    double blk[36];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) blk[6 * i + j] = __ldcg(A.S + (size_t)(6 * bi + i) * n + 6 * bi + j);
    }
    if (!spd6_inverse(blk)) fail = 1;
#pragma unroll
    for (int i = 0; i < 36; ++i) minv[36 * bi + i] = blk[i];
  }
  rs[lane] = 0.0; rs[lane + 32] = 0.0; us[lane] = 0.0; us[lane + 32] = 0.0;
  fail = __syncthreads_or(fail);
  double m0[6], m1[6];
  const int b0 = h0 ? i0 / 6 : 0, b1 = h1 ? i1 / 6 : 0;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    m0[k] = h0 ? minv[36 * b0 + 6 * (i0 - 6 * b0) + k] : 0.0;
    m1[k] = h1 ? minv[36 * b1 + 6 * (i1 - 6 * b1) + k] : 0.0;
  }
  const double* ra = rs + 6 * b0;
  const double* rb = rs + 6 * b1;
  int buf = 0;
  // y = M^-1 v (rows of this lane) through the warp's own copy of v; then this warp's part of S y into part[buf]
  auto apply_minv = [&](double v0, double v1, double& y0, double& y1) {
    rs[i0] = v0; rs[i1] = v1;
    __syncwarp();
    y0 = (m0[0] * ra[0] + m0[1] * ra[1]) + (m0[2] * ra[2] + m0[3] * ra[3]) + (m0[4] * ra[4] + m0[5] * ra[5]);
    y1 = (m1[0] * rb[0] + m1[1] * rb[1]) + (m1[2] * rb[2] + m1[3] * rb[3]) + (m1[4] * rb[4] + m1[5] * rb[5]);
  };
  auto publish = [&](double y0, double y1) {
    us[i0] = y0; us[i1] = y1;
    __syncwarp();
  };
  auto product_half = [&](int qa, int qb, double& wa, double& wb, double& wc, double& wd) {
#pragma unroll
    for (int q = qa; q < qb; ++q) {                 // (entries beyond the column range are zero, the vector beyond n is zero)
      const int j = min(j0 + 2 * q, 62);
      const double ua = us[j], ub = us[j + 1];
      wa = fma(sr0[2 * q], ua, wa); wb = fma(sr0[2 * q + 1], ub, wb);
      wc = fma(sr1[2 * q], ua, wc); wd = fma(sr1[2 * q + 1], ub, wd);
    }
  };
  auto store_partial = [&](double wa, double wb, double wc, double wd) {
    double* pw = part + (buf * kPcgSmallWarps + wid) * 64;
    pw[i0] = h0 ? wa + wb : 0.0;
    pw[i1] = h1 ? wc + wd : 0.0;
  };
  auto summed = [&](double& y0, double& y1) {        // (after the block barrier) the four partials, in warp order
    const double* pr = part + buf * kPcgSmallWarps * 64;
    buf ^= 1;
    y0 = (pr[i0] + pr[64 + i0]) + (pr[128 + i0] + pr[192 + i0]);
    y1 = (pr[i1] + pr[64 + i1]) + (pr[128 + i1] + pr[192 + i1]);
  };
  double x0 = 0.0, x1 = 0.0, p0 = 0.0, p1 = 0.0, s0 = 0.0, s1 = 0.0, q0 = 0.0, q1 = 0.0, z0 = 0.0, z1 = 0.0;
  double r0 = h0 ? A.b[i0] : 0.0, r1 = h1 ? A.b[i1] : 0.0;
  double u0 = 0.0, u1 = 0.0, w0 = 0.0, w1 = 0.0;
  int status = fail ? 1 : 2, it = 0;
  double inv_gamma_old = 0.0, pap_old = 0.0, thresh = 0.0;
  if (!fail) {
    apply_minv(r0, r1, u0, u1);                      // u = M^-1 r
    __syncwarp();
    publish(u0, u1);
    {
      double wa = 0.0, wb = 0.0, wc = 0.0, wd = 0.0;
      product_half(0, 8, wa, wb, wc, wd);
      store_partial(wa, wb, wc, wd);
    }
    __syncthreads();
    summed(w0, w1);                                  // w = S u
    for (it = 0;; ++it) {
      double gamma = r0 * u0 + r1 * u1, delta = w0 * u0 + w1 * u1, rr = r0 * r0 + r1 * r1;
      double ta, tb, tc, mm0, mm1, wa = 0.0, wb = 0.0, wc = 0.0, wd = 0.0;
#define VSLAM_BFLY_ISSUE(o) ta = __shfl_xor_sync(0xffffffffu, gamma, o); tb = __shfl_xor_sync(0xffffffffu, delta, o); \
                            tc = __shfl_xor_sync(0xffffffffu, rr, o);
#define VSLAM_BFLY_ADD() gamma += ta; delta += tb; rr += tc;
      __syncwarp();                                  // (the previous iteration's reads of rs / us are done)
      VSLAM_BFLY_ISSUE(16)
      apply_minv(w0, w1, mm0, mm1);                  // m = M^-1 w
      VSLAM_BFLY_ADD()
      VSLAM_BFLY_ISSUE(8)
      publish(mm0, mm1);
      VSLAM_BFLY_ADD()
      VSLAM_BFLY_ISSUE(4)
      product_half(0, 4, wa, wb, wc, wd);
      VSLAM_BFLY_ADD()
      VSLAM_BFLY_ISSUE(2)
      product_half(4, 8, wa, wb, wc, wd);
      VSLAM_BFLY_ADD()
      VSLAM_BFLY_ISSUE(1)
      store_partial(wa, wb, wc, wd);
      VSLAM_BFLY_ADD()
#undef VSLAM_BFLY_ISSUE
#undef VSLAM_BFLY_ADD
      __syncthreads();
      double n0, n1;
      summed(n0, n1);                                // n = S m
      if (it == 0) {
        if (!(rr > 0.0)) { status = 0; break; }      // b == 0: x = 0
        thresh = A.tol * A.tol * rr;
      }
      if (rr <= thresh && it > 0) { status = 0; break; }
      if (it == A.max_iter) break;
      double beta = 0.0, pap = delta;
      if (it > 0) {
        beta = gamma * inv_gamma_old;
        pap = delta - beta * beta * pap_old;           // = p^T S p in exact arithmetic
      }
      if (!(pap > 1e-290) || !(pap < 1e290) || !(gamma > 1e-290) || !(gamma < 1e290)) { status = 1; break; }
      const double alpha = gamma * fast_rcp(pap);
      inv_gamma_old = fast_rcp(gamma); pap_old = pap;
      z0 = n0 + beta * z0; z1 = n1 + beta * z1;
      q0 = mm0 + beta * q0; q1 = mm1 + beta * q1;
      s0 = w0 + beta * s0; s1 = w1 + beta * s1;
      p0 = u0 + beta * p0; p1 = u1 + beta * p1;
      x0 += alpha * p0; x1 += alpha * p1;
      r0 -= alpha * s0; r1 -= alpha * s1;
      u0 -= alpha * q0; u1 -= alpha * q1;
      w0 -= alpha * z0; w1 -= alpha * z1;
    }
  }
  if (wid == 0) {
    if (h0) A.x[i0] = (status == 1) ? 0.0 : x0;
    if (h1) A.x[i1] = (status == 1) ? 0.0 : x1;
    if (lane == 0) { *A.status = status; *A.iters = it; }
  }
}

// Systems of 11 .. 21 free cameras (64 < n <= kPcgMidMax): the single-CTA kernel above with four rows per lane and
// eight warps, i.e. S (<= 128 x 128) still lives in registers for the whole solve -- 64 doubles per thread -- and an
// iteration still costs one block barrier.  The cooperative multi-CTA kernel pays a grid barrier and a gather through
// L2 per iteration (~3.2 us) whatever the size: at n = 66 that was 70 of the 120 us of an LM iteration.  The
// preconditioner rows are read from shared memory here (no registers left for them); the diagonal blocks are inverted
// BEFORE the loads of S are issued, because the inversion's live set and the 128 registers of S do not fit together.
constexpr int kPcgMidMax = 128;
constexpr int kPcgMidThreads = 256;
constexpr int kPcgMidWarps = kPcgMidThreads / 32;
constexpr int kPcgMidRows = 4;                         // rows per lane: lane, lane + 32, ...
constexpr int kPcgMidCols = 16;                        // columns per warp, at most
constexpr size_t kPcgMidSmemDoubles = 6 * kPcgMidMax + 4 * kPcgMidWarps * kPcgMidMax;   // M^-1, r / u copies, two sets of partials

__global__ void __launch_bounds__(kPcgMidThreads, 1) ba_pcg_mid_kernel(PcgArgs A) {
  extern __shared__ __align__(16) double sm[];
  pdl_wait();
  pdl_trigger();
  if (A.done != nullptr && *A.done) return;
  constexpr int NR = kPcgMidRows, W = kPcgMidWarps, NM = kPcgMidMax;
  const int n = A.n, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double* minv = sm;                                // [n/6][36]
  double* rs = sm + 6 * NM + wid * NM;              // [W][NM] each warp's own copy of r
  double* us = sm + 6 * NM + (W + wid) * NM;        // [W][NM] ... of u
  double* part = sm + 6 * NM + 2 * W * NM;          // [2][W][NM]
  int fail = 0;
  if (threadIdx.x < n / 6) {                        // n <= 128: at most 21 blocks, all in warp 0
    const int bi = threadIdx.x;
    double blk[36];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) blk[6 * i + j] = __ldcg(A.S + (size_t)(6 * bi + i) * n + 6 * bi + j);
    }
    if (!spd6_inverse(blk)) fail = 1;
#pragma unroll
    for (int i = 0; i < 36; ++i) minv[36 * bi + i] = blk[i];
  }
  // this lane's entries of S for the warp's column range: four rows x <= 16 columns (S is symmetric, so a warp reads
  // consecutive addresses of row j for its "column" j)
  int ri[NR], rb6[NR], rm[NR];
  bool h[NR];
#pragma unroll
  for (int k = 0; k < NR; ++k) {
    ri[k] = lane + 32 * k;
    h[k] = ri[k] < n;
    const int b = h[k] ? ri[k] / 6 : 0;
    rb6[k] = 6 * b;                                 // first row of the diagonal block
    rm[k] = h[k] ? 36 * b + 6 * (ri[k] - 6 * b) : 0;   // this row of M^-1
  }
  const int cols = ((n + 2 * W - 1) / (2 * W)) * 2;   // column range of this warp, <= 16 (multiples of 2)
  const int j0 = min(n, wid * cols), j1 = min(n, j0 + cols);
  double sr[NR][kPcgMidCols];
#pragma unroll
  for (int q = 0; q < kPcgMidCols; ++q) {
    const int j = j0 + q;
#pragma unroll
    for (int k = 0; k < NR; ++k) sr[k][q] = (j < j1 && h[k]) ? __ldcg(A.S + (size_t)j * n + ri[k]) : 0.0;
  }
#pragma unroll
  for (int k = 0; k < NR; ++k) { rs[ri[k]] = 0.0; us[ri[k]] = 0.0; }
  fail = __syncthreads_or(fail);
  double x[NR], p[NR], sv[NR], r[NR];
#pragma unroll
  for (int k = 0; k < NR; ++k) { x[k] = 0.0; p[k] = 0.0; sv[k] = 0.0; r[k] = h[k] ? A.b[ri[k]] : 0.0; }
  int status = fail ? 1 : 2, it = 0, buf = 0;
  double inv_gamma_old = 0.0, pap_old = 0.0, thresh = 0.0;
  if (!fail) {
    for (it = 0;; ++it) {
      // u = M^-1 r through the warp's own shared-memory copy of r
#pragma unroll
      for (int k = 0; k < NR; ++k) rs[ri[k]] = r[k];
      __syncwarp();
      double u[NR];
#pragma unroll
      for (int k = 0; k < NR; ++k) {
        const double* ra = rs + rb6[k];
        const double* m = minv + rm[k];
        const double t = (m[0] * ra[0] + m[1] * ra[1]) + (m[2] * ra[2] + m[3] * ra[3]) + (m[4] * ra[4] + m[5] * ra[5]);
        u[k] = h[k] ? t : 0.0;
        us[ri[k]] = u[k];
      }
      __syncwarp();
      {   // partial w = S[:, j0:j1] u[j0:j1] for this lane's rows
        double wa[NR], wb[NR];
#pragma unroll
        for (int k = 0; k < NR; ++k) { wa[k] = 0.0; wb[k] = 0.0; }
#pragma unroll
        for (int q = 0; q < kPcgMidCols / 2; ++q) {   // (entries beyond the column range are zero, u beyond n is zero)
          const int j = min(j0 + 2 * q, NM - 2);
          const double ua = us[j], ub = us[j + 1];
#pragma unroll
          for (int k = 0; k < NR; ++k) { wa[k] = fma(sr[k][2 * q], ua, wa[k]); wb[k] = fma(sr[k][2 * q + 1], ub, wb[k]); }
        }
        double* pw = part + (buf * W + wid) * NM;
#pragma unroll
        for (int k = 0; k < NR; ++k) pw[ri[k]] = h[k] ? wa[k] + wb[k] : 0.0;
      }
      __syncthreads();
      const double* pr = part + buf * W * NM;
      buf ^= 1;
      double w[NR];
      double g_p = 0.0, d_p = 0.0, r_p = 0.0;
#pragma unroll
      for (int k = 0; k < NR; ++k) {
        const double* q = pr + ri[k];
        w[k] = ((q[0] + q[NM]) + (q[2 * NM] + q[3 * NM])) + ((q[4 * NM] + q[5 * NM]) + (q[6 * NM] + q[7 * NM]));
        g_p = fma(r[k], u[k], g_p); d_p = fma(w[k], u[k], d_p); r_p = fma(r[k], r[k], r_p);
      }
      double gamma = g_p, delta = d_p, rr = r_p;
      warp_sum3(gamma, delta, rr);
      if (it == 0) {
        if (!(rr > 0.0)) { status = 0; break; }      // b == 0: x = 0
        thresh = A.tol * A.tol * rr;
      }
      if (rr <= thresh && it > 0) { status = 0; break; }
      if (it == A.max_iter) break;
      double beta = 0.0, pap = delta;
      if (it > 0) {
        beta = gamma * inv_gamma_old;
        pap = delta - beta * beta * pap_old;           // = p^T S p in exact arithmetic
      }
      if (!(pap > 1e-290) || !(pap < 1e290) || !(gamma > 1e-290) || !(gamma < 1e290)) { status = 1; break; }
      const double alpha = gamma * fast_rcp(pap);
      inv_gamma_old = fast_rcp(gamma); pap_old = pap;
#pragma unroll
      for (int k = 0; k < NR; ++k) {
        p[k] = u[k] + beta * p[k];
        sv[k] = w[k] + beta * sv[k];
        x[k] += alpha * p[k];
        r[k] -= alpha * sv[k];
      }
    }
  }
  if (wid == 0) {
#pragma unroll
    for (int k = 0; k < NR; ++k)
      if (h[k]) A.x[ri[k]] = (status == 1) ? 0.0 : x[k];
    if (lane == 0) { *A.status = status; *A.iters = it; }
  }
}

// Systems of 22 .. 64 free cameras (kPcgMidMax < n <= kPcgClusterMaxN): the pipelined iteration of ba_pcg_kernel on ONE
// thread-block cluster of up to eight CTAs, with S in registers as in the single-CTA kernels.  A CTA owns a slab of
// <= 48 rows (two per lane); its eight warps split the COLUMNS (S is symmetric: a warp loads consecutive addresses), so
// a thread keeps 2 x <= 48 entries for the whole solve.  What the cooperative kernel exchanges through L2 behind a grid
// barrier -- the CTA's rows of m = M^-1 w and its three partial dot products -- is stored straight into every CTA's
// shared memory (DSMEM) ahead of one cluster barrier: an iteration has no global-memory round trip, nothing to gather,
// and every CTA adds the same partials in the same order (identical decisions, deterministic).  Warp 0 owns the
// vector recurrences of the slab; exchange buffers alternate with the iteration's parity (a CTA that has left the
// barrier may publish the next iteration's values while a slower one still reads this iteration's).
constexpr int kPcgClusterMaxCtas = 8;                  // <2, 48, 8>: two rows per lane, portable cluster size
constexpr int kPcgClusterMaxRows = 48;                 // rows of S per CTA
constexpr int kPcgClusterMaxCols = 48;                 // columns of S per warp
constexpr int kPcgClusterMaxN = kPcgClusterMaxCtas * kPcgClusterMaxRows;
constexpr int kPcgCluster16MaxCtas = 16;               // <1, 64, 16>: one row per lane, sixteen CTAs
constexpr int kPcgCluster16MaxRows = 30;
constexpr int kPcgCluster16MaxCols = 64;
constexpr int kPcgCluster16MaxN = kPcgCluster16MaxCtas * kPcgCluster16MaxRows;
constexpr int kPcgClusterWarps = kPcgThreads / 32;
__host__ __device__ inline size_t pcg_cluster_smem_doubles(int n, int rows, int max_ctas) {
  return 2 * (size_t)n + 7 * (size_t)max_ctas + 2 + 6 * kPcgClusterMaxRows + 10 * 64 + kPcgClusterWarps * 64 +
         (size_t)n * rows;   // (the last term: staging of the slab)
}

// DSMEM exchange of the cluster kernel: a value goes into a peer's shared memory by an asynchronous store that also
// counts its bytes on the peer's mbarrier (st.async ... mbarrier::complete_tx::bytes); the peer waits on its OWN barrier
// for the bytes of a phase.  No cluster-wide barrier per iteration: a cluster barrier cost ~600 cycles plus ~900 more
// before the first shared-memory read behind its acquire returned (measured with clock64 stamps).
__device__ __forceinline__ uint32_t cluster_map_u32(const void* local_smem, int rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(local_smem)), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_async_f64(uint32_t remote_addr, double v, uint32_t remote_bar) {
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr),
               "l"(__double_as_longlong(v)), "r"(remote_bar)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// (bounded: a protocol error must end as a reported breakdown, not as a hung device)
__device__ __forceinline__ bool mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
  for (int k = 0; k < (1 << 24); ++k)
    if (mbar_try_wait(bar, parity)) return true;
  return false;
}

// NRL rows per lane, CM columns per warp at most, MAXG CTAs per cluster at most: <2, 48, 8> for n <= 384 (portable cluster
// size), <1, 64, 16> for n <= 480 (sixteen CTAs of <= 30 rows: a non-portable cluster size, used where the device places it).
template <int NRL, int CM, int MAXG>
__global__ void __launch_bounds__(kPcgThreads, 1) ba_pcg_cluster_kernel(PcgArgs A) {
  extern __shared__ __align__(16) double sm[];
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  pdl_wait();
  pdl_trigger();
  if (A.done != nullptr && *A.done) return;          // (uniform over the cluster)
  constexpr int W = kPcgClusterWarps;
  const int n = A.n, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int g = (int)cluster.block_rank(), G = (int)cluster.num_blocks();
  const int row0 = g * A.rows, R = max(0, min(A.rows, n - row0));
  // identical layout in every CTA of the cluster (remote stores address it by offset)
  double* u_buf = sm;                                 // [2][n]   exchange: m of every CTA
  double* pch = sm + 2 * (size_t)n;                   // [2][3][MAXG] exchange: partial dot products
  double* failv = pch + 6 * MAXG;                     // [MAXG]
  uint64_t* bars = reinterpret_cast<uint64_t*>(failv + MAXG);   // [2] one mbarrier per exchange buffer
  double* minv = failv + MAXG + 2;                    // [<= 8][36]
  double* vec = minv + 6 * kPcgClusterMaxRows;        // [10][64]  the slab's vectors (warp 0)
  double* part = vec + 10 * 64;                       // [W][64]   partial products of the warps
  double *xl = vec, *rl = vec + 64, *pl = vec + 128, *sl = vec + 192, *ul = vec + 256, *wl = vec + 320, *ml = vec + 384,
         *zl = vec + 448, *ql = vec + 512;
  double* stage = part + W * 64;                      // [n][rows]  columns row0 .. of S, until they are in registers
  const int ia = lane, ib = lane + 32;                // this lane's rows of the slab
  const bool ha = ia < R, hb = NRL == 2 && ib < R;
  // The slab arrives by 16-byte asynchronous copies -- S is symmetric, so the CTA's rows are read as a column slab:
  // R consecutive doubles of every row j -- while warp 0 inverts the diagonal blocks.  (Loading the registers straight
  // from global memory, 96 eight-byte loads per thread, cost 2-4 us per solve.)
  {
    const int per_row = R / 2;                        // (row0, R and n are even: every piece is 16-byte aligned)
    const int cnt = n * per_row;
    for (int i = tid; i < cnt; i += kPcgThreads) {
      const int j = i / per_row, k = i - j * per_row;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage + (size_t)j * A.rows + 2 * k)),
                   "l"(A.S + (size_t)j * n + row0 + 2 * k) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // block-Jacobi preconditioner: the local diagonal 6x6 blocks, straight from global memory
  int fail = 0;
  if (tid < R / 6) {
    const int bi = tid;
    double blk[36];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) blk[6 * i + j] = __ldcg(A.S + (size_t)(row0 + 6 * bi + i) * n + row0 + 6 * bi + j);
    }
    if (!spd6_inverse(blk)) fail = 1;
#pragma unroll
    for (int i = 0; i < 36; ++i) minv[36 * bi + i] = blk[i];
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  // this lane's entries of S for the warp's column range (after the inversion: its live set and these do not fit in
  // the register file together)
  const int cols = ((n + 2 * W - 1) / (2 * W)) * 2;  // <= 48, even
  const int j0 = min(n, wid * cols), j1 = min(n, j0 + cols);
  double sa[CM], sb[NRL == 2 ? CM : 2];
#pragma unroll
  for (int q = 0; q < CM; ++q) {
    const int j = j0 + q;
    sa[q] = (j < j1 && ha) ? stage[(size_t)j * A.rows + ia] : 0.0;
    if constexpr (NRL == 2) sb[q] = (j < j1 && hb) ? stage[(size_t)j * A.rows + ib] : 0.0;
  }
  if (wid == 0) {
    xl[ia] = 0.0; xl[ib] = 0.0; pl[ia] = 0.0; pl[ib] = 0.0; sl[ia] = 0.0; sl[ib] = 0.0; zl[ia] = 0.0; zl[ib] = 0.0;
    ql[ia] = 0.0; ql[ib] = 0.0; wl[ia] = 0.0; wl[ib] = 0.0; ul[ia] = 0.0; ul[ib] = 0.0; ml[ia] = 0.0; ml[ib] = 0.0;
    rl[ia] = ha ? A.b[row0 + ia] : 0.0; rl[ib] = hb ? A.b[row0 + ib] : 0.0;
  }
  if (tid == 0) { mbar_init(bars, 1); mbar_init(bars + 1, 1); mbar_fence_init(); }
  fail = __syncthreads_or(fail);
  auto minv_row = [&](int i, const double* v) {       // (M^-1 v)_i for a local row i
    const int bi = i / 6, ii = i - 6 * bi;
    const double* m = minv + 36 * bi + 6 * ii;
    const double* q = v + 6 * bi;
    return (m[0] * q[0] + m[1] * q[1]) + (m[2] * q[2] + m[3] * q[3]) + (m[4] * q[4] + m[5] * q[5]);
  };
  // every CTA's window onto this shared memory layout (the mapping only adds a per-CTA base: one mapa per peer, once,
  // instead of two in front of every store)
  uint32_t rbase[MAXG];
#pragma unroll
  for (int k = 0; k < MAXG; ++k) rbase[k] = cluster_map_u32(sm, min(k, G - 1)) - smem_u32(sm);
  auto push_to = [&](int buf, const double* dst, double v) {   // a value into every CTA's copy of *dst, counted on its barrier
    const uint32_t a = smem_u32(dst), b = smem_u32(bars + buf);
#pragma unroll
    for (int k = 0; k < MAXG; ++k)
      if (k < G) st_async_f64(a + rbase[k], v, b + rbase[k]);
  };
  auto push = [&](int buf, int i, double v) { push_to(buf, u_buf + (size_t)buf * n + row0 + i, v); };
  bool lost = false;                                  // a wait ran out: reported as a breakdown
  // this warp's part of (S v)[slab rows], v = exchange buffer `buf`, into part[wid]
  auto partial_product = [&](int buf) {
    const double* v = u_buf + (size_t)buf * n;
    double wa0 = 0.0, wa1 = 0.0, wb0 = 0.0, wb1 = 0.0;
    // eight columns per (uniform) step, their four 128-bit loads issued together: with a branch around every pair of
    // columns each load's latency was exposed (80 cycles per pair, two warps per scheduler).  Entries of S beyond the
    // column range are zero and the index is clamped, so a partly used step multiplies finite values by zero.
#pragma unroll
    for (int c = 0; c < CM / 8; ++c) {
      if (8 * c < cols) {
        double2 uu[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) uu[k] = *reinterpret_cast<const double2*>(v + min(j0 + 8 * c + 2 * k, n - 2));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int q = 8 * c + 2 * k;
          wa0 = fma(sa[q], uu[k].x, wa0); wa1 = fma(sa[q + 1], uu[k].y, wa1);
          if constexpr (NRL == 2) { wb0 = fma(sb[q], uu[k].x, wb0); wb1 = fma(sb[q + 1], uu[k].y, wb1); }
        }
      }
    }
    part[wid * 64 + ia] = wa0 + wa1;
    if constexpr (NRL == 2) part[wid * 64 + ib] = wb0 + wb1;
  };
  auto summed_product = [&](int i) {                  // (after a block barrier) the eight partials of row i, in warp order
    const double* q = part + i;
    return ((q[0] + q[64]) + (q[128] + q[192])) + ((q[256] + q[320]) + (q[384] + q[448]));
  };
  cluster.sync();                                     // every CTA's barriers are initialised
  if (wid == 0) {
    if (lane == 0) mbar_expect_tx(bars + 1, (uint32_t)(8 * (n + G)));            // (the initial exchange uses buffer 1)
    if (ha) { const double u = minv_row(ia, rl); ul[ia] = u; push(1, ia, u); }
    if (hb) { const double u = minv_row(ib, rl); ul[ib] = u; push(1, ib, u); }
    if (lane == 0) push_to(1, failv + g, fail ? 1.0 : 0.0);
  }
  lost = !mbar_wait_bounded(bars + 1, 0);
  bool any_fail = lost;
  for (int k = 0; k < G; ++k) any_fail = any_fail || failv[k] != 0.0;
  int status = 2, it = 0;
  double thresh = 0.0;
  if (any_fail) {
    status = 1;
  } else {
    partial_product(1);                                // w = S u
    __syncthreads();
    if (wid == 0) { wl[ia] = ha ? summed_product(ia) : 0.0; wl[ib] = hb ? summed_product(ib) : 0.0; }
    __syncthreads();                                   // (part is rewritten after the next cluster barrier only; wl: warp 0)
    double inv_gamma_old = 0.0, pap_old = 0.0;
    for (it = 0; it <= A.max_iter; ++it) {
      const int buf = it & 1;
      // The serial part of an iteration on FIVE warps (every dependent latency of it is exposed, so it is spread, not
      // shortened): warp 1 forms the three dot products; warps 0 / 3 form m for the lane rows 0..31 and publish it to
      // the peers 0..3 / 4..7, warps 2 / 4 do the same for the rows 32..47.  All of them read what the recurrences of
      // the previous iteration (warps 0 and 2, below) left in shared memory: a 160-thread named barrier orders that.
      if (wid <= 4) {
        asm volatile("bar.sync 1, 160;" ::: "memory");
        if (wid == 1) {
          const double ra = rl[ia], rb = rl[ib], ua = ul[ia], ub = ul[ib], wa = wl[ia], wb = wl[ib];
          double gam_p = fma(ra, ua, rb * ub), del_p = fma(wa, ua, wb * ub), rr_p = fma(ra, ra, rb * rb);
          warp_sum3(gam_p, del_p, rr_p);
#pragma unroll
          for (int e0 = 0; e0 < 3 * MAXG; e0 += 32) {
            const int e = e0 + lane;
            if (e < 3 * G) {
              const int k = e / 3, which = e - 3 * k;
              const double v = which == 0 ? gam_p : (which == 1 ? del_p : rr_p);
              const uint32_t rk = cluster_map_u32(sm, k) - smem_u32(sm);
              st_async_f64(smem_u32(pch + buf * 3 * MAXG + which * MAXG + g) + rk, v, smem_u32(bars + buf) + rk);
            }
          }
        } else {
          if (wid == 0 && lane == 0) mbar_expect_tx(bars + buf, (uint32_t)(8 * (n + 3 * G)));
          // publisher p = 0..3 (warps 0, 2, 3, 4).  Two rows per lane: rows lane (p even) / lane + 32 (p odd), peers in
          // two halves; one row per lane: the peers in four quarters
          const int pidx = wid == 0 ? 0 : wid - 1;
          constexpr int PN = NRL == 2 ? MAXG / 2 : MAXG / 4;   // peers per publisher
          const int i = (NRL == 2 && (pidx & 1)) ? ib : ia;
          const int k0 = (NRL == 2 ? (pidx >> 1) : pidx) * PN;
          if (i < R && k0 < G) {
            const double m = minv_row(i, wl);              // m = M^-1 w
            if (NRL == 2 ? pidx < 2 : pidx == 0) ml[i] = m;
            const uint32_t a = smem_u32(u_buf + (size_t)buf * n + row0 + i), bb = smem_u32(bars + buf);
#pragma unroll
            for (int k = 0; k < MAXG; ++k)
              if (k >= k0 && k < k0 + PN && k < G) st_async_f64(a + rbase[k], m, bb + rbase[k]);
          }
        }
      }
      // phase of this buffer's barrier: buffer 1 was used once by the initial exchange
      if (!mbar_wait_bounded(bars + buf, (uint32_t)(((it + buf) >> 1) & 1))) { lost = true; status = 1; break; }
      // (all 24 loads first, whatever G: with a branch per CTA the loop was eight load -> add round trips)
      double gamma = 0.0, delta = 0.0, rr = 0.0;
      {
        double pv[3 * MAXG];
#pragma unroll
        for (int k = 0; k < 3 * MAXG; ++k) pv[k] = pch[buf * 3 * MAXG + k];
#pragma unroll
        for (int k = 0; k < MAXG; ++k) {
          gamma += k < G ? pv[k] : 0.0; delta += k < G ? pv[MAXG + k] : 0.0; rr += k < G ? pv[2 * MAXG + k] : 0.0;
        }
      }
      if (it == 0) {
        if (!(rr > 0.0)) { status = 0; break; }        // b == 0: x = 0
        thresh = A.tol * A.tol * rr;
      }
      if (rr <= thresh && it > 0) { status = 0; break; }
      if (it == A.max_iter) break;
      partial_product(buf);                            // n = S m
      double beta = 0.0, pap = delta;
      if (it > 0) {
        beta = gamma * inv_gamma_old;
        pap = delta - beta * beta * pap_old;
      }
      if (!(pap > 1e-290) || !(pap < 1e290) || !(gamma > 1e-290) || !(gamma < 1e290)) status = 1;
      const double alpha = gamma * fast_rcp(pap);
      inv_gamma_old = fast_rcp(gamma); pap_old = pap;
      __syncthreads();
      if (status == 1) break;                          // (uniform: every thread of every CTA has the same scalars)
      if (wid == 0 || (NRL == 2 && wid == 2)) {        // the recurrences: warp 0 rows 0..31, warp 2 rows 32..47
        // (every load ahead of the first store: the arrays may alias as far as the compiler knows, and
        //  load -> FMA -> store -> load ... chains made this 640 cycles)
        const int i = wid == 0 ? ia : ib;
        if (i < R) {
          const double nv = summed_product(i);
          const double zo = zl[i], mo = ml[i], qo = ql[i], wo = wl[i], so = sl[i], uo = ul[i], po = pl[i], xo = xl[i],
                       ro = rl[i];
          const double z = nv + beta * zo;
          const double q = mo + beta * qo;
          const double sv = wo + beta * so;
          const double pv = uo + beta * po;
          zl[i] = z; ql[i] = q; sl[i] = sv; pl[i] = pv;
          xl[i] = xo + alpha * pv;
          rl[i] = ro - alpha * sv;
          ul[i] = uo - alpha * q;
          wl[i] = wo - alpha * z;
        }
      }
    }
  }
  if (wid == 0) {
    if (ha) A.x[row0 + ia] = (status == 1) ? 0.0 : xl[ia];
    if (hb) A.x[row0 + ib] = (status == 1) ? 0.0 : xl[ib];
    if (g == 0 && lane == 0) { *A.status = status; *A.iters = it; }
  }
  cluster.sync();                                      // nobody leaves while its shared memory may still be written
}

}  // namespace vslam
