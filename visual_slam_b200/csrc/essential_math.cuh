// Device math for the two-view path: five-point essential-matrix solver (Nister), Sampson error with
// OpenCV's float32 semantics, one-sided Jacobi SVD (3x3 decomposition of E, 4x4 DLT triangulation).
// Restates what cv2.findEssentialMat / cv2.recoverPose do behind reference
// src/v2/helper_functions.py:52 and :176 (see oracle/essential.py).  All FP64.
#pragma once
#include "common.cuh"

#ifndef VSLAM_HD
#define VSLAM_HD __host__ __device__
#endif

namespace vslam {

// ---- trivariate polynomials in (x, y, z) ------------------------------------------------------
// degree 1: [x, y, z, 1]
// degree 2: [x^2, xy, xz, x, y^2, yz, y, z^2, z, 1]
// degree 3 (Nister's elimination order): [x^3, y^3, x^2y, xy^2, x^2z, x^2, y^2z, y^2, xyz, xy,
//                                         xz^2, xz, x, yz^2, yz, y, z^3, z^2, z, 1]
VSLAM_HD __forceinline__ int mul11_idx(int a, int b) {
  constexpr int t[4][4] = {{0, 1, 2, 3}, {1, 4, 5, 6}, {2, 5, 7, 8}, {3, 6, 8, 9}};
  return t[a][b];
}
VSLAM_HD __forceinline__ int mul21_idx(int a, int b) {
  constexpr int t[10][4] = {{0, 2, 4, 5},    {2, 3, 8, 9},     {4, 8, 10, 11},   {5, 9, 11, 12},
                            {3, 1, 6, 7},    {8, 6, 13, 14},   {9, 7, 14, 15},   {10, 13, 16, 17},
                            {11, 14, 17, 18}, {12, 15, 18, 19}};
  return t[a][b];
}
// c(deg 2) += s * a(deg 1) * b(deg 1)
VSLAM_HD __forceinline__ void pmul11_acc(const double* a, const double* b, double s, double* c) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) c[mul11_idx(i, j)] += s * a[i] * b[j];
  }
}
// c(deg 3) += s * a(deg 2) * b(deg 1)
VSLAM_HD __forceinline__ void pmul21_acc(const double* a, const double* b, double s, double* c) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) c[mul21_idx(i, j)] += s * a[i] * b[j];
  }
}

// Orthonormal basis of the null space of the 5x9 epipolar system (rows q2 (x) q1, row-major vec E).
// Rows are orthonormalised by modified Gram-Schmidt (twice), then a pivoted Cholesky of the projector
// I - sum q q^T yields four orthonormal null vectors.  basis[4][9].  Returns false when degenerate.
VSLAM_HD inline bool epipolar_null_space(const double (*x1)[2], const double (*x2)[2], double (*basis)[9]) {
  double q[5][9];
  for (int i = 0; i < 5; ++i) {
    const double a = x1[i][0], b = x1[i][1], c = x2[i][0], d = x2[i][1];
    q[i][0] = c * a; q[i][1] = c * b; q[i][2] = c;
    q[i][3] = d * a; q[i][4] = d * b; q[i][5] = d;
    q[i][6] = a;     q[i][7] = b;     q[i][8] = 1.0;
  }
  for (int i = 0; i < 5; ++i) {
    double n0 = 0.0;
    for (int k = 0; k < 9; ++k) n0 += q[i][k] * q[i][k];
    for (int pass = 0; pass < 2; ++pass) {
      for (int j = 0; j < i; ++j) {
        double dot = 0.0;
        for (int k = 0; k < 9; ++k) dot += q[i][k] * q[j][k];
        for (int k = 0; k < 9; ++k) q[i][k] -= dot * q[j][k];
      }
    }
    double n1 = 0.0;
    for (int k = 0; k < 9; ++k) n1 += q[i][k] * q[i][k];
    if (!(n1 > 1e-20 * n0) || !(n1 > 0.0)) return false;  // rank-deficient sample
    const double inv = 1.0 / sqrt(n1);
    for (int k = 0; k < 9; ++k) q[i][k] *= inv;
  }
  double P[9][9];
  for (int r = 0; r < 9; ++r)
    for (int c = 0; c < 9; ++c) {
      double s = (r == c) ? 1.0 : 0.0;
      for (int i = 0; i < 5; ++i) s -= q[i][r] * q[i][c];
      P[r][c] = s;
    }
  for (int k = 0; k < 4; ++k) {
    int piv = 0;
    for (int j = 1; j < 9; ++j)
      if (P[j][j] > P[piv][piv]) piv = j;
    const double d = P[piv][piv];
    if (!(d > 1e-12)) return false;
    const double inv = 1.0 / sqrt(d);
    double v[9];
    for (int r = 0; r < 9; ++r) v[r] = P[r][piv] * inv;
    // re-orthogonalise against the row space and the vectors found so far, renormalise
    for (int pass = 0; pass < 2; ++pass) {
      for (int i = 0; i < 5; ++i) {
        double dot = 0.0;
        for (int r = 0; r < 9; ++r) dot += v[r] * q[i][r];
        for (int r = 0; r < 9; ++r) v[r] -= dot * q[i][r];
      }
      for (int j = 0; j < k; ++j) {
        double dot = 0.0;
        for (int r = 0; r < 9; ++r) dot += v[r] * basis[j][r];
        for (int r = 0; r < 9; ++r) v[r] -= dot * basis[j][r];
      }
    }
    double n = 0.0;
    for (int r = 0; r < 9; ++r) n += v[r] * v[r];
    n = 1.0 / sqrt(n);
    for (int r = 0; r < 9; ++r) { v[r] *= n; basis[k][r] = v[r]; }
    for (int r = 0; r < 9; ++r)
      for (int c = 0; c < 9; ++c) P[r][c] -= v[r] * v[c];
  }
  return true;
}

// The ten cubic constraints det(E) = 0, E E^T E - 1/2 tr(E E^T) E = 0 for E = x E1 + y E2 + z E3 + E4,
// as a 10x20 coefficient matrix in the degree-3 monomial order above.
VSLAM_HD inline void essential_constraints(const double (*basis)[9], double (*A)[20]) {
  double e[9][4];  // entry (r,c) as a degree-1 polynomial
  for (int i = 0; i < 9; ++i)
    for (int k = 0; k < 4; ++k) e[i][k] = basis[k][i];
  for (int r = 0; r < 10; ++r)
    for (int c = 0; c < 20; ++c) A[r][c] = 0.0;
  // det(E) by the first row
  {
    double m[10];
    for (int i = 0; i < 10; ++i) m[i] = 0.0;
    pmul11_acc(e[4], e[8], 1.0, m); pmul11_acc(e[5], e[7], -1.0, m);
    pmul21_acc(m, e[0], 1.0, A[0]);
    for (int i = 0; i < 10; ++i) m[i] = 0.0;
    pmul11_acc(e[5], e[6], 1.0, m); pmul11_acc(e[3], e[8], -1.0, m);
    pmul21_acc(m, e[1], 1.0, A[0]);
    for (int i = 0; i < 10; ++i) m[i] = 0.0;
    pmul11_acc(e[3], e[7], 1.0, m); pmul11_acc(e[4], e[6], -1.0, m);
    pmul21_acc(m, e[2], 1.0, A[0]);
  }
  // L = E E^T - 1/2 tr(E E^T) I  (symmetric, degree 2)
  double L[6][10];  // (0,0) (0,1) (0,2) (1,1) (1,2) (2,2)
  {
    int k = 0;
    for (int r = 0; r < 3; ++r)
      for (int c = r; c < 3; ++c, ++k) {
        for (int i = 0; i < 10; ++i) L[k][i] = 0.0;
        for (int j = 0; j < 3; ++j) pmul11_acc(e[3 * r + j], e[3 * c + j], 1.0, L[k]);
      }
    for (int i = 0; i < 10; ++i) {
      const double h = 0.5 * (L[0][i] + L[3][i] + L[5][i]);
      L[0][i] -= h; L[3][i] -= h; L[5][i] -= h;
    }
  }
  const int sym[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c)
      for (int k = 0; k < 3; ++k) pmul21_acc(L[sym[r][k]], e[3 * k + c], 1.0, A[1 + 3 * r + c]);
}

// Gauss-Jordan with partial pivoting on the ten leading columns of A (10x20), in place.
VSLAM_HD inline bool gauss_jordan_10x20(double (*A)[20]) {
  for (int col = 0; col < 10; ++col) {
    int piv = col;
    double best = fabs(A[col][col]);
    for (int r = col + 1; r < 10; ++r)
      if (fabs(A[r][col]) > best) { best = fabs(A[r][col]); piv = r; }
    if (!(best > 1e-300)) return false;
    if (piv != col)
      for (int c = col; c < 20; ++c) { const double t = A[col][c]; A[col][c] = A[piv][c]; A[piv][c] = t; }
    const double inv = 1.0 / A[col][col];
    for (int c = col; c < 20; ++c) A[col][c] *= inv;
    for (int r = 0; r < 10; ++r) {
      if (r == col) continue;
      const double f = A[r][col];
      if (f == 0.0) continue;
      for (int c = col; c < 20; ++c) A[r][c] -= f * A[col][c];
    }
  }
  return true;
}

// ---- univariate polynomials (ascending coefficients) and the degree-10 root finder -------------
struct cplx { double re, im; };
VSLAM_HD __forceinline__ cplx cmul(cplx a, cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
VSLAM_HD __forceinline__ cplx cdiv(cplx a, cplx b) {
  // Smith's algorithm
  if (fabs(b.re) >= fabs(b.im)) {
    const double r = b.im / b.re, d = b.re + b.im * r;
    return {(a.re + a.im * r) / d, (a.im - a.re * r) / d};
  }
  const double r = b.re / b.im, d = b.re * r + b.im;
  return {(a.re * r + a.im) / d, (a.im * r - a.re) / d};
}

// All complex roots of c[0] + c[1] z + ... + c[deg] z^deg by Aberth-Ehrlich iteration.
// Returns the number of iterations used, or -1 if the polynomial is degenerate.
template <int MAXDEG>
VSLAM_HD inline int aberth_roots(const double* c, int deg, cplx* z) {
  while (deg > 0 && c[deg] == 0.0) --deg;
  if (deg <= 0) return -1;
  // starting points on a circle of the Cauchy-bound radius, slightly rotated off the axes
  double rad = 0.0;
  for (int i = 0; i < deg; ++i) rad = fmax(rad, fabs(c[i] / c[deg]));
  rad = 1.0 + rad;
  // a tighter start: geometric-mean radius |c0/cn|^(1/n), clamped to the bound
  double r0 = (c[0] != 0.0) ? pow(fabs(c[0] / c[deg]), 1.0 / deg) : 1.0;
  if (!(r0 > 1e-8)) r0 = 1e-8;
  if (r0 > rad) r0 = rad;
  for (int i = 0; i < deg; ++i) {
    double s, co;
    sincos(2.0 * 3.14159265358979323846 * i / deg + 0.4, &s, &co);
    z[i] = {r0 * co, r0 * s};
  }
  // A root is FINISHED (and frozen) once its correction is at rounding level relative to the root, or once p(z) is at
  // the rounding level of its own evaluation (backward-stable root: |p(z)| <= 32 eps sum |c_k| |z|^k) -- the second
  // test is what ends the iteration for roots whose corrections keep dithering in the last bits (a test on the
  // largest relative step alone let about a fifth of all five-point polynomials run into the iteration cap, and
  // in a warp one such lane holds the other 31).  The real roots are polished afterwards (polish_xyz).
  unsigned finished = 0;
  const unsigned all = (1u << deg) - 1u;
  int it = 0;
  for (; it < 200 && finished != all; ++it) {
    for (int i = 0; i < deg; ++i) {
      if ((finished >> i) & 1u) continue;
      // Horner for p and p', and for the evaluation bound sum |c_k| |z|^k
      const cplx zi = z[i];
      const double az = sqrt(zi.re * zi.re + zi.im * zi.im);
      cplx p = {c[deg], 0.0}, dp = {0.0, 0.0};
      double bound = fabs(c[deg]);
      for (int k = deg - 1; k >= 0; --k) {
        dp = cmul(dp, zi); dp.re += p.re; dp.im += p.im;
        p = cmul(p, zi); p.re += c[k];
        bound = bound * az + fabs(c[k]);
      }
      if (fabs(p.re) + fabs(p.im) <= 7.2e-15 * bound) { finished |= 1u << i; continue; }
      cplx w = cdiv(p, dp);
      if (!(isfinite(w.re) && isfinite(w.im))) { z[i].re += 1e-3 * r0; z[i].im += 1e-3 * r0; continue; }
      cplx s = {0.0, 0.0};
      for (int j = 0; j < deg; ++j) {
        if (j == i) continue;
        // 1 / (z_i - z_j) = conj(d) / |d|^2: one division (coincident or absurdly distant roots give a non-finite
        // term, which is left out exactly like before)
        const double dr = zi.re - z[j].re, di = zi.im - z[j].im;
        const double q = 1.0 / (dr * dr + di * di);
        const double ir = dr * q, ii = -di * q;
        if (isfinite(ir) && isfinite(ii)) { s.re += ir; s.im += ii; }
      }
      const cplx ws = cmul(w, s);
      const cplx step = cdiv(w, {1.0 - ws.re, -ws.im});
      if (!(isfinite(step.re) && isfinite(step.im))) continue;
      z[i].re -= step.re; z[i].im -= step.im;
      const double m = fmax(fabs(step.re), fabs(step.im)) / fmax(1e-300, fmax(fabs(z[i].re), fabs(z[i].im)));
      if (m < 1e-15) finished |= 1u << i;
    }
  }
#ifdef VSLAM_ABERTH_HOOK
  VSLAM_ABERTH_HOOK(it);
#endif
  return it;
}

VSLAM_HD __forceinline__ double polyval_asc(const double* p, int deg, double z) {
  double s = p[deg];
  for (int k = deg - 1; k >= 0; --k) s = s * z + p[k];
  return s;
}

// p (deg a+b) = q (deg a) * r (deg b), ascending
VSLAM_HD __forceinline__ void polymul_asc(const double* q, int a, const double* r, int b, double* p) {
  for (int i = 0; i <= a + b; ++i) p[i] = 0.0;
  for (int i = 0; i <= a; ++i)
    for (int j = 0; j <= b; ++j) p[i + j] += q[i] * r[j];
}

VSLAM_HD __forceinline__ void mat3_mul(const double* A, const double* B, double* C) {
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c)
      C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}
VSLAM_HD __forceinline__ void mat3_mul_nt(const double* A, const double* B, double* C) {  // A B^T
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c)
      C[3 * r + c] = A[3 * r] * B[3 * c] + A[3 * r + 1] * B[3 * c + 1] + A[3 * r + 2] * B[3 * c + 2];
}
VSLAM_HD __forceinline__ double det3(const double* M) {
  return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) +
         M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// Residual of the ten constraints at E (det, then E E^T E - 1/2 tr(E E^T) E row-major).
VSLAM_HD inline void essential_residual(const double* E, double* r) {
  double EEt[9], T[9];
  mat3_mul_nt(E, E, EEt);
  const double h = 0.5 * (EEt[0] + EEt[4] + EEt[8]);
  EEt[0] -= h; EEt[4] -= h; EEt[8] -= h;
  mat3_mul(EEt, E, T);
  r[0] = det3(E);
  for (int i = 0; i < 9; ++i) r[1 + i] = T[i];
}

// Gauss-Newton polish of (x, y, z) on the ten constraints (consistent, over-determined system):
// removes the conditioning loss of the degree-10 polynomial route.  Central differences are exact
// enough here because the residual is a cubic polynomial evaluated in FP64 and the result is only
// accepted when the residual norm decreases.
VSLAM_HD inline void polish_xyz(const double (*basis)[9], double* xyz) {
  auto eval = [&](const double* v, double* r) {
    double E[9];
    for (int i = 0; i < 9; ++i) E[i] = v[0] * basis[0][i] + v[1] * basis[1][i] + v[2] * basis[2][i] + basis[3][i];
    essential_residual(E, r);
  };
  double r0[10];
  eval(xyz, r0);
  double n0 = 0.0;
  for (int i = 0; i < 10; ++i) n0 += r0[i] * r0[i];
  for (int it = 0; it < 4; ++it) {
    // analytic Jacobian: d r / d v_k with dE = basis[k]
    double E[9];
    for (int i = 0; i < 9; ++i) E[i] = xyz[0] * basis[0][i] + xyz[1] * basis[1][i] + xyz[2] * basis[2][i] + basis[3][i];
    double J[10][3];
    double EEt[9];
    mat3_mul_nt(E, E, EEt);
    const double tr = EEt[0] + EEt[4] + EEt[8];
    for (int k = 0; k < 3; ++k) {
      const double* D = basis[k];
      // d det = sum cof(E)_ij D_ij
      const double cof[9] = {E[4] * E[8] - E[5] * E[7], E[5] * E[6] - E[3] * E[8], E[3] * E[7] - E[4] * E[6],
                             E[2] * E[7] - E[1] * E[8], E[0] * E[8] - E[2] * E[6], E[1] * E[6] - E[0] * E[7],
                             E[1] * E[5] - E[2] * E[4], E[2] * E[3] - E[0] * E[5], E[0] * E[4] - E[1] * E[3]};
      double dd = 0.0;
      for (int i = 0; i < 9; ++i) dd += cof[i] * D[i];
      J[0][k] = dd;
      // d(E E^T E) = D E^T E + E D^T E + E E^T D ;  d(1/2 tr(EE^T) E) = tr(E D^T) E + 1/2 tr(EE^T) D
      double DEt[9], EDt[9], A1[9], A2[9], A3[9];
      mat3_mul_nt(D, E, DEt); mat3_mul_nt(E, D, EDt);
      mat3_mul(DEt, E, A1); mat3_mul(EDt, E, A2); mat3_mul(EEt, D, A3);
      const double trd = EDt[0] + EDt[4] + EDt[8];
      for (int i = 0; i < 9; ++i) J[1 + i][k] = A1[i] + A2[i] + A3[i] - trd * E[i] - 0.5 * tr * D[i];
    }
    double H[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
    for (int i = 0; i < 10; ++i) {
      H[0] += J[i][0] * J[i][0]; H[1] += J[i][0] * J[i][1]; H[2] += J[i][0] * J[i][2];
      H[3] += J[i][1] * J[i][1]; H[4] += J[i][1] * J[i][2]; H[5] += J[i][2] * J[i][2];
      g[0] += J[i][0] * r0[i]; g[1] += J[i][1] * r0[i]; g[2] += J[i][2] * r0[i];
    }
    // solve the symmetric 3x3 by cofactors
    const double c00 = H[3] * H[5] - H[4] * H[4], c01 = H[2] * H[4] - H[1] * H[5], c02 = H[1] * H[4] - H[2] * H[3];
    const double det = H[0] * c00 + H[1] * c01 + H[2] * c02;
    if (!(fabs(det) > 0.0) || !isfinite(det)) break;
    const double c11 = H[0] * H[5] - H[2] * H[2], c12 = H[1] * H[2] - H[0] * H[4], c22 = H[0] * H[3] - H[1] * H[1];
    const double d0 = -(c00 * g[0] + c01 * g[1] + c02 * g[2]) / det;
    const double d1 = -(c01 * g[0] + c11 * g[1] + c12 * g[2]) / det;
    const double d2 = -(c02 * g[0] + c12 * g[1] + c22 * g[2]) / det;
    const double trial[3] = {xyz[0] + d0, xyz[1] + d1, xyz[2] + d2};
    double r1[10];
    eval(trial, r1);
    double n1 = 0.0;
    for (int i = 0; i < 10; ++i) n1 += r1[i] * r1[i];
    if (!(n1 < n0)) break;
    xyz[0] = trial[0]; xyz[1] = trial[1]; xyz[2] = trial[2];
    for (int i = 0; i < 10; ++i) r0[i] = r1[i];
    n0 = n1;
  }
}

// Canonical sign and order of a model set (neither cv2's nor numpy's root order is canonical):
// sign such that <E, g1> > 0, ascending <E, g2>.
VSLAM_HD __forceinline__ double canon_key1(const double* E) {
  return 0.9 * E[0] - 0.5 * E[1] + 0.3 * E[2] + 0.7 * E[3] - 0.2 * E[4] + 0.6 * E[5] - 0.8 * E[6] + 0.4 * E[7] + 0.1 * E[8];
}
VSLAM_HD __forceinline__ double canon_key2(const double* E) {
  return 0.2 * E[0] + 0.7 * E[1] - 0.6 * E[2] - 0.1 * E[3] + 0.9 * E[4] + 0.3 * E[5] + 0.5 * E[6] - 0.4 * E[7] + 0.8 * E[8];
}

// Five-point solver.  x1, x2: five normalised correspondences.  models[10][9] (unit Frobenius norm,
// canonical sign and order).  Returns the number of models (0..10).
VSLAM_HD inline int five_point(const double (*x1)[2], const double (*x2)[2], double (*models)[9]) {
  double basis[4][9];
  if (!epipolar_null_space(x1, x2, basis)) return 0;
  double A[10][20];
  essential_constraints(basis, A);
  if (!gauss_jordan_10x20(A)) return 0;
  double B[3][3][5];
  for (int i = 0; i < 3; ++i) {
    const double* a = A[4 + 2 * i];
    const double* b = A[5 + 2 * i];
    B[i][0][0] = a[12]; B[i][0][1] = a[11] - b[12]; B[i][0][2] = a[10] - b[11]; B[i][0][3] = -b[10]; B[i][0][4] = 0.0;
    B[i][1][0] = a[15]; B[i][1][1] = a[14] - b[15]; B[i][1][2] = a[13] - b[14]; B[i][1][3] = -b[13]; B[i][1][4] = 0.0;
    B[i][2][0] = a[19]; B[i][2][1] = a[18] - b[19]; B[i][2][2] = a[17] - b[18]; B[i][2][3] = a[16] - b[17]; B[i][2][4] = -b[16];
  }
  // det B(z) = sum_i B[i][2] * cofactor_i, cofactor_i from columns 0 and 1 (degree 6)
  double det[11];
  for (int i = 0; i < 11; ++i) det[i] = 0.0;
  for (int i = 0; i < 3; ++i) {
    const int j = (i + 1) % 3, k = (i + 2) % 3;
    double m1[7], m2[7], t[11];
    polymul_asc(B[j][0], 3, B[k][1], 3, m1);
    polymul_asc(B[j][1], 3, B[k][0], 3, m2);
    for (int q = 0; q < 7; ++q) m1[q] -= m2[q];
    polymul_asc(m1, 6, B[i][2], 4, t);
    for (int q = 0; q < 11; ++q) det[q] += t[q];
  }
  bool finite = true, nonzero = false;
  for (int i = 0; i < 11; ++i) { finite = finite && isfinite(det[i]); nonzero = nonzero || det[i] != 0.0; }
  if (!finite || !nonzero) return 0;
  cplx roots[10];
  int deg = 10;
  while (deg > 0 && det[deg] == 0.0) --deg;
  if (aberth_roots<10>(det, deg, roots) < 0) return 0;
  int count = 0;
  for (int i = 0; i < deg; ++i) {
    if (fabs(roots[i].im) > 1e-10) continue;
    const double z = roots[i].re;
    double Bz[3][3];
    for (int r = 0; r < 3; ++r) {
      Bz[r][0] = polyval_asc(B[r][0], 3, z);
      Bz[r][1] = polyval_asc(B[r][1], 3, z);
      Bz[r][2] = polyval_asc(B[r][2], 4, z);
    }
    // null vector of Bz: the largest cross product of two rows
    double best[3] = {0, 0, 0}, bn = -1.0;
    for (int r = 0; r < 3; ++r) {
      const double* u = Bz[r];
      const double* v = Bz[(r + 1) % 3];
      const double c[3] = {u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]};
      const double n = c[0] * c[0] + c[1] * c[1] + c[2] * c[2];
      if (n > bn) { bn = n; best[0] = c[0]; best[1] = c[1]; best[2] = c[2]; }
    }
    if (!(bn > 0.0)) continue;
    if (fabs(best[2]) < 1e-10 * sqrt(bn)) continue;
    double xyz[3] = {best[0] / best[2], best[1] / best[2], z};
    polish_xyz(basis, xyz);
    double E[9], n = 0.0;
    for (int q = 0; q < 9; ++q) {
      E[q] = xyz[0] * basis[0][q] + xyz[1] * basis[1][q] + xyz[2] * basis[2][q] + basis[3][q];
      n += E[q] * E[q];
    }
    if (!(n > 0.0) || !isfinite(n)) continue;
    n = 1.0 / sqrt(n);
    if (canon_key1(E) < 0.0) n = -n;
    for (int q = 0; q < 9; ++q) models[count][q] = E[q] * n;
    ++count;
  }
  // insertion sort by the second key
  for (int i = 1; i < count; ++i) {
    double cur[9];
    for (int q = 0; q < 9; ++q) cur[q] = models[i][q];
    const double kc = canon_key2(cur);
    int j = i - 1;
    while (j >= 0 && canon_key2(models[j]) > kc) {
      for (int q = 0; q < 9; ++q) models[j + 1][q] = models[j][q];
      --j;
    }
    for (int q = 0; q < 9; ++q) models[j + 1][q] = cur[q];
  }
  return count;
}

// ---- one-sided Jacobi SVD (Hestenes) for small square matrices ------------------------------------
// On return the columns of A are U * diag(sigma) (mutually orthogonal), V holds the right singular
// vectors as columns, sig2[j] = |A[:,j]|^2.  Unsorted.
template <int N>
VSLAM_HD inline void jacobi_svd(double (*A)[N], double (*V)[N], double* sig2) {
  for (int r = 0; r < N; ++r)
    for (int c = 0; c < N; ++c) V[r][c] = (r == c) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 40; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < N - 1; ++p) {
      for (int q = p + 1; q < N; ++q) {
        double alpha = 0.0, beta = 0.0, gamma = 0.0;
        for (int k = 0; k < N; ++k) {
          alpha += A[k][p] * A[k][p]; beta += A[k][q] * A[k][q]; gamma += A[k][p] * A[k][q];
        }
        if (gamma == 0.0 || fabs(gamma) <= 1e-16 * sqrt(alpha * beta)) continue;
        rotated = true;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int k = 0; k < N; ++k) {
          const double a = A[k][p], b = A[k][q];
          A[k][p] = c * a - s * b; A[k][q] = s * a + c * b;
          const double va = V[k][p], vb = V[k][q];
          V[k][p] = c * va - s * vb; V[k][q] = s * va + c * vb;
        }
      }
    }
    if (!rotated) break;
  }
  for (int j = 0; j < N; ++j) {
    double s = 0.0;
    for (int k = 0; k < N; ++k) s += A[k][j] * A[k][j];
    sig2[j] = s;
  }
}

// cv2.decomposeEssentialMat: R1 = U W V^T, R2 = U W^T V^T, t = u3 with det U = det V = +1.
VSLAM_HD inline void decompose_essential(const double* E, double* R1, double* R2, double* t) {
  double A[3][3], V[3][3], s2[3];
  for (int i = 0; i < 9; ++i) A[i / 3][i % 3] = E[i];
  jacobi_svd<3>(A, V, s2);
  // order the two largest singular values first
  int o[3] = {0, 1, 2};
  for (int i = 0; i < 2; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (s2[o[j]] > s2[o[i]]) { const int tmp = o[i]; o[i] = o[j]; o[j] = tmp; }
  double U[9], Vm[9];
  for (int c = 0; c < 2; ++c) {
    const double inv = 1.0 / sqrt(s2[o[c]]);
    for (int r = 0; r < 3; ++r) U[3 * r + c] = A[r][o[c]] * inv;
  }
  // u2 re-orthogonalised against u1 (equal singular values leave them orthogonal only to rounding)
  {
    double d = U[0] * U[1] + U[3] * U[4] + U[6] * U[7];
    U[1] -= d * U[0]; U[4] -= d * U[3]; U[7] -= d * U[6];
    d = 1.0 / sqrt(U[1] * U[1] + U[4] * U[4] + U[7] * U[7]);
    U[1] *= d; U[4] *= d; U[7] *= d;
  }
  U[2] = U[3] * U[7] - U[6] * U[4];
  U[5] = U[6] * U[1] - U[0] * U[7];
  U[8] = U[0] * U[4] - U[3] * U[1];
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) Vm[3 * r + c] = V[r][o[c]];
  if (det3(Vm) < 0.0)
    for (int i = 0; i < 9; ++i) Vm[i] = -Vm[i];
  // U W = [-u2, u1, u3],  U W^T = [u2, -u1, u3]
  double UW[9], UWt[9];
  for (int r = 0; r < 3; ++r) {
    UW[3 * r] = -U[3 * r + 1]; UW[3 * r + 1] = U[3 * r]; UW[3 * r + 2] = U[3 * r + 2];
    UWt[3 * r] = U[3 * r + 1]; UWt[3 * r + 1] = -U[3 * r]; UWt[3 * r + 2] = U[3 * r + 2];
  }
  mat3_mul_nt(UW, Vm, R1);
  mat3_mul_nt(UWt, Vm, R2);
  t[0] = U[2]; t[1] = U[5]; t[2] = U[8];
}

// cv2.triangulatePoints for P0 = [I|0], P1 = [R|t]: right singular vector of the smallest singular
// value of the 4x4 DLT system.  X[4] is a unit vector with free sign.
VSLAM_HD inline void triangulate_dlt(const double* R, const double* t, double x1, double y1, double x2,
                                     double y2, double* X) {
  double A[4][4], V[4][4], s2[4];
  A[0][0] = -1.0; A[0][1] = 0.0;  A[0][2] = x1; A[0][3] = 0.0;
  A[1][0] = 0.0;  A[1][1] = -1.0; A[1][2] = y1; A[1][3] = 0.0;
  for (int c = 0; c < 3; ++c) {
    A[2][c] = x2 * R[6 + c] - R[c];
    A[3][c] = y2 * R[6 + c] - R[3 + c];
  }
  A[2][3] = x2 * t[2] - t[0];
  A[3][3] = y2 * t[2] - t[1];
  jacobi_svd<4>(A, V, s2);
  int m = 0;
  for (int j = 1; j < 4; ++j)
    if (s2[j] < s2[m]) m = j;
  for (int r = 0; r < 4; ++r) X[r] = V[r][m];
}

// Cheirality / distance test of cv2.recoverPose for one point and one candidate.
VSLAM_HD inline bool cheirality_ok(const double* R, const double* t, const double* Q, double dist) {
  bool ok = (Q[2] * Q[3]) > 0.0;
  const double X = Q[0] / Q[3], Y = Q[1] / Q[3], Z = Q[2] / Q[3], W = Q[3] / Q[3];
  ok = ok && (Z < dist);
  const double z2 = R[6] * X + R[7] * Y + R[8] * Z + t[2] * W;
  ok = ok && (z2 > 0.0) && (z2 < dist);
  return ok;
}

}  // namespace vslam
