// Device math for the RANSAC PnP path: Rodrigues, quartic roots, P3P (+1), all FP64.
// Restates what cv2.solvePnPRansac does around reference src/v2/main.py:196-197 (see oracle/pnp.py).
#pragma once
#include "common.cuh"

namespace vslam {

__device__ __forceinline__ void rodrigues(const double* r, double* R) {
  const double th2 = r[0] * r[0] + r[1] * r[1] + r[2] * r[2];
  const double th = sqrt(th2);
  if (th < 1e-300) {
    R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
    return;
  }
  const double kx = r[0] / th, ky = r[1] / th, kz = r[2] / th;
  double s, c;
  sincos(th, &s, &c);
  const double v = 1.0 - c;
  R[0] = c + kx * kx * v;      R[1] = kx * ky * v - kz * s; R[2] = kx * kz * v + ky * s;
  R[3] = ky * kx * v + kz * s; R[4] = c + ky * ky * v;      R[5] = ky * kz * v - kx * s;
  R[6] = kz * kx * v - ky * s; R[7] = kz * ky * v + kx * s; R[8] = c + kz * kz * v;
}

// Rotation matrix -> rotation vector.
__device__ inline void rodrigues_inv(const double* R, double* r) {
  double c = (R[0] + R[4] + R[8] - 1.0) * 0.5;
  c = fmin(1.0, fmax(-1.0, c));
  const double th = acos(c);
  const double vx = R[7] - R[5], vy = R[2] - R[6], vz = R[3] - R[1];
  const double s = 0.5 * sqrt(vx * vx + vy * vy + vz * vz);
  if (th < 1e-12) { r[0] = 0.5 * vx; r[1] = 0.5 * vy; r[2] = 0.5 * vz; return; }
  if (3.14159265358979323846 - th < 1e-6) {
    const double a[3] = {0.5 * (R[0] + 1.0), 0.5 * (R[4] + 1.0), 0.5 * (R[8] + 1.0)};
    int i = 0;
    if (a[1] > a[i]) i = 1;
    if (a[2] > a[i]) i = 2;
    const double d = sqrt(fmax(a[i], 0.0));
    double ax[3] = {0.5 * (R[0 + i] + (i == 0 ? 1.0 : 0.0)), 0.5 * (R[3 + i] + (i == 1 ? 1.0 : 0.0)),
                    0.5 * (R[6 + i] + (i == 2 ? 1.0 : 0.0))};
    ax[0] /= d; ax[1] /= d; ax[2] /= d;
    if (ax[0] * vx + ax[1] * vy + ax[2] * vz < 0) { ax[0] = -ax[0]; ax[1] = -ax[1]; ax[2] = -ax[2]; }
    const double n = sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]);
    r[0] = ax[0] / n * th; r[1] = ax[1] / n * th; r[2] = ax[2] / n * th;
    return;
  }
  const double f = th / (2.0 * s);
  r[0] = vx * f; r[1] = vy * f; r[2] = vz * f;
}

// One real root of x^3 + a x^2 + b x + c (the largest), Cardano / trigonometric.
__device__ inline double cubic_largest_real_root(double a, double b, double c) {
  const double q = (a * a - 3.0 * b) / 9.0;
  const double r = (2.0 * a * a * a - 9.0 * a * b + 27.0 * c) / 54.0;
  const double r2 = r * r, q3 = q * q * q;
  double x;
  if (r2 < q3) {
    const double t = acos(fmin(1.0, fmax(-1.0, r / sqrt(q3))));
    const double m = -2.0 * sqrt(q);
    const double x0 = m * cos(t / 3.0) - a / 3.0;
    const double x1 = m * cos((t + 2.0 * 3.14159265358979323846) / 3.0) - a / 3.0;
    const double x2 = m * cos((t - 2.0 * 3.14159265358979323846) / 3.0) - a / 3.0;
    x = fmax(x0, fmax(x1, x2));
  } else {
    const double A = -copysign(cbrt(fabs(r) + sqrt(r2 - q3)), r);
    const double B = (A != 0.0) ? q / A : 0.0;
    x = (A + B) - a / 3.0;
  }
  // two Newton steps
  for (int i = 0; i < 2; ++i) {
    const double f = ((x + a) * x + b) * x + c, d = (3.0 * x + 2.0 * a) * x + b;
    if (d != 0.0) x -= f / d;
  }
  return x;
}

// Real roots of k[0] x^4 + k[1] x^3 + k[2] x^2 + k[3] x + k[4]; returns their number (<= 4).
// Ferrari through the resolvent cubic, each root polished by Newton on the original polynomial.
__device__ inline int quartic_real_roots(const double* k, double* roots) {
  double scale = fmax(fmax(fabs(k[0]), fabs(k[1])), fmax(fabs(k[2]), fmax(fabs(k[3]), fabs(k[4]))));
  if (!(scale > 0.0)) return 0;
  int n = 0;
  if (fabs(k[0]) < 1e-14 * scale) return 0;  // degenerate sample: no hypothesis
  const double a = k[1] / k[0], b = k[2] / k[0], c = k[3] / k[0], d = k[4] / k[0];
  const double a2 = a * a;
  const double p = b - 0.375 * a2;
  const double q = c - 0.5 * a * b + 0.125 * a2 * a;
  const double r = d - 0.25 * a * c + 0.0625 * a2 * b - (3.0 / 256.0) * a2 * a2;
  double y[4];
  const double tol = 1e-11 * (1.0 + fabs(p) * fabs(p) + fabs(r));
  if (fabs(q) < 1e-14 * (1.0 + fabs(p) + fabs(r))) {
    double disc = p * p - 4.0 * r;
    if (disc < 0.0 && disc > -tol) disc = 0.0;
    if (disc >= 0.0) {
      const double sd = sqrt(disc);
      const double z[2] = {0.5 * (-p + sd), 0.5 * (-p - sd)};
      for (int i = 0; i < 2; ++i) {
        if (z[i] >= 0.0) { const double s = sqrt(z[i]); y[n++] = s; y[n++] = -s; }
      }
    }
  } else {
    // 8 m^3 + 8 p m^2 + (2 p^2 - 8 r) m - q^2 = 0  ->  m^3 + p m^2 + (p^2/4 - r) m - q^2/8 = 0
    const double m = cubic_largest_real_root(p, 0.25 * p * p - r, -0.125 * q * q);
    if (m > 0.0) {
      const double s = sqrt(2.0 * m);
      const double h = 0.5 * p + m, g = q / (2.0 * s);
      // y^2 + s y + (h - g) = 0   and   y^2 - s y + (h + g) = 0
      const double cs[2] = {h - g, h + g}, ss[2] = {s, -s};
      for (int i = 0; i < 2; ++i) {
        double disc = ss[i] * ss[i] - 4.0 * cs[i];
        if (disc < 0.0 && disc > -tol) disc = 0.0;
        if (disc >= 0.0) {
          const double sd = sqrt(disc);
          const double t = -0.5 * (ss[i] + copysign(sd, ss[i]));  // stable quadratic formula
          y[n++] = t;
          y[n++] = (t != 0.0) ? cs[i] / t : 0.0;
        }
      }
    }
  }
  for (int i = 0; i < n; ++i) {
    double x = y[i] - 0.25 * a;
    for (int it = 0; it < 3; ++it) {
      const double f = (((k[0] * x + k[1]) * x + k[2]) * x + k[3]) * x + k[4];
      const double df = ((4.0 * k[0] * x + 3.0 * k[1]) * x + 2.0 * k[2]) * x + k[3];
      if (df != 0.0 && isfinite(f / df)) x -= f / df;
    }
    roots[i] = x;
  }
  return n;
}

__device__ __forceinline__ void cross3(const double* a, const double* b, double* o) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}

// Orthonormal frame (columns e1, e3 x e1, e3) of a triangle; false when degenerate.
__device__ inline bool tri_frame(const double (*A)[3], double* F /* 3x3 row-major, columns = axes */) {
  double e1[3] = {A[1][0] - A[0][0], A[1][1] - A[0][1], A[1][2] - A[0][2]};
  const double d2[3] = {A[2][0] - A[0][0], A[2][1] - A[0][1], A[2][2] - A[0][2]};
  double e3[3], e2[3];
  cross3(e1, d2, e3);
  const double n1 = sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]);
  const double n3 = sqrt(e3[0] * e3[0] + e3[1] * e3[1] + e3[2] * e3[2]);
  if (!(n1 > 1e-300) || !(n3 > 1e-300)) return false;
  for (int i = 0; i < 3; ++i) { e1[i] /= n1; e3[i] /= n3; }
  cross3(e3, e1, e2);
  for (int i = 0; i < 3; ++i) { F[3 * i] = e1[i]; F[3 * i + 1] = e2[i]; F[3 * i + 2] = e3[i]; }
  return true;
}

// P3P on the first three correspondences, the fourth selects among the (<= 4) solutions by
// reprojection error.  X4 world points, uv4 pixels.  Writes [R|t] row-major 3x4; false if none.
__device__ inline bool p3p_plus1(const double (*X4)[3], const double (*uv4)[2], const double4 K,
                                 double* Rt) {
  double f[3][3];
  for (int i = 0; i < 3; ++i) {
    const double x = (uv4[i][0] - K.z) / K.x, y = (uv4[i][1] - K.w) / K.y;
    const double inv = 1.0 / sqrt(x * x + y * y + 1.0);
    f[i][0] = x * inv; f[i][1] = y * inv; f[i][2] = inv;
  }
  auto d2 = [&](int i, int j) {
    const double a = X4[i][0] - X4[j][0], b = X4[i][1] - X4[j][1], c = X4[i][2] - X4[j][2];
    return a * a + b * b + c * c;
  };
  auto dot = [&](int i, int j) { return f[i][0] * f[j][0] + f[i][1] * f[j][1] + f[i][2] * f[j][2]; };
  const double a2 = d2(1, 2), b2 = d2(0, 2), c2 = d2(0, 1);
  const double ca = dot(1, 2), cb = dot(0, 2), cg = dot(0, 1);
  // polynomials in v, highest power first:  q = v^2 - 2 cb v + 1, N = (a2-c2) q - b2 (v^2 - 1),
  // D = 2 b2 (cg - ca v);  quartic = b2 (D^2 + N^2 - 2 cg N D) - c2 q D^2
  const double q[3] = {1.0, -2.0 * cb, 1.0};
  const double N[3] = {(a2 - c2) - b2, (a2 - c2) * q[1], (a2 - c2) + b2};
  const double D[2] = {-2.0 * b2 * ca, 2.0 * b2 * cg};
  const double DD[3] = {D[0] * D[0], 2.0 * D[0] * D[1], D[1] * D[1]};
  double NN[5] = {0, 0, 0, 0, 0}, ND[4] = {0, 0, 0, 0}, qDD[5] = {0, 0, 0, 0, 0};
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { NN[i + j] += N[i] * N[j]; qDD[i + j] += q[i] * DD[j]; }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 2; ++j) ND[i + j] += N[i] * D[j];
  double k[5];
  for (int i = 0; i < 5; ++i) {
    double v = NN[i];
    if (i >= 2) v += DD[i - 2];
    if (i >= 1) v -= 2.0 * cg * ND[i - 1];
    k[i] = b2 * v - c2 * qDD[i];
  }
  double roots[4];
  const int nr = quartic_real_roots(k, roots);
  double Fw[9];
  if (!tri_frame(X4, Fw)) return false;
  bool found = false;
  double best_e = 1e300;
  for (int ri = 0; ri < nr; ++ri) {
    const double v = roots[ri];
    const double Dv = D[0] * v + D[1];
    if (!(fabs(Dv) > 1e-12)) continue;
    const double u = ((N[0] * v + N[1]) * v + N[2]) / Dv;
    const double den = 1.0 + v * v - 2.0 * v * cb;
    if (!(den > 0.0) || !(u > 0.0) || !(v > 0.0)) continue;
    const double s1 = sqrt(b2 / den), s2 = u * s1, s3 = v * s1;
    double C[3][3];
    for (int j = 0; j < 3; ++j) { C[0][j] = s1 * f[0][j]; C[1][j] = s2 * f[1][j]; C[2][j] = s3 * f[2][j]; }
    double Fc[9];
    if (!tri_frame(C, Fc)) continue;
    double R[9];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        R[3 * i + j] = Fc[3 * i] * Fw[3 * j] + Fc[3 * i + 1] * Fw[3 * j + 1] + Fc[3 * i + 2] * Fw[3 * j + 2];
    double t[3];
    for (int i = 0; i < 3; ++i)
      t[i] = C[0][i] - (R[3 * i] * X4[0][0] + R[3 * i + 1] * X4[0][1] + R[3 * i + 2] * X4[0][2]);
    double pc[3];
    for (int i = 0; i < 3; ++i)
      pc[i] = R[3 * i] * X4[3][0] + R[3 * i + 1] * X4[3][1] + R[3 * i + 2] * X4[3][2] + t[i];
    if (!(pc[2] > 0.0)) continue;
    const double eu = K.x * pc[0] / pc[2] + K.z - uv4[3][0], ev = K.y * pc[1] / pc[2] + K.w - uv4[3][1];
    const double e = eu * eu + ev * ev;
    if (e < best_e && isfinite(e)) {
      best_e = e;
      found = true;
      for (int i = 0; i < 3; ++i) {
        Rt[4 * i] = R[3 * i]; Rt[4 * i + 1] = R[3 * i + 1]; Rt[4 * i + 2] = R[3 * i + 2]; Rt[4 * i + 3] = t[i];
      }
    }
  }
  return found;
}

// cv2 float semantics of PnPRansacCallback::computeError: double projection rounded to float32,
// float32 difference, float32 sum of squares without FMA contraction.
__device__ __forceinline__ float reproj_err_f32(const double* Rt, float X, float Y, float Z, float u,
                                                float v, const double4 K) {
  const double x = Rt[0] * X + Rt[1] * Y + Rt[2] * Z + Rt[3];
  const double y = Rt[4] * X + Rt[5] * Y + Rt[6] * Z + Rt[7];
  double z = Rt[8] * X + Rt[9] * Y + Rt[10] * Z + Rt[11];
  z = (z != 0.0) ? 1.0 / z : 1.0;
  const float pu = (float)(__dadd_rn(__dmul_rn(__dmul_rn(x, z), K.x), K.z));
  const float pv = (float)(__dadd_rn(__dmul_rn(__dmul_rn(y, z), K.y), K.w));
  const float du = __fsub_rn(u, pu), dv = __fsub_rn(v, pv);
  return __fadd_rn(__fmul_rn(du, du), __fmul_rn(dv, dv));
}

}  // namespace vslam
