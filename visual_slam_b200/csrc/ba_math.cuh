// Per-observation arithmetic of the SBA edge (g2o EdgeProjectP2MC / SBACam semantics restated in
// SURVEY.md 8a-3; reference call sites src/v2/LocalBA.py:56-94).  All FP64.
#pragma once
#include "common.cuh"

namespace vslam {

constexpr double kHuberDelta2 = 5.991;  // LocalBA.py:82: RobustKernelHuber(sqrt(5.991))

// Camera record prepared once per state: R (camera->world, row-major 9) + t (3).
struct CamRT {
  double R[9];
  double t[3];
};

__device__ __forceinline__ void quat_to_R(const double* q, double* R) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - z * w);     R[2] = 2 * (x * z + y * w);
  R[3] = 2 * (x * y + z * w);     R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - x * w);
  R[6] = 2 * (x * z - y * w);     R[7] = 2 * (y * z + x * w);     R[8] = 1 - 2 * (x * x + y * y);
}

// pc = R^T (X - t)   (w2n applied to the homogeneous point)
__device__ __forceinline__ void to_camera(const CamRT& c, const double* X, double* pc) {
  const double dx = X[0] - c.t[0], dy = X[1] - c.t[1], dz = X[2] - c.t[2];
  pc[0] = c.R[0] * dx + c.R[3] * dy + c.R[6] * dz;
  pc[1] = c.R[1] * dx + c.R[4] * dy + c.R[7] * dz;
  pc[2] = c.R[2] * dx + c.R[5] * dy + c.R[8] * dz;
}

// e = projected - measured  (EdgeProjectP2MC::computeError)
__device__ __forceinline__ void residual(const double* pc, const double4 K, double u, double v,
                                         double* e) {
  const double iz = 1.0 / pc[2];
  e[0] = K.x * pc[0] * iz + K.z - u;
  e[1] = K.y * pc[1] * iz + K.w - v;
}

// Huber IRLS weight and rho (robust_kernel_impl: rho[0], rho[1]); Omega = I2 (LocalBA.py:81).
__device__ __forceinline__ void huber(double chi2, double& rho, double& w) {
  if (chi2 <= kHuberDelta2) {
    rho = chi2;
    w = 1.0;
  } else {
    const double s = sqrt(chi2), d = sqrt(kHuberDelta2);
    rho = 2.0 * s * d - kHuberDelta2;
    w = d / s;
  }
}

// DCS with phi = 1 (RobustKernelDCS default, LocalBA.py:115)
__device__ __forceinline__ void dcs(double chi2, double& rho, double& w) {
  const double s = 2.0 / (1.0 + chi2);
  if (s >= 1.0) { rho = chi2; w = 1.0; }
  else { rho = s * s * chi2; w = s * s; }
}

// Analytic Jacobians (EdgeProjectP2MC::linearizeOplus).
//   Jp[2][3]: d e / d X;  Jc[2][6]: columns 0..2 translation (= -Jp), 3..5 rotation.
// g(dp) = ((pz dp0 - px dp2) ipz2 fx, (pz dp1 - py dp2) ipz2 fy);  dRdk * (X - t) = A_k * pc.
__device__ __forceinline__ void jacobians(const CamRT& c, const double* pc, const double4 K,
                                          double (&Jc)[2][6], double (&Jp)[2][3]) {
  const double px = pc[0], py = pc[1], pz = pc[2];
  const double ipz2 = 1.0 / (pz * pz);
  const double a = ipz2 * K.x, b = ipz2 * K.y;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    // w2n column k = row k of R
    const double d0 = c.R[3 * k + 0], d1 = c.R[3 * k + 1], d2 = c.R[3 * k + 2];
    Jp[0][k] = (pz * d0 - px * d2) * a;
    Jp[1][k] = (pz * d1 - py * d2) * b;
    Jc[0][k] = -Jp[0][k];
    Jc[1][k] = -Jp[1][k];
  }
  // A_x pc = (0, 2pz, -2py); A_y pc = (-2pz, 0, 2px); A_z pc = (2py, -2px, 0)
  Jc[0][3] = (pz * 0.0 - px * (-2.0 * py)) * a;
  Jc[1][3] = (pz * (2.0 * pz) - py * (-2.0 * py)) * b;
  Jc[0][4] = (pz * (-2.0 * pz) - px * (2.0 * px)) * a;
  Jc[1][4] = (pz * 0.0 - py * (2.0 * px)) * b;
  Jc[0][5] = (pz * (2.0 * py)) * a;
  Jc[1][5] = (pz * (-2.0 * px)) * b;
}

// Inverse of a symmetric 3x3 given as (xx, xy, xz, yy, yz, zz); same storage out.
__device__ __forceinline__ void sym3_inverse(const double* h, double* o) {
  const double a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5];
  const double c00 = d * f - e * e, c01 = c * e - b * f, c02 = b * e - c * d;
  const double det = a * c00 + b * c01 + c * c02;
  const double id = 1.0 / det;
  o[0] = c00 * id; o[1] = c01 * id; o[2] = c02 * id;
  o[3] = (a * f - c * c) * id; o[4] = (b * c - a * e) * id; o[5] = (a * d - b * b) * id;
}

__device__ __forceinline__ void sym3_mul(const double* h, const double* v, double* o) {
  o[0] = h[0] * v[0] + h[1] * v[1] + h[2] * v[2];
  o[1] = h[1] * v[0] + h[3] * v[1] + h[4] * v[2];
  o[2] = h[2] * v[0] + h[4] * v[1] + h[5] * v[2];
}

// Deterministic block reduction of NV doubles per thread: result valid in thread 0.
template <int NV, int THREADS>
__device__ __forceinline__ void block_reduce(double (&v)[NV], double* smem /* [THREADS/32][NV] */) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) smem[wid * NV + i] = v[i];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      double s = 0.0;
      for (int w = 0; w < THREADS / 32; ++w) s += smem[w * NV + i];
      v[i] = s;
    }
  }
  __syncthreads();
}

}  // namespace vslam
