// Batched two-view triangulation + cheirality/depth filter.
// Replaces the per-point Python loop of reference src/v2/helper_functions.py:281-291 (`triangulate`:
// 4x4 DLT matrix, np.linalg.svd, last right singular vector) and the filter that follows it at
// src/v2/main.py:284-306 (de-homogenise, depth of the point in both cameras, 0 < z < 1).
// One thread per correspondence: the 4x4 system is built in registers and its right singular vectors
// come from a one-sided Jacobi SVD (Hestenes rotations on the columns, V accumulated), which keeps
// the accuracy of an SVD of A (no squaring of the condition number as with eig(A^T A)).  FP64.
#include "common.cuh"

namespace vslam {

// One-sided Jacobi on the 4 columns of A (a[c][r]); returns in v the right singular vector of the
// smallest singular value, normalised, sign fixed so that the largest component is positive.
__device__ __forceinline__ void smallest_right_singular_vector(double (&a)[4][4], double (&v)[4]) {
  double V[4][4];   // V[c][r]: column c of V
#pragma unroll
  for (int c = 0; c < 4; ++c) {
#pragma unroll
    for (int r = 0; r < 4; ++r) V[c][r] = (r == c) ? 1.0 : 0.0;
  }
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double alpha = 0.0, beta = 0.0, gamma = 0.0;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          alpha += a[p][r] * a[p][r];
          beta += a[q][r] * a[q][r];
          gamma += a[p][r] * a[q][r];
        }
        const double lim = 1e-15 * sqrt(alpha * beta);
        if (fabs(gamma) > lim && fabs(gamma) > 1e-300) {
          off = fmax(off, fabs(gamma) / fmax(sqrt(alpha * beta), 1e-300));
          const double zeta = (beta - alpha) / (2.0 * gamma);
          const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const double ap = a[p][r], aq = a[q][r];
            a[p][r] = cs * ap - sn * aq;
            a[q][r] = sn * ap + cs * aq;
            const double vp = V[p][r], vq = V[q][r];
            V[p][r] = cs * vp - sn * vq;
            V[q][r] = sn * vp + cs * vq;
          }
        }
      }
    }
    if (off < 1e-15) break;
  }
  int best = 0;
  double nb = 1.7976931348623157e308;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    double n2 = 0.0;
#pragma unroll
    for (int r = 0; r < 4; ++r) n2 += a[c][r] * a[c][r];
    if (n2 < nb) { nb = n2; best = c; }
  }
  double vv[4], nrm = 0.0;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    vv[r] = best == 0 ? V[0][r] : best == 1 ? V[1][r] : best == 2 ? V[2][r] : V[3][r];
    nrm += vv[r] * vv[r];
  }
  nrm = 1.0 / sqrt(nrm);
  int big = 0;
#pragma unroll
  for (int r = 1; r < 4; ++r) if (fabs(vv[r]) > fabs(vv[big])) big = r;
  const double s = (big == 0 ? vv[0] : big == 1 ? vv[1] : big == 2 ? vv[2] : vv[3]) < 0.0 ? -nrm : nrm;
#pragma unroll
  for (int r = 0; r < 4; ++r) v[r] = vv[r] * s;
}

// proj   [batch][2][12]  3x4 projection matrices of the two views (row-major), as passed to `triangulate`
// extr   [batch][2][12]  3x4 world->camera matrices used for the depth test (may be null: no filter)
// x1,x2  [batch][n_max][2] pixel coordinates;  n_pts [batch]
// Xh     [batch][n_max][4] unit homogeneous points (vt[3] up to sign);  X3 [batch][n_max][3] = Xh[:3]/Xh[3]
// depth  [batch][n_max][2] z of X3 in both cameras;  good u8 [batch][n_max]: 0 < z1 < zmax && 0 < z2 < zmax
__global__ void __launch_bounds__(128)
triangulate_kernel(const double* __restrict__ proj, const double* __restrict__ extr,
                   const double* __restrict__ x1, const double* __restrict__ x2,
                   const int32_t* __restrict__ n_pts, int n_max, double z_max,
                   double* __restrict__ Xh, double* __restrict__ X3, double* __restrict__ depth,
                   uint8_t* __restrict__ good) {
  __shared__ double P[24], T[24];
  const int b = blockIdx.y;
  if (threadIdx.x < 24) {
    P[threadIdx.x] = proj[24 * (size_t)b + threadIdx.x];
    T[threadIdx.x] = extr ? extr[24 * (size_t)b + threadIdx.x] : 0.0;
  }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= min(n_pts[b], n_max)) return;
  const size_t o = (size_t)b * n_max + i;
  const double u1 = x1[2 * o], v1 = x1[2 * o + 1], u2 = x2[2 * o], v2 = x2[2 * o + 1];
  // rows of A (helper_functions.py:285-288): u * P[2] - P[0], v * P[2] - P[1] for both views;
  // stored by column: a[c][r]
  double a[4][4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    a[c][0] = u1 * P[8 + c] - P[c];
    a[c][1] = v1 * P[8 + c] - P[4 + c];
    a[c][2] = u2 * P[12 + 8 + c] - P[12 + c];
    a[c][3] = v2 * P[12 + 8 + c] - P[12 + 4 + c];
  }
  double v[4];
  smallest_right_singular_vector(a, v);
#pragma unroll
  for (int r = 0; r < 4; ++r) Xh[4 * o + r] = v[r];
  const double iw = 1.0 / v[3];
  const double X[3] = {v[0] * iw, v[1] * iw, v[2] * iw};
  if (X3) { X3[3 * o] = X[0]; X3[3 * o + 1] = X[1]; X3[3 * o + 2] = X[2]; }
  if (extr) {
    const double z1 = T[8] * X[0] + T[9] * X[1] + T[10] * X[2] + T[11];
    const double z2 = T[12 + 8] * X[0] + T[12 + 9] * X[1] + T[12 + 10] * X[2] + T[12 + 11];
    if (depth) { depth[2 * o] = z1; depth[2 * o + 1] = z2; }
    if (good) good[o] = (z1 > 0.0 && z2 > 0.0 && z1 < z_max && z2 < z_max) ? 1 : 0;
  }
}

}  // namespace vslam

extern "C" int vslam_triangulate(const double* proj, const double* extr, const double* x1, const double* x2,
                                 const int32_t* n_pts, int n_max, int batch, double z_max, double* Xh,
                                 double* X3, double* depth, uint8_t* good, void* stream) {
  if (n_max < 0 || batch < 0) return VSLAM_ERR_BAD_ARG;
  if (n_max == 0 || batch == 0) return VSLAM_OK;   // nothing to do: the reference returns an empty (0, 4) array
  if (!proj || !x1 || !x2 || !n_pts || !Xh) return VSLAM_ERR_BAD_ARG;
  if (!extr && (depth || good)) return VSLAM_ERR_BAD_ARG;
  dim3 grid((n_max + 127) / 128, batch);
  vslam::triangulate_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      proj, extr, x1, x2, n_pts, n_max, z_max, Xh, X3, depth, good);
  return vslam::check_launch("vslam_triangulate");
}
