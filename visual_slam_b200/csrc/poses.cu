// Pose plumbing around the PnP call of the tracking loop (reference src/v2/main.py:191-204 and
// src/v2/helper_functions.py:269-278): cv2.Rodrigues both ways, [R|t] -> camera->world inversion, the
// Stage-B guess from the previous frame's camera->world pose, and the (q, t) form bundle adjustment consumes
// (LocalBA.py:57-58), batched, one thread per pose, so that tracking results feed the map / BA arrays without
// a host round trip.  Follows OpenCV's cv::Rodrigues operation by operation (the matrix -> vector direction
// first projects the input on SO(3), U V^T of its SVD, here by the Newton iteration for the polar factor).
#include "common.cuh"

namespace vslam {

// cv::Rodrigues, vector -> matrix: theta < DBL_EPSILON -> I; else c I + (1 - c) r r^T + s [r]x with r / theta.
__device__ __forceinline__ void cv_rodrigues(const double* rv, double* R) {
  const double theta = sqrt(rv[0] * rv[0] + rv[1] * rv[1] + rv[2] * rv[2]);
  if (theta < 2.220446049250313e-16) {
    R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
    return;
  }
  double s, c;
  sincos(theta, &s, &c);
  const double c1 = 1.0 - c, it = 1.0 / theta;
  const double x = rv[0] * it, y = rv[1] * it, z = rv[2] * it;
  R[0] = c + c1 * (x * x);          R[1] = c1 * (x * y) + s * (-z);  R[2] = c1 * (x * z) + s * y;
  R[3] = c1 * (x * y) + s * z;      R[4] = c + c1 * (y * y);         R[5] = c1 * (y * z) + s * (-x);
  R[6] = c1 * (x * z) + s * (-y);   R[7] = c1 * (y * z) + s * x;     R[8] = c + c1 * (z * z);
}

__device__ __forceinline__ void inv3_transposed(const double* A, double* o) {   // o = A^-T
  const double c0 = A[4] * A[8] - A[5] * A[7], c1 = A[5] * A[6] - A[3] * A[8], c2 = A[3] * A[7] - A[4] * A[6];
  const double id = 1.0 / (A[0] * c0 + A[1] * c1 + A[2] * c2);
  o[0] = c0 * id; o[1] = c1 * id; o[2] = c2 * id;
  o[3] = (A[2] * A[7] - A[1] * A[8]) * id; o[4] = (A[0] * A[8] - A[2] * A[6]) * id; o[5] = (A[1] * A[6] - A[0] * A[7]) * id;
  o[6] = (A[1] * A[5] - A[2] * A[4]) * id; o[7] = (A[2] * A[3] - A[0] * A[5]) * id; o[8] = (A[0] * A[4] - A[1] * A[3]) * id;
}

// Orthogonal polar factor of a nonsingular 3x3 matrix (= U V^T of its SVD): X <- (X + X^-T) / 2, quadratic.
__device__ inline void polar_factor(const double* A, double* X) {
#pragma unroll
  for (int i = 0; i < 9; ++i) X[i] = A[i];
  for (int it = 0; it < 30; ++it) {
    double Y[9], d = 0.0;
    inv3_transposed(X, Y);
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      const double n = 0.5 * (X[i] + Y[i]);
      d = fmax(d, fabs(n - X[i]));
      X[i] = n;
    }
    if (!(d > 1e-16)) break;      // converged (also leaves on NaN: a singular input stays NaN, cv2 would return garbage too)
  }
}

// cv::Rodrigues, matrix -> vector.
__device__ inline void cv_rodrigues_inv(const double* Rin, double* rv) {
  double R[9];
  polar_factor(Rin, R);
  double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
  const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
  double c = (R[0] + R[4] + R[8] - 1.0) * 0.5;
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  double theta = acos(c);
  if (s < 1e-5) {
    if (c > 0) { rv[0] = rv[1] = rv[2] = 0.0; return; }
    double t = (R[0] + 1.0) * 0.5;
    rx = sqrt(fmax(t, 0.0));
    t = (R[4] + 1.0) * 0.5;
    ry = sqrt(fmax(t, 0.0)) * (R[1] < 0 ? -1.0 : 1.0);
    t = (R[8] + 1.0) * 0.5;
    rz = sqrt(fmax(t, 0.0)) * (R[2] < 0 ? -1.0 : 1.0);
    if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0) != (ry * rz > 0)) rz = -rz;
    theta /= sqrt(rx * rx + ry * ry + rz * rz);
    rv[0] = rx * theta; rv[1] = ry * theta; rv[2] = rz * theta;
    return;
  }
  const double vth = (1.0 / (2.0 * s)) * theta;
  rv[0] = rx * vth; rv[1] = ry * vth; rv[2] = rz * vth;
}

// Eigen's matrix -> quaternion followed by the SBACam constructor (normalised, w >= 0): (x, y, z, w).
__device__ inline void R_to_unit_quat(const double* R, double* q) {
  const double tr = R[0] + R[4] + R[8];
  if (tr > 0) {
    double s = sqrt(tr + 1.0);
    q[3] = 0.5 * s;
    s = 0.5 / s;
    q[0] = (R[7] - R[5]) * s; q[1] = (R[2] - R[6]) * s; q[2] = (R[3] - R[1]) * s;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (i + 2) % 3;
    double s = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    q[i] = 0.5 * s;
    s = 0.5 / s;
    q[3] = (R[3 * k + j] - R[3 * j + k]) * s;
    q[j] = (R[3 * j + i] + R[3 * i + j]) * s;
    q[k] = (R[3 * k + i] + R[3 * i + k]) * s;
  }
  const double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const double sg = q[3] < 0 ? -1.0 : 1.0;
#pragma unroll
  for (int a = 0; a < 4; ++a) q[a] = (sg * q[a]) / n;
}

__global__ void rodrigues_kernel(const double* __restrict__ rvec, double* __restrict__ R, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double r[9];
  cv_rodrigues(rvec + 3 * (size_t)i, r);
#pragma unroll
  for (int k = 0; k < 9; ++k) R[9 * (size_t)i + k] = r[k];
}

__global__ void rodrigues_inv_kernel(const double* __restrict__ R, double* __restrict__ rvec, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double m[9], r[3];
#pragma unroll
  for (int k = 0; k < 9; ++k) m[k] = R[9 * (size_t)i + k];
  cv_rodrigues_inv(m, r);
  rvec[3 * (size_t)i] = r[0]; rvec[3 * (size_t)i + 1] = r[1]; rvec[3 * (size_t)i + 2] = r[2];
}

// main.py:201-203: T = [Rodrigues(rvec) | tvec] (world -> camera), W_T_curr = [R^T | -R^T t] (camera -> world).
__global__ void pose_camera_to_world_kernel(const double* __restrict__ pose6, double* __restrict__ w2c,
                                            double* __restrict__ c2w, double* __restrict__ cam7, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* p = pose6 + 6 * (size_t)i;
  double R[9];
  cv_rodrigues(p, R);
  const double t[3] = {p[3], p[4], p[5]};
  double Rt[9], ti[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) Rt[3 * r + c] = R[3 * c + r];
  }
#pragma unroll
  for (int r = 0; r < 3; ++r) ti[r] = -(Rt[3 * r] * t[0] + Rt[3 * r + 1] * t[1] + Rt[3 * r + 2] * t[2]);
  if (w2c) {
    double* o = w2c + 16 * (size_t)i;
#pragma unroll
    for (int r = 0; r < 3; ++r) { o[4 * r] = R[3 * r]; o[4 * r + 1] = R[3 * r + 1]; o[4 * r + 2] = R[3 * r + 2]; o[4 * r + 3] = t[r]; }
    o[12] = 0; o[13] = 0; o[14] = 0; o[15] = 1;
  }
  if (c2w) {
    double* o = c2w + 16 * (size_t)i;
#pragma unroll
    for (int r = 0; r < 3; ++r) { o[4 * r] = Rt[3 * r]; o[4 * r + 1] = Rt[3 * r + 1]; o[4 * r + 2] = Rt[3 * r + 2]; o[4 * r + 3] = ti[r]; }
    o[12] = 0; o[13] = 0; o[14] = 0; o[15] = 1;
  }
  if (cam7) {
    double q[4];
    R_to_unit_quat(Rt, q);
    double* o = cam7 + 7 * (size_t)i;
    o[0] = q[0]; o[1] = q[1]; o[2] = q[2]; o[3] = q[3]; o[4] = ti[0]; o[5] = ti[1]; o[6] = ti[2];
  }
}

// main.py:191-194: rvec_guess = Rodrigues(R of the previous frame's camera->world pose), tvec_guess = its translation.
__global__ void pose_guess_kernel(const double* __restrict__ c2w, double* __restrict__ pose6, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* T = c2w + 16 * (size_t)i;
  const double R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
  double r[3];
  cv_rodrigues_inv(R, r);
  double* o = pose6 + 6 * (size_t)i;
  o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = T[3]; o[4] = T[7]; o[5] = T[11];
}

__global__ void pose_guess_cam7_kernel(const double* __restrict__ cam7, double* __restrict__ pose6, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* c = cam7 + 7 * (size_t)i;
  const double x = c[0], y = c[1], z = c[2], w = c[3];
  const double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                       2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                       2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)};
  double r[3];
  cv_rodrigues_inv(R, r);
  double* o = pose6 + 6 * (size_t)i;
  o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = c[4]; o[4] = c[5]; o[5] = c[6];
}

}  // namespace vslam

using namespace vslam;

extern "C" int vslam_pose_guess_cam7(const double* cam7, double* pose6, int batch, void* stream) {
  if (batch < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!cam7 || !pose6) return VSLAM_ERR_BAD_ARG;
  pose_guess_cam7_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(cam7, pose6, batch);
  return check_launch("vslam_pose_guess_cam7");
}

extern "C" int vslam_rodrigues(const double* rvec, double* R, int batch, void* stream) {
  if (batch < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!rvec || !R) return VSLAM_ERR_BAD_ARG;
  rodrigues_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(rvec, R, batch);
  return check_launch("vslam_rodrigues");
}

extern "C" int vslam_rodrigues_inverse(const double* R, double* rvec, int batch, void* stream) {
  if (batch < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!rvec || !R) return VSLAM_ERR_BAD_ARG;
  rodrigues_inv_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(R, rvec, batch);
  return check_launch("vslam_rodrigues_inverse");
}

extern "C" int vslam_pose_camera_to_world(const double* pose6, double* w2c, double* c2w, double* cam7, int batch,
                                          void* stream) {
  if (batch < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!pose6 || (!w2c && !c2w && !cam7)) return VSLAM_ERR_BAD_ARG;
  pose_camera_to_world_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(pose6, w2c, c2w,
                                                                                                  cam7, batch);
  return check_launch("vslam_pose_camera_to_world");
}

extern "C" int vslam_pose_guess(const double* c2w, double* pose6, int batch, void* stream) {
  if (batch < 0) return VSLAM_ERR_BAD_ARG;
  if (batch == 0) return VSLAM_OK;
  if (!c2w || !pose6) return VSLAM_ERR_BAD_ARG;
  pose_guess_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(c2w, pose6, batch);
  return check_launch("vslam_pose_guess");
}
