/*
 * vslam_b200.h -- C ABI of libvslam_b200.so (sm_100a only).
 *
 * This is the drop-in boundary for the match -> RANSAC pose -> bundle adjustment
 * hot path of juuso-oskari/visual_slam src/v2.  The reference has no FFI layer of
 * its own (it is pure Python over the cv2 and g2o wheels), so every entry point
 * below cites the reference call site whose third-party native call it replaces.
 *
 * Conventions
 *   - every pointer marked "dev" is a DEVICE pointer (HBM); the caller owns all
 *     buffers; the library allocates nothing on the data path except the opaque
 *     BA problem object (vslam_ba_create / vslam_ba_destroy).
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).
 *   - every function returns a vslam_status; 0 is success.  Nothing falls back
 *     to the CPU: without a B200 the calls fail with VSLAM_ERR_CUDA.
 */
#ifndef VSLAM_B200_H
#define VSLAM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  VSLAM_OK = 0,
  VSLAM_ERR_BAD_ARG = 1,    /* null pointer, negative size, size over a documented limit */
  VSLAM_ERR_ALIGNMENT = 2,  /* descriptor buffers must be 16-byte aligned */
  VSLAM_ERR_CUDA = 3,       /* a CUDA runtime call failed; see vslam_last_cuda_error() */
  VSLAM_ERR_NUMERIC = 4,    /* solver breakdown reported in-band (PCG indefinite, ...) */
  VSLAM_ERR_NO_DEVICE = 5,
  VSLAM_ERR_TOO_LARGE = 6   /* a no-allocation entry point was given a size beyond its static scratch; see vslam_last_cuda_error() */
} vslam_status;

const char* vslam_status_string(int status);
const char* vslam_last_cuda_error(void);
/* ABI version of this header; bumped on any signature change. */
int vslam_abi_version(void);
/* Number of SMs of the current device (148 on B200), or <0 on error. */
int vslam_device_sm_count(void);

/* ------------------------------------------------------------------------- */
/* 1. Descriptor matching                                                     */
/* ------------------------------------------------------------------------- */

/* Limits of the packed (distance, index) keys used by the kNN kernels. */
#define VSLAM_HAMMING_MAX_N (1 << 22)
#define VSLAM_DESC_BYTES 32

/*
 * Brute-force 256-bit Hamming kNN, k = 2, for `batch` independent frame pairs.
 * Replaces cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(desc1, desc2, k=2)
 * (reference: src/v2/frame.py:23, ORB variant of frame.py:8/18).
 *
 *   q        dev  uint8 [batch][nq][32]   query descriptors (desc1)
 *   t        dev  uint8 [batch][nt][32]   train descriptors (desc2)
 *   idx      dev  int32 [batch][nq][2]    trainIdx of best / second best
 *   dist     dev  int32 [batch][nq][2]    their Hamming distances
 *   rev_key  dev  uint32[batch][nt] or NULL.  When non-NULL the same pass also
 *            produces, for every train row, the packed key (distance << 22 | q)
 *            of its nearest query with first-min tie-break (the cross-check
 *            direction, reference semantics src/slam.py:40-47).  The kernel
 *            initialises it; the caller need not.
 *
 * Order and ties are bit-exact with OpenCV: ascending (distance, trainIdx).
 * If nt < 2 the missing entries are idx = -1, dist = -1.
 * nq, nt <= VSLAM_HAMMING_MAX_N.
 */
int vslam_hamming_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt,
                       int32_t* idx, int32_t* dist, uint32_t* rev_key,
                       int batch, void* stream);

/*
 * The same for RAGGED batches (frames with different key-point counts, as vslam_extract_features produces
 * them): q / t are padded to nq_stride / nt_stride rows per pair, nq_arr / nt_arr (dev i32 [batch]) give the
 * number of real rows of each pair.  Rows beyond the count never match and are never matched: their idx / dist
 * entries are -1 (so vslam_match_filter drops them) and their rev_key entries stay 0xFFFFFFFF.  A pair with
 * fewer than 2 train rows yields -1 second entries exactly like nt < 2 above.
 */
int vslam_hamming_knn2_ragged(const uint8_t* q, int nq_stride, const int32_t* nq_arr, const uint8_t* t,
                              int nt_stride, const int32_t* nt_arr, int32_t* idx, int32_t* dist,
                              uint32_t* rev_key, int batch, void* stream);

/*
 * Lowe ratio test (+ optional mutual-nearest cross-check) and order-preserving
 * compaction.  Replaces the Python loop src/v2/frame.py:30-42.
 *
 *   idx, dist  dev  outputs of vslam_hamming_knn2
 *   rev_key    dev  as produced by vslam_hamming_knn2 (required iff cross_check)
 *   ratio      the reference compares float32 distances in float64:
 *              keep iff (double)d0 < ratio * (double)d1     (frame.py:33)
 *   keep_q     dev  int32 [batch][nq]  kept query indices, ascending, padded -1
 *   keep_t     dev  int32 [batch][nq]  their trainIdx,               padded -1
 *   n_keep     dev  int32 [batch]
 */
int vslam_match_filter(const int32_t* idx, const int32_t* dist,
                       const uint32_t* rev_key, int nq, int nt, double ratio,
                       int cross_check, int32_t* keep_q, int32_t* keep_t,
                       int32_t* n_keep, int batch, void* stream);

/*
 * Integer-pipe microbenchmark used for the matching roofline:
 * runs `iters` dependent-free popc.b32 per thread on a full grid and returns
 * the achieved popc/s in *popc_per_s (host pointer).
 */
int vslam_microbench_popc(int iters, double* popc_per_s, void* stream);
/*
 * The same for the pipe that actually binds the kNN kernel (ncu: ALU 83 % busy): dependent-free three-input logic
 * operations (lop3.b32) per second in *lop3_per_s.  The kernel issues 16 LOP3 + 4 min/max per 256-bit pair
 * distance (carry-save popcount tree), all on that pipe.
 */
int vslam_microbench_lop3(int iters, double* lop3_per_s, void* stream);


/* ------------------------------------------------------------------------- */
/* 2. RANSAC pose (PnP)                                                       */
/* ------------------------------------------------------------------------- */
/*
 * Together these replace cv2.solvePnPRansac(objectPoints, imagePoints, cameraMatrix, distCoeffs=[],
 * rvec, tvec, useExtrinsicGuess=True) as called at reference src/v2/main.py:196-197 (defaults:
 * 100 iterations, reprojectionError 8.0, confidence 0.99), batched over frame pairs.
 *
 *   X      dev f32 [batch][n_max][3]   object points (world), first n_pts[b] rows valid
 *   uv     dev f32 [batch][n_max][2]   image points (pixels)
 *   n_pts  dev i32 [batch]
 *   K      host f64[4]                 fx, fy, cx, cy (no distortion, as the reference passes [])
 *   words  dev u32 [batch][H][4]       harness-drawn random words; sample index j of hypothesis h is
 *                                      (word * n) >> 32 with duplicates stepped to the next free index
 *   hyps   dev f64 [batch][H][12]      world->camera [R|t], row-major 3x4
 *   valid  dev u8  [batch][H]
 */
/* Minimal solver: P3P on three correspondences, the fourth picks among the solutions. */
int vslam_pnp_hypotheses(const float* X, const float* uv, const int32_t* n_pts, int n_max, int batch,
                         const uint32_t* words, int H, const double* K, double* hyps, uint8_t* valid,
                         void* stream);
/* Inlier count of every hypothesis; inlier iff |uv - proj|^2 <= thr^2 with OpenCV's float32
 * semantics (PnPRansacCallback::computeError).  One warp per hypothesis.  counts dev i32 [batch][H]. */
int vslam_pnp_score(const float* X, const float* uv, const int32_t* n_pts, int n_max, int batch,
                    const double* hyps, const uint8_t* valid, int H, const double* K, double thr,
                    int32_t* counts, void* stream);
/*
 * Selection with OpenCV's shrinking iteration budget (RANSACUpdateNumIters), inlier mask of the
 * winner, then iterative refinement on the inliers from the caller's guess (cv2 "Stage B").
 *   guess, pose  dev f64 [batch][6]   rvec (3) + tvec (3), world->camera
 *   best         dev i32 [batch]      winning hypothesis or -1
 *   inlier_bits  dev u32 [batch][ceil(n_max/32)]
 *   n_inliers    dev i32 [batch];  status dev i32 [batch]: 0 ok, 1 no model, 2 refinement breakdown
 */
int vslam_pnp_select_refine(const float* X, const float* uv, const int32_t* n_pts, int n_max, int batch,
                            const double* hyps, const uint8_t* valid, const int32_t* counts, int H,
                            const double* K, double thr, double confidence, const double* guess,
                            int max_refine_iters, int32_t* best, uint32_t* inlier_bits,
                            int32_t* n_inliers, double* pose, int32_t* status, void* stream);

/*
 * Correspondence gather between matching and PnP (reference src/v2/main.py:187-188, frame.py:36-37):
 * X[b][i] = map_xyz[b][keep_q[b][i]], uv[b][i] = kp_cur[b][keep_t[b][i]] for i < n_keep[b], zero after.
 *   map_xyz dev f32 [batch][nq][3]  3-D map points of the query (key-frame) descriptors
 *   kp_cur  dev f32 [batch][nt][2]  keypoints of the train (current-frame) descriptors
 *   X dev f32 [batch][nq][3], uv dev f32 [batch][nq][2]  (n_max of the PnP calls = nq)
 */
int vslam_gather_correspondences(const int32_t* keep_q, const int32_t* keep_t, const int32_t* n_keep,
                                 int nq, int nt, int batch, const float* map_xyz, const float* kp_cur,
                                 float* X, float* uv, void* stream);

/* ------------------------------------------------------------------------- */
/* 2a. Pose plumbing around the PnP call                                      */
/* ------------------------------------------------------------------------- */
/*
 * cv2.Rodrigues both ways, batched (reference src/v2/helper_functions.py:269-278: transformMatrix builds
 * [Rodrigues(rvec) | tvec], Rtorvec is the matrix -> vector direction).  Operation order follows OpenCV's
 * cv::Rodrigues; the matrix -> vector direction first projects its input on SO(3) (U V^T of the SVD) as
 * OpenCV does.
 *   rvec dev f64 [batch][3];  R dev f64 [batch][9] row-major
 */
int vslam_rodrigues(const double* rvec, double* R, int batch, void* stream);
int vslam_rodrigues_inverse(const double* R, double* rvec, int batch, void* stream);
/*
 * reference src/v2/main.py:199-204: T = transformMatrix(rvec, tvec) (world -> camera) and its inverse
 * W_T_curr = [R^T | -R^T t] (camera -> world, what Frame.pose stores), plus the (qx, qy, qz, qw, tx, ty, tz)
 * form of W_T_curr that bundle adjustment consumes (SBACam(orientation, position), LocalBA.py:57-58), so that
 * vslam_pnp_select_refine's output feeds the map / BA arrays on the device.  Any of the three outputs may be
 * NULL (not all).
 *   pose6 dev f64 [batch][6] (rvec, tvec);  w2c, c2w dev f64 [batch][16] row-major 4x4;  cam7 dev f64 [batch][7]
 */
int vslam_pose_camera_to_world(const double* pose6, double* w2c, double* c2w, double* cam7, int batch,
                               void* stream);
/*
 * reference src/v2/main.py:191-194: the extrinsic guess handed to solvePnPRansac is the PREVIOUS frame's
 * camera -> world pose, rvec_guess = Rodrigues(W_T_prev[0:3, 0:3]), tvec_guess = W_T_prev[0:3, 3] (the
 * convention mismatch is the reference's and only seeds Stage B).
 *   c2w dev f64 [batch][16];  pose6 dev f64 [batch][6]
 */
int vslam_pose_guess(const double* c2w, double* pose6, int batch, void* stream);
/* The same from the (qx,qy,qz,qw,tx,ty,tz) form of the pose (the rows of the BA / tracking-window camera arrays). */
int vslam_pose_guess_cam7(const double* cam7, double* pose6, int batch, void* stream);

/* ------------------------------------------------------------------------- */
/* 2b. Two-view geometry (map initialisation)                                 */
/* ------------------------------------------------------------------------- */
/*
 * cv2.undistortPoints(pts, cameraMatrix=K, distCoeffs=None) as called at reference
 * src/v2/helper_functions.py:49-50: out = (float)(((double)u - cx) * (1/fx)), bit-exact.
 *   uv, out  dev f32 [n][2];  K host f64[4] = fx, fy, cx, cy
 */
int vslam_undistort_points(const float* uv, int64_t n, const double* K, float* out, void* stream);

/*
 * Together these replace cv2.findEssentialMat(pts1, pts2, method=RANSAC, prob=0.999, threshold=thr)
 * (reference src/v2/helper_functions.py:52; focal 1, pp 0, at most 1000 iterations), batched.
 *   x1, x2   dev f32 [batch][n_max][2]  normalised coordinates (outputs of vslam_undistort_points)
 *   n_pts    dev i32 [batch]
 *   words    dev u32 [batch][H][5]      harness-drawn random words, sampling rule as for PnP (k = 5)
 *   models   dev f64 [batch][H][10][9]  essential matrices of every sample (unit Frobenius norm,
 *                                       canonical sign and order), row-major, zero-padded
 *   n_models dev i32 [batch][H]         0..10
 * Five-point solver: Nister (null space, ten cubic constraints, degree-10 polynomial) + a
 * Gauss-Newton polish of every root on the constraints.
 */
int vslam_essential_hypotheses(const float* x1, const float* x2, const int32_t* n_pts, int n_max,
                               int batch, const uint32_t* words, int H, double* models,
                               int32_t* n_models, void* stream);
/* Inlier count of every model: inlier iff (float)Sampson <= (float)(thr*thr), OpenCV's semantics
 * (EMEstimatorCallback::computeError).  One warp per model.  counts dev i32 [batch][H][10]. */
int vslam_essential_score(const float* x1, const float* x2, const int32_t* n_pts, int n_max, int batch,
                          const double* models, const int32_t* n_models, int H, double thr,
                          int32_t* counts, void* stream);
/*
 * Selection with OpenCV's shrinking budget (RANSACUpdateNumIters, 5 model points), inlier mask of the
 * winner and the epiline score of estimateEssential (helper_functions.py:54-68).
 *   best dev i32 [batch] (sample*10 + model, or -1);  E dev f64 [batch][9];
 *   mask dev u8 [batch][n_max] in {0,1};  n_inliers dev i32 [batch];  score dev f64 [batch]
 */
int vslam_essential_select(const float* x1, const float* x2, const int32_t* n_pts, int n_max, int batch,
                           const double* models, const int32_t* n_models, const int32_t* counts, int H,
                           double thr, double confidence, int32_t* best, double* E, uint8_t* mask,
                           int32_t* n_inliers, double* score, void* stream);
/*
 * The iteration budget cv2's RANSAC loop (RANSACUpdateNumIters) has left after the first H hypotheses of a run
 * that started with max_iters iterations.  A caller that generates hypotheses in rounds may stop at the first
 * prefix with budget <= H: vslam_essential_select on that prefix returns the model of the full max_iters run.
 *   n_models dev i32 [batch][H], counts dev i32 [batch][H][10], budget dev i32 [batch]
 */
int vslam_essential_budget(const int32_t* n_pts, int n_max, int batch, const int32_t* n_models,
                           const int32_t* counts, int H, int max_iters, double confidence, int32_t* budget,
                           void* stream);

/*
 * cv2.recoverPose(E, p1, p2, cameraMatrix=K, distanceThresh) (reference helper_functions.py:176).
 *   E dev f64 [batch][9];  p1, p2 dev f32 [batch][n_max][2] PIXEL coordinates;  K host f64[4]
 *   R dev f64 [batch][9];  t dev f64 [batch][3] (unit);  mask dev u8 [batch][n_max] in {0,255}
 *   tri dev f64 [batch][4][n_max]  homogeneous points in the first camera's frame (unit columns,
 *       sign free -- callers de-homogenise, main.py:123-124)
 *   good dev i32 [batch][5]  survivors of the chosen candidate, then of the four candidates
 */
int vslam_recover_pose(const double* E, const float* p1, const float* p2, const int32_t* n_pts, int n_max,
                       int batch, const double* K, double distance_thresh, double* R, double* t,
                       uint8_t* mask, double* tri, int32_t* good, void* stream);

/*
 * Batched two-view triangulation + cheirality/depth filter: the per-point loop of
 * reference src/v2/helper_functions.py:281-291 (`triangulate`: rows u*P[2]-P[0], v*P[2]-P[1] of both
 * views, np.linalg.svd, vt[3]) and the filter that follows it at src/v2/main.py:284-306.
 *   proj  dev f64 [batch][2][12]  the two 3x4 projection matrices (row-major)
 *   extr  dev f64 [batch][2][12]  world->camera 3x4 matrices for the depth test, or NULL (then depth
 *                                 and good must be NULL)
 *   x1,x2 dev f64 [batch][n_max][2] pixel coordinates;  n_pts dev i32 [batch]
 *   Xh    dev f64 [batch][n_max][4] unit homogeneous point (sign: largest component positive; the
 *                                   reference's sign is LAPACK's, callers divide by w at main.py:285)
 *   X3    dev f64 [batch][n_max][3] Xh[:3]/Xh[3] (may be NULL);  depth dev f64 [batch][n_max][2] (may be NULL)
 *   good  dev u8  [batch][n_max]    1 iff 0 < z1 < z_max and 0 < z2 < z_max (main.py:306 uses z_max = 1)
 */
int vslam_triangulate(const double* proj, const double* extr, const double* x1, const double* x2,
                      const int32_t* n_pts, int n_max, int batch, double z_max, double* Xh, double* X3,
                      double* depth, uint8_t* good, void* stream);

/*
 * Unmatched-feature selection: indices of the rows of kp1 that are "not in" kp2, reference
 * src/v2/helper_functions.py:316-326 (`GetListDiff`, used at src/v2/main.py:239).
 *   kp1   dev f64 [n1][2], kp2 dev f64 [n2][2]   pixel coordinates (exact comparison)
 *   mode  0: the reference's effective test, numpy `x not in kp2` = no row of kp2 shares EITHER coordinate
 *            with x;  1: row-wise test (both coordinates equal), what the reference's dead loop spells out
 *   keep      dev u8  [n1]  scratch / 1 where the row is kept
 *   keep_idx  dev i32 [n1]  ascending kept indices, first *n_keep valid;  n_keep dev i32 [1]
 */
int vslam_list_diff(const double* kp1, int n1, const double* kp2, int n2, int mode, uint8_t* keep,
                    int32_t* keep_idx, int32_t* n_keep, void* stream);

/*
 * Feature extraction: reference src/v2/frame.py:5-14 (`FeatureExtractor.compute_features`,
 * ORB variant of frame.py:8) = cv2.goodFeaturesToTrack(mean-of-channels grey, max_corners, quality,
 * min_distance) -> cv2.KeyPoint(x, y, size 20) [angle -1] -> cv2.ORB_create().compute -> KeyPoint_convert.
 * Bit-exact against OpenCV 4.13 (corner order, key points, descriptors).
 *   img        dev u8  [batch][H][W][3]  BGR, interleaved (cv2.imread layout)
 *   max_corners, quality, min_distance   goodFeaturesToTrack arguments (frame.py:11: 3000, 0.01, 7);
 *                                        min_distance must be an integer value or < 1
 *   cos_a, sin_a                         float(cos), float(sin) of float(angle) * float(pi/180), the key point
 *                                        angle ORB rotates its pattern by (KeyPoint default: -1 degree)
 *   pattern    dev i8  [512][2]          the 256 rBRIEF test pairs (x, y), OpenCV's published table
 *   workspace  dev, vslam_features_workspace_bytes(batch, H, W, min_distance) bytes, 256-byte aligned
 *   eig        dev f32 [batch][H][W]     cv2.cornerMinEigenVal(grey, 3, ksize=3)         (output)
 *   blur       dev u8  [batch][H][W]     the blurred level image ORB samples             (output)
 *   corners    dev f32 [batch][max_corners][2], n_corners dev i32 [batch]   goodFeaturesToTrack result
 *   kp         dev f32 [batch][max_corners][2], n_kp dev i32 [batch]        key points ORB keeps (31-px border)
 *   desc       dev u8  [batch][max_corners][32]                             their descriptors
 *   status     dev i32 [batch] or NULL   1 if an internal list overflowed (cannot happen for valid input)
 */
long long vslam_features_workspace_bytes(int batch, int H, int W, double min_distance);
int vslam_extract_features(const uint8_t* img, int batch, int H, int W, int max_corners, double quality,
                           double min_distance, float cos_a, float sin_a, const int8_t* pattern,
                           void* workspace, long long workspace_bytes, float* eig, uint8_t* blur,
                           float* corners, int32_t* n_corners, float* kp, uint8_t* desc, int32_t* n_kp,
                           int32_t* status, void* stream);

/* ------------------------------------------------------------------------- */
/* 3. Bundle adjustment                                                       */
/* ------------------------------------------------------------------------- */
/*
 * Opaque problem object = what reference src/v2/LocalBA.py builds as a g2o graph on every call
 * (add_pose :56-65, add_point :68-77, add_edge :79-94, AddScalingEdge :115-131) plus the solver
 * stack it selects (:22-25: BlockSolverSE3 + Levenberg).  Structure arrays are HOST pointers
 * (graph construction is host work in the reference too); state arrays are DEVICE pointers.
 *
 *   cams   [n_cams][7] f64  camera->world pose, quaternion (x,y,z,w) + translation (SBACam)
 *   pts    [n_pts][3]  f64
 *   obs_uv [n_obs][2]  f64  pixel measurements, in the order of obs_cam / obs_pt
 *   K      host f64[4]      fx, fy, cx, cy (LocalBA.py:59 set_cam, baseline 0)
 * Fixed vertices (cam_fixed / pt_fixed, uint8 flags, may be NULL) get no column; all points fixed
 * = motionOnlyBundleAdjustement (:195-229).  Scale edges are EdgeSBAScale with DCS (:115-131).
 */
typedef struct vslam_ba vslam_ba;

int vslam_ba_create(vslam_ba** out, int n_cams, int n_pts, int n_obs, int n_scale,
                    const int32_t* obs_cam, const int32_t* obs_pt,
                    const uint8_t* cam_fixed, const uint8_t* pt_fixed,
                    const int32_t* scale_a, const int32_t* scale_b);
/* Re-uses a problem object for another structure (same arguments as vslam_ba_create): device buffers are kept
 * while the new sizes fit (capacity doubles otherwise), so the per-call cudaMalloc / cudaFree of building a fresh
 * optimiser for every BA call -- what the reference does (src/v2/main.py:146,213,322) -- disappears.  Synchronises
 * the device; vslam_ba_set_state must follow. */
int vslam_ba_set_structure(vslam_ba* ba, int n_cams, int n_pts, int n_obs, int n_scale,
                           const int32_t* obs_cam, const int32_t* obs_pt, const uint8_t* cam_fixed,
                           const uint8_t* pt_fixed, const int32_t* scale_a, const int32_t* scale_b);
int vslam_ba_destroy(vslam_ba* ba);
int vslam_ba_set_state(vslam_ba* ba, const double* cams, const double* pts, const double* obs_uv,
                       const double* scale_meas, const double* K, void* stream);
int vslam_ba_get_state(vslam_ba* ba, double* cams, double* pts, void* stream);
/* PCG relative residual tolerance (default 1e-10) and iteration cap (<=0 keeps the default). */
int vslam_ba_set_solver(vslam_ba* ba, double pcg_tol, int pcg_max_iter);
/* Linear solver of the reduced camera system: 0 = automatic = 1 = block-Jacobi PCG (north_star), 2 = in-CTA direct
 * solve for n = 6 * free cameras <= 96 (what g2o's CHOLMOD path does for a local window; rejected when n > 96):
 * register-resident Gauss-Jordan elimination for n <= 63, shared-memory LDL^T above; 3 = the LDL^T kernel at every
 * n <= 96.  A non-positive pivot / curvature is reported the same way by all (rejected LM step).  Measured on B200
 * at n = 54 (C3): PCG 8-12 us per solve (its matrix entries live in registers), Gauss-Jordan 17 us, LDL^T 33 us;
 * C3 as a whole 20.6k / 18.7k / 12k LM iterations per second.  The PCG runs one of four kernels by n: one CTA with the
 * matrix in registers (n <= 64, n <= 128), one thread-block cluster exchanging through distributed shared memory
 * (n <= 480 where the device places the cluster), a cooperative multi-CTA kernel beyond. */
int vslam_ba_set_linear_solver(vslam_ba* ba, int kind);

/*
 * SparseOptimizer::optimize(max_iterations) with OptimizationAlgorithmLevenberg
 * (LocalBA.py:39-42 calls it with 10).  stats8 (host): initial cost, final cost, outer
 * iterations, trials, final lambda, total PCG iterations, rejected trials, PCG breakdowns.
 * Cost = sum of robustified chi2 (Huber sqrt(5.991) on projections, DCS on scale edges).
 */
int vslam_ba_optimize(vslam_ba* ba, int max_iterations, double* stats8, void* stream);

/* The stages of one LM iteration, exposed for parity tests and profiling. */
/* vslam_ba_optimize for `count` INDEPENDENT problems at once (one local window per sequence, SURVEY 8e: "multiple
 * independent BA problems shard like pairs"): their LM slots are enqueued side by side on internal streams and
 * overlap on the device; results are identical to optimising them one after the other.  stats8: host f64 [count][8].
 * Synchronises the device. */
int vslam_ba_optimize_batch(vslam_ba** bas, int count, int max_iterations, double* stats8);

/* buildSystem(): out2 (host) = {robust cost, max |diag H|}.  Synchronises the stream. */
int vslam_ba_linearize(vslam_ba* ba, double* out2, void* stream);
/* Profiling: times one stage with CUDA events recorded on `stream` around the stage's own kernels:
 * 0 = fused linearisation kernel (buildSystem), 1 = Schur assembly, 2 = one whole LM trial: each of the
 * `reps` launches bracketed on its own after flush_buf/flush_bytes (device, may be NULL/0) are
 * overwritten to evict L2, ms_out[reps];  3 = the fused linearisation kernel `reps` times back to back
 * between one pair of events, ms_out[0] = average per launch.  ms_out on the host.  Synchronises. */
int vslam_ba_time_stage(vslam_ba* ba, int stage, int reps, void* flush_buf, int64_t flush_bytes,
                        float* ms_out, void* stream);
/* Profiling: one linearisation with per-CTA %globaltimer stamps; out[3*i..] (host) = {kind (0 camera
 * chunk, 1 point block, 2 scale edges), start_ns, end_ns} for CTA i of the fused grid, *n = CTA count. */
int vslam_ba_trace_linearize(vslam_ba* ba, int64_t* out, int64_t capacity, int32_t* n, void* stream);
/* One damped solve + trial update: out4 (host) = {trial cost, computeScale(), PCG status
 * (0 converged, 1 breakdown, 2 iteration cap), PCG iterations}.  Synchronises the stream. */
int vslam_ba_trial(vslam_ba* ba, double lambda, double* out4, void* stream);
/* Make the last trial state current (g2o discardTop()). */
int vslam_ba_accept(vslam_ba* ba);
/* out8 (host): free cams, free pts, n = 6*free cams, Schur items, work units, S blocks, PCG CTAs, rows/CTA */
int vslam_ba_dims(vslam_ba* ba, int32_t* out8);
/* Copies an internal buffer to dst (device): 0 Hpp[Cf][36] 1 bp 2 Hll[Pf][6] 3 bl 4 S[n][n]
 * 5 bhat 6 dp 7 dl 8 trial cams 9 trial pts. */
int vslam_ba_read(vslam_ba* ba, int what, double* dst, int64_t capacity, void* stream);

/* The small-system solver on its own on a dense symmetric system, n a multiple of 6, n <= 96 (device pointers):
 * Gauss-Jordan for n <= 63, LDL^T above; pass -n to force the LDL^T kernel.  *status (host) = 0 solved, 1 = a pivot
 * was not positive (x = 0): g2o's "Cholesky failure". */
int vslam_direct_solve(const double* S, const double* b, double* x, int n, int32_t* status, void* stream);

/* Block-Jacobi PCG on a caller-supplied dense symmetric system (n multiple of 6, device pointers);
 * status_iters (host int32[2]) = {status, iterations}.  Status 1 = not positive definite, the case
 * g2o reports as "Cholesky failure" (reference debug.txt). */
int vslam_pcg_solve(const double* S, const double* b, double* x, int n, double tol, int max_iter,
                    int32_t* status_iters, void* stream);

/*
 * vslam_pcg_solve WITHOUT any allocation.  This entry point has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernels and results as vslam_pcg_solve; its scratch (grid-barrier
 * counter, flags, partial sums, two vectors of n doubles) lives in a file-scope device array of 24.25 KiB that every
 * GPU loading the library holds until exit.  It calls no cudaMalloc / cudaFree and does not synchronise.
 *   n                 multiple of 6, at most vslam_pcg_solve_static_max_n() = 1536 unknowns (256 free cameras);
 *                     beyond that it returns VSLAM_ERR_TOO_LARGE (message in vslam_last_cuda_error()) and launches nothing
 *   status_iters_dev  DEVICE int32[2] = {status, iterations}, written by the kernel (stream-ordered)
 * All calls share the one scratch array, which keeps the last call's contents (what must start at zero is cleared
 * on `stream` before every launch): calls that overlap on the GPU -- two streams, two host threads -- race on it
 * and must be serialised by the caller.
 */
int vslam_pcg_solve_static(const double* S, const double* b, double* x, int n, double tol, int max_iter,
                           int32_t* status_iters_dev, void* stream);
int vslam_pcg_solve_static_max_n(void);

/* ------------------------------------------------------------------------- */
/* 5. Device-resident tracking window                          */
/* ------------------------------------------------------------------------- */
/*
 * The local map of the reference's tracking loop (src/v2/main.py:153-214) as device arrays that frames are
 * appended to, and the motion-only bundle adjustment that runs after every tracked frame
 * (src/v2/LocalBA.py:195-229: key frame fixed, all points fixed, every tracked frame free, observation edges with
 * Huber sqrt(5.991), 10 Levenberg iterations) as ONE cooperative kernel launch.  Replaces, per frame, the
 * reference's graph rebuild (LocalBA.py:198-214: O(points x frames) dict probes, one pybind call per vertex and
 * edge) and a fresh optimiser.  A frame's observations arrive together, so appending a frame keeps camera-major
 * order: nothing is sorted, nothing is re-uploaded.
 *
 *   create   max_frames <= number of SMs (one co-resident CTA per frame, key frame included);
 *            slot = observations a frame can hold (the nq of the matching call); max_points = rows of the point list
 *   reset    a new key frame: pts dev f64 [n_pts][3] the map points it sees (main.py:183 / map.py:28-44),
 *            kf_uv dev f32 [n_pts][2] their pixels in the key frame, kf_cam7 dev f64 [7] its camera->world pose as
 *            (qx,qy,qz,qw,tx,ty,tz), K host f64[4]
 *   append   main.py:208-210: one tracked frame -- cam7 dev f64 [7] (vslam_pose_camera_to_world's output) and one
 *            observation per ratio match: point row keep_q[i], pixel kp_cur[keep_t[i]], i < *n_keep (all device
 *            pointers, exactly vslam_match_filter's outputs; n_keep is read on the device: no host round trip)
 *   motion_ba  LocalBA.py:195-229 on the window; stats8 (host, may be NULL = no synchronisation) as vslam_ba_optimize
 *   info     out4 = {frames, points, max_frames, slot}; device pointers of the camera array [max_frames][7]
 *            (updated in place by motion_ba) and of the stats
 */
typedef struct vslam_window vslam_window;
int vslam_window_create(vslam_window** out, int max_frames, int slot, int max_points);
int vslam_window_destroy(vslam_window* w);
int vslam_window_reset(vslam_window* w, const double* pts, const float* kf_uv, int n_pts, const double* kf_cam7,
                       const double* K, void* stream);
int vslam_window_append(vslam_window* w, const double* cam7, const int32_t* keep_q, const int32_t* keep_t,
                        const int32_t* n_keep, const float* kp_cur, void* stream);
int vslam_window_motion_ba(vslam_window* w, int max_iterations, double* stats8, void* stream);
int vslam_window_info(vslam_window* w, int32_t* out4, double** cams_dev, double** stats_dev);
/* Copies `count` camera rows starting at frame `first` into dst (dev f64 [count][7]), stream-ordered. */
int vslam_window_read_cams(vslam_window* w, double* dst, int first, int count, void* stream);

/* The library holds no mutable state besides the device-attribute cache and the last-error string; the kernel
 * variant selectors used by tools/ and tests/ are declared in visual_slam_b200/csrc/tuning.h, not here. */

#ifdef __cplusplus
}
#endif
#endif /* VSLAM_B200_H */
