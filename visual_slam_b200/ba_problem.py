"""Structure-of-arrays bundle-adjustment problem (what LocalBA.py builds as a g2o graph).

Pose convention follows the reference: camera->world, quaternion (x, y, z, w) + translation,
exactly what add_pose feeds SBACam (reference src/v2/LocalBA.py:56-65).
"""
import numpy as np


def R_to_quat(R):
    """Rotation matrix -> UNIT quaternion (x, y, z, w) with w >= 0: Eigen's conversion followed by the
    normalisation and sign choice g2o's SBACam / SE3Quat constructors apply (LocalBA.py:57-58), so that a
    slightly non-orthonormal R (cv2.recoverPose output, products of 4x4 poses) still gives a rotation."""
    R = np.asarray(R, dtype=np.float64)
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    q = [0.0, 0.0, 0.0, 0.0]
    if tr > 0:                                   # operation order of Eigen::Quaternion(Matrix3)
        s = np.sqrt(tr + 1.0)
        q[3] = 0.5 * s
        s = 0.5 / s
        q[0], q[1], q[2] = (R[2, 1] - R[1, 2]) * s, (R[0, 2] - R[2, 0]) * s, (R[1, 0] - R[0, 1]) * s
    else:
        i = 0
        if R[1, 1] > R[0, 0]:
            i = 1
        if R[2, 2] > R[i, i]:
            i = 2
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0)
        q[i] = 0.5 * s
        s = 0.5 / s
        q[3] = (R[k, j] - R[j, k]) * s
        q[j] = (R[j, i] + R[i, j]) * s
        q[k] = (R[k, i] + R[i, k]) * s
    q = np.asarray(q, dtype=np.float64)
    if q[3] < 0:
        q = -q
    return q / np.linalg.norm(q)


def quat_to_R(q):
    x, y, z, w = [float(v) for v in q]
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


class BAProblem:
    """cams [C,7] f64 (qx,qy,qz,qw,tx,ty,tz); pts [P,3] f64; observations (cam, pt, uv);
    scale edges (cam a, cam b, measured baseline); fixed flags; K = (fx, fy, cx, cy)."""

    def __init__(self, cams, pts, obs_cam, obs_pt, obs_uv, K, cam_fixed=None, pt_fixed=None,
                 scale_a=None, scale_b=None, scale_meas=None):
        self.cams = np.ascontiguousarray(cams, dtype=np.float64).reshape(-1, 7)
        self.pts = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 3)
        self.obs_cam = np.ascontiguousarray(obs_cam, dtype=np.int32)
        self.obs_pt = np.ascontiguousarray(obs_pt, dtype=np.int32)
        self.obs_uv = np.ascontiguousarray(obs_uv, dtype=np.float64).reshape(-1, 2)
        self.K = tuple(float(k) for k in K)
        C, P = self.cams.shape[0], self.pts.shape[0]
        self.cam_fixed = (np.zeros(C, bool) if cam_fixed is None else np.asarray(cam_fixed, bool))
        self.pt_fixed = (np.zeros(P, bool) if pt_fixed is None else np.asarray(pt_fixed, bool))
        e = np.zeros(0, np.int32)
        self.scale_a = e if scale_a is None else np.ascontiguousarray(scale_a, dtype=np.int32)
        self.scale_b = e if scale_b is None else np.ascontiguousarray(scale_b, dtype=np.int32)
        self.scale_meas = (np.zeros(0) if scale_meas is None
                           else np.ascontiguousarray(scale_meas, dtype=np.float64))
        n = len(self.obs_cam)
        if len(self.obs_pt) != n or self.obs_uv.shape[0] != n:
            raise ValueError("observation arrays disagree in length")
        if n and (self.obs_cam.min() < 0 or self.obs_cam.max() >= C or
                  self.obs_pt.min() < 0 or self.obs_pt.max() >= P):
            raise ValueError("observation index out of range")

    n_cams = property(lambda s: s.cams.shape[0])
    n_pts = property(lambda s: s.pts.shape[0])
    n_obs = property(lambda s: len(s.obs_cam))
    n_scale = property(lambda s: len(s.scale_a))

    @property
    def cam_col(self):
        col = np.full(self.n_cams, -1, np.int64)
        free = ~self.cam_fixed
        col[free] = np.arange(int(free.sum()))
        return col

    @property
    def pt_col(self):
        col = np.full(self.n_pts, -1, np.int64)
        free = ~self.pt_fixed
        col[free] = np.arange(int(free.sum()))
        return col

    n_free_cams = property(lambda s: int((~s.cam_fixed).sum()))
    n_free_pts = property(lambda s: int((~s.pt_fixed).sum()))

    def copy(self):
        return BAProblem(self.cams.copy(), self.pts.copy(), self.obs_cam, self.obs_pt, self.obs_uv,
                         self.K, self.cam_fixed, self.pt_fixed, self.scale_a, self.scale_b,
                         self.scale_meas)

    def linearize_bytes(self):
        """Algorithmic bytes of one linearisation."""
        return 184 * self.n_obs + 96 * self.n_pts + 392 * self.n_cams


ICL_K = (481.2, 480.0, 319.5, 239.5)


def synthetic_ba(n_cams=10, n_pts=5000, obs_per_pt=6.0, seed=0, K=ICL_K, pix_sigma=1.0,
                 outlier_frac=0.05, pose_noise=(1.0, 0.05), pt_noise=0.02, width=640, height=480,
                 motion_only=False, with_scale=True):
    """C3 / C4 generator: cameras on an arc looking at a 6x4x3 m cloud, frustum
    visibility with dropout tuned to ~obs_per_pt, sigma px noise, gross outliers, perturbed init,
    camera 0 fixed, scale edges between consecutive cameras (measurement = true baseline)."""
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = K
    pts_true = np.stack([rng.uniform(-3, 3, n_pts), rng.uniform(-2, 2, n_pts),
                         rng.uniform(4.0, 7.0, n_pts)], axis=1)
    ang = np.linspace(-0.35, 0.35, n_cams) if n_cams > 1 else np.zeros(1)
    centre = np.array([0.0, 0.0, 5.5])
    cams_true = np.zeros((n_cams, 7))
    Rs, ts = [], []
    for i, a in enumerate(ang):
        radius = 5.5 + 0.3 * np.sin(3 * a)
        pos = centre + radius * np.array([np.sin(a), 0.15 * np.sin(2 * a + i), -np.cos(a)])
        z = centre - pos
        z /= np.linalg.norm(z)
        x = np.cross(np.array([0.0, 1.0, 0.0]), z)
        x /= np.linalg.norm(x)
        y = np.cross(z, x)
        R = np.stack([x, y, z], axis=1)  # camera->world
        cams_true[i, 0:4] = R_to_quat(R)
        cams_true[i, 4:7] = pos
        Rs.append(R); ts.append(pos)
    # visibility: inside the image for camera c, then random dropout to the target density
    vis_c, vis_p, uvs = [], [], []
    for c in range(n_cams):
        pc = (pts_true - ts[c]) @ Rs[c]
        u = fx * pc[:, 0] / pc[:, 2] + cx
        v = fy * pc[:, 1] / pc[:, 2] + cy
        ok = (pc[:, 2] > 0.5) & (u >= 0) & (u < width) & (v >= 0) & (v < height)
        idx = np.nonzero(ok)[0]
        vis_c.append(np.full(len(idx), c)); vis_p.append(idx)
        uvs.append(np.stack([u[idx], v[idx]], axis=1))
    vis_c = np.concatenate(vis_c); vis_p = np.concatenate(vis_p); uvs = np.concatenate(uvs)
    keep_p = min(1.0, obs_per_pt * n_pts / max(len(vis_c), 1))
    keep = rng.random(len(vis_c)) < keep_p
    obs_cam, obs_pt, uv = vis_c[keep], vis_p[keep], uvs[keep]
    # every point needs >= 2 observations to be constrained; drop the rest from the problem
    cnt = np.bincount(obs_pt, minlength=n_pts)
    good = cnt[obs_pt] >= 2
    obs_cam, obs_pt, uv = obs_cam[good], obs_pt[good], uv[good]
    used = np.nonzero(np.bincount(obs_pt, minlength=n_pts) > 0)[0]
    remap = np.full(n_pts, -1); remap[used] = np.arange(len(used))
    obs_pt = remap[obs_pt]
    pts_true = pts_true[used]
    uv = uv + rng.normal(0, pix_sigma, uv.shape)
    out = rng.random(len(uv)) < outlier_frac
    uv[out] += rng.uniform(-50, 50, (int(out.sum()), 2))
    order = np.lexsort((obs_cam, obs_pt))
    obs_cam, obs_pt, uv = obs_cam[order], obs_pt[order], uv[order]
    # perturbed initial state
    cams = cams_true.copy()
    rot_sig = np.deg2rad(pose_noise[0])
    for c in range(1, n_cams):
        axis_angle = rng.normal(0, rot_sig / np.sqrt(3), 3)
        th = np.linalg.norm(axis_angle)
        dq = np.concatenate([np.sin(th / 2) * axis_angle / max(th, 1e-12), [np.cos(th / 2)]])
        x1, y1, z1, w1 = cams_true[c, 0:4]; x2, y2, z2, w2 = dq
        cams[c, 0:4] = [w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                        w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2]
        cams[c, 4:7] += rng.normal(0, pose_noise[1] / np.sqrt(3), 3)
    cam_fixed = np.zeros(n_cams, bool); cam_fixed[0] = True
    if motion_only:
        pts, pt_fixed = pts_true.copy(), np.ones(len(pts_true), bool)
    else:
        pts, pt_fixed = pts_true + rng.normal(0, pt_noise / np.sqrt(3), pts_true.shape), None
    sa = sb = sm = None
    if with_scale and n_cams > 1 and not motion_only:
        sa = np.arange(0, n_cams - 1); sb = np.arange(1, n_cams)
        sm = np.linalg.norm(cams_true[sb, 4:7] - cams_true[sa, 4:7], axis=1)
    return BAProblem(cams, pts, obs_cam, obs_pt, uv, K, cam_fixed, pt_fixed, sa, sb, sm)
