"""Seeded synthetic workloads of the shapes BASELINE.json names."""
import numpy as np


def descriptors_pair(nq=2048, nt=2048, seed=0, planted=0.7, flip=0.10):
    """C2: train = uniform bits; `planted` of the queries are distinct train rows with each
    bit flipped w.p. `flip`; the rest are fresh uniform rows.  Returns (q, t, truth[nq])."""
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, size=(nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, size=(nq, 32), dtype=np.uint8)
    truth = np.full(nq, -1, dtype=np.int32)
    n_pl = min(int(round(planted * nq)), nt)
    if n_pl > 0:
        src = rng.permutation(nt)[:n_pl]
        rows = rng.permutation(nq)[:n_pl]
        noise = np.packbits(rng.random((n_pl, 256)) < flip, axis=1)
        q[rows] = t[src] ^ noise
        truth[rows] = src
    return q, t, truth


def descriptors_batch(batch, nq=2048, nt=2048, seed=0, **kw):
    qs, ts = [], []
    for b in range(batch):
        q, t, _ = descriptors_pair(nq, nt, seed=seed * 1000003 + b, **kw)
        qs.append(q); ts.append(t)
    return np.stack(qs), np.stack(ts)


def low_entropy_pair(nq, nt, seed=0):
    """Adversarial tie set: every byte is 0 or 1, so distances collide massively."""
    rng = np.random.default_rng(seed)
    return (rng.integers(0, 2, size=(nq, 32), dtype=np.uint8),
            rng.integers(0, 2, size=(nt, 32), dtype=np.uint8))


ICL_K = (481.2, 480.0, 319.5, 239.5)


def _rodrigues(r):
    th = float(np.linalg.norm(r))
    if th < 1e-300:
        return np.eye(3)
    k = np.asarray(r, dtype=np.float64) / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)


def tracking_pair(seed, n=2048, outlier_frac=0.3, pix_sigma=0.5, K=ICL_K, width=640, height=480):
    """C5 geometry: n world points in the frustum of the current camera (depth 1-8 m),
    world->camera pose = small motion away from identity (<= 2 deg, <= 5 cm), pixel noise sigma,
    `outlier_frac` of the correspondences get uniform-random pixels (wrong matches).
    Returns X [n,3] f32, uv [n,2] f32, rvec [3], tvec [3] (true world->camera), inlier mask [n]."""
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = K
    rvec = rng.normal(size=3); rvec *= np.deg2rad(rng.uniform(0.2, 2.0)) / np.linalg.norm(rvec)
    tvec = rng.normal(size=3); tvec *= rng.uniform(0.005, 0.05) / np.linalg.norm(tvec)
    R = _rodrigues(rvec)
    z = rng.uniform(1.0, 8.0, n)
    u = rng.uniform(0, width, n); v = rng.uniform(0, height, n)
    Xc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], axis=1)
    X = (Xc - tvec) @ R          # world = R^T (Xc - t)
    uv = np.stack([u, v], axis=1) + rng.normal(0, pix_sigma, (n, 2))
    inl = rng.random(n) >= outlier_frac
    n_out = int((~inl).sum())
    uv[~inl] = np.stack([rng.uniform(0, width, n_out), rng.uniform(0, height, n_out)], axis=1)
    return X.astype(np.float32), uv.astype(np.float32), rvec, tvec, inl


def tracking_frame_pair(seed, n=2048, planted=0.7, flip=0.10, wrong_pix=0.10, pix_sigma=0.5, K=ICL_K,
                        width=640, height=480):
    """One C5 frame pair: a key frame's n mapped points (descriptor + 3-D position) and the current
    frame's n features (descriptor + keypoint).  `planted` of the key-frame points reappear in the
    current frame with noisy descriptors; `wrong_pix` of those carry a wrong pixel (PnP outliers).
    Returns desc_kf u8[n,32], desc_cur u8[n,32], map_xyz f32[n,3], kp_cur f32[n,2], rvec, tvec."""
    rng = np.random.default_rng(seed + 7919)
    X, uv, rvec, tvec, _ = tracking_pair(seed, n, outlier_frac=0.0, pix_sigma=pix_sigma, K=K,
                                         width=width, height=height)
    desc_kf = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    desc_cur = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    kp_cur = np.stack([rng.uniform(0, width, n), rng.uniform(0, height, n)], axis=1).astype(np.float32)
    n_pl = int(round(planted * n))
    src = rng.permutation(n)[:n_pl]      # key-frame rows that reappear
    dst = rng.permutation(n)[:n_pl]      # their rows in the current frame
    noise = np.packbits(rng.random((n_pl, 256)) < flip, axis=1)
    desc_cur[dst] = desc_kf[src] ^ noise
    kp_cur[dst] = uv[src]
    bad = rng.random(n_pl) < wrong_pix
    kp_cur[dst[bad]] = np.stack([rng.uniform(0, width, int(bad.sum())),
                                 rng.uniform(0, height, int(bad.sum()))], axis=1).astype(np.float32)
    return desc_kf, desc_cur, X, kp_cur, rvec, tvec


def two_view_pair(seed, n=400, outlier_frac=0.2, pix_sigma=0.5, K=ICL_K, width=640, height=480):
    """Map-initialisation geometry (reference src/v2/main.py:88-138): n points seen from the first camera
    (identity) and from a second one that moved sideways by 0.2-0.5 m with a small rotation, pixel noise
    sigma, `outlier_frac` wrong matches (uniform pixels in image 2).
    Returns p1 [n,2] f32, p2 [n,2] f32 (pixels), R [3,3], t [3] (x2 ~ R x1 + t), inlier mask [n]."""
    rng = np.random.default_rng(seed + 104729)
    fx, fy, cx, cy = K
    rvec = rng.normal(size=3); rvec *= np.deg2rad(rng.uniform(0.5, 4.0)) / np.linalg.norm(rvec)
    t = rng.normal(size=3) * np.array([1.0, 0.4, 0.4]); t *= rng.uniform(0.2, 0.5) / np.linalg.norm(t)
    R = _rodrigues(rvec)
    z = rng.uniform(1.5, 8.0, n)
    u = rng.uniform(20, width - 20, n); v = rng.uniform(20, height - 20, n)
    X = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], axis=1)
    Xc = X @ R.T + t
    p1 = np.stack([u, v], axis=1) + rng.normal(0, pix_sigma, (n, 2))
    p2 = np.stack([fx * Xc[:, 0] / Xc[:, 2] + cx, fy * Xc[:, 1] / Xc[:, 2] + cy], axis=1) + rng.normal(0, pix_sigma, (n, 2))
    inl = rng.random(n) >= outlier_frac
    n_out = int((~inl).sum())
    p2[~inl] = np.stack([rng.uniform(0, width, n_out), rng.uniform(0, height, n_out)], axis=1)
    return p1.astype(np.float32), p2.astype(np.float32), R, t, inl


def frames(n, H=480, W=640, seed=0, block=8):
    """Synthetic BGR frames for the feature-extraction figures: random 8x8-pixel colour blocks (corners at the
    block junctions, ~4800 candidates so that goodFeaturesToTrack's 3000 cap of frame.py:11 binds) plus +-3
    grey levels of pixel noise.  uint8 (n, H, W, 3)."""
    rng = np.random.default_rng(seed)
    hb, wb = (H + block - 1) // block, (W + block - 1) // block
    out = np.empty((n, H, W, 3), np.uint8)
    for i in range(n):
        coarse = rng.integers(20, 236, (hb, wb, 3), dtype=np.int16)
        img = np.repeat(np.repeat(coarse, block, axis=0), block, axis=1)[:H, :W]
        img = img + rng.integers(-3, 4, (H, W, 3), dtype=np.int16)
        out[i] = np.clip(img, 0, 255).astype(np.uint8)
    return out


# ---- C5: 8 sequences x 1000 frame pairs, every pair seeded by hash(seq, i) ------------------------
# Counter-based generation (splitmix64 on int64 tensors, wrapping arithmetic) so that any pair of any sequence can
# be produced on its own, on the CPU or on the device, with identical bytes: ranks generate their shard in place.
_M64 = (1 << 64)
_C1, _C2, _GOLD = 0xBF58476D1CE4E5B9 - _M64, 0x94D049BB133111EB - _M64, 0x9E3779B97F4A7C15 - _M64


def _wrap(x):
    """Python int -> the same value modulo 2^64 as a signed 64-bit number."""
    x &= _M64 - 1
    return x - _M64 if x >= (1 << 63) else x


def _lsr(x, s):
    return (x >> s) & ((1 << (64 - s)) - 1)


def _mix64(x):
    x = (x ^ _lsr(x, 30)) * _C1
    x = (x ^ _lsr(x, 27)) * _C2
    return x ^ _lsr(x, 31)


def c5_pair_key(seq, i):
    """hash(seq, i) as a Python int in [-2^63, 2^63)."""
    import torch
    return int(_mix64(torch.tensor([_wrap(((seq * 1000003 + i) * 2 + 1) * _GOLD)], dtype=torch.int64))[0])


def _bits(keys, stream, count):
    """int64 [B, count] of hashed counters: keys [B] (pair keys), stream = which array of the pair."""
    import torch
    ctr = torch.arange(count, dtype=torch.int64, device=keys.device)[None, :]
    return _mix64(_mix64(keys[:, None] + _wrap(stream * _GOLD)) + ctr * _GOLD)


def _uniform(keys, stream, count):
    return _lsr(_bits(keys, stream, count), 11).double() * (1.0 / (1 << 53))


def c5_poses(seq, n_pairs=1000):
    """World->camera poses of frames 0..n_pairs of sequence `seq` (host, float64): a smooth random walk, per frame a
    rotation <= 2 deg and a translation <= 5 cm whose directions drift slowly.  Returns rvec [n+1,3], tvec [n+1,3];
    pair i tracks frame i+1 with frame i's pose as the guess."""
    rng = np.random.default_rng(abs(c5_pair_key(seq, -1)) % (1 << 63))
    R, t = np.eye(3), np.zeros(3)
    w, v = rng.normal(size=3), rng.normal(size=3)
    rv, tv = [np.zeros(3)], [np.zeros(3)]
    for _ in range(n_pairs):
        w = 0.95 * w + 0.3 * rng.normal(size=3)
        v = 0.95 * v + 0.3 * rng.normal(size=3)
        dr = w / np.linalg.norm(w) * np.deg2rad(rng.uniform(0.2, 2.0))
        dt = v / np.linalg.norm(v) * rng.uniform(0.005, 0.05)
        dR = _rodrigues(dr)
        R, t = dR @ R, dR @ t + dt
        c = np.clip((np.trace(R) - 1) / 2, -1, 1)
        th = np.arccos(c)
        ax = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
        rv.append(ax * (0.5 if th < 1e-9 else th / (2 * np.sin(th))))
        tv.append(t.copy())
    return np.array(rv), np.array(tv)


def c5_pairs(seq, lo, hi, n=2048, device="cpu", planted=0.7, flip=0.10, wrong_pix=0.10, pix_sigma=0.5, K=ICL_K,
             width=640, height=480, poses=None):
    """Frame pairs lo..hi-1 of sequence `seq` as tensors on `device`: desc_kf u8 [B,n,32], desc_cur u8 [B,n,32],
    map_xyz f32 [B,n,3], kp_cur f32 [B,n,2], pose_true f64 [B,6] (rvec, tvec of the current frame, world->camera),
    guess f64 [B,6] (the previous frame's pose).  n world points uniform in the current camera's frustum (depth
    1-8 m); `planted` of the key-frame points reappear in the current frame with each descriptor bit flipped w.p.
    `flip` and pixel noise sigma; `wrong_pix` of those carry a uniform-random pixel (wrong matches)."""
    import torch
    rv, tv = poses if poses is not None else c5_poses(seq, hi)
    B = hi - lo
    keys = torch.tensor([c5_pair_key(seq, i) for i in range(lo, hi)], dtype=torch.int64, device=device)
    fx, fy, cx, cy = K
    pose_true = torch.from_numpy(np.concatenate([rv[lo + 1:hi + 1], tv[lo + 1:hi + 1]], axis=1)).to(device)
    guess = torch.from_numpy(np.concatenate([rv[lo:hi], tv[lo:hi]], axis=1)).to(device)
    Rm = torch.from_numpy(np.stack([_rodrigues(r) for r in rv[lo + 1:hi + 1]])).to(device)          # [B,3,3]
    tm = pose_true[:, 3:]
    u = _uniform(keys, 1, n) * width
    v = _uniform(keys, 2, n) * height
    z = 1.0 + 7.0 * _uniform(keys, 3, n)
    Xc = torch.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], dim=2)                              # [B,n,3]
    Xw = torch.einsum("bji,bnj->bni", Rm, Xc - tm[:, None, :])                                      # R^T (Xc - t)
    desc_kf = (_bits(keys, 4, n * 4).view(torch.uint8).view(B, n, 32)).clone()
    desc_cur = (_bits(keys, 5, n * 4).view(torch.uint8).view(B, n, 32)).clone()
    kp_cur = torch.stack([_uniform(keys, 6, n) * width, _uniform(keys, 7, n) * height], dim=2)
    n_pl = int(round(planted * n))
    src = torch.argsort(_bits(keys, 8, n), dim=1)[:, :n_pl]                                          # key-frame rows that reappear
    dst = torch.argsort(_bits(keys, 9, n), dim=1)[:, :n_pl]                                          # their rows in the current frame
    flips = (_uniform(keys, 10, n_pl * 256) < flip).view(B, n_pl, 32, 8)
    weights = torch.tensor([128, 64, 32, 16, 8, 4, 2, 1], dtype=torch.int64, device=device)
    noise = (flips.long() * weights).sum(dim=3).to(torch.uint8)                                     # packbits, MSB first
    bi = torch.arange(B, device=device)[:, None]
    desc_cur[bi, dst] = desc_kf[bi, src] ^ noise
    r1, r2 = _uniform(keys, 11, n_pl * 2).view(B, n_pl, 2).clamp_min(1e-300), _uniform(keys, 12, n_pl * 2).view(B, n_pl, 2)
    gauss = torch.sqrt(-2.0 * torch.log(r1)) * torch.cos(2.0 * np.pi * r2)                           # Box-Muller
    pix = torch.stack([u, v], dim=2)[bi, src] + pix_sigma * gauss
    bad = _uniform(keys, 13, n_pl) < wrong_pix
    rnd = torch.stack([_uniform(keys, 14, n_pl) * width, _uniform(keys, 15, n_pl) * height], dim=2)
    kp_cur[bi, dst] = torch.where(bad[:, :, None], rnd, pix)
    return {"desc_kf": desc_kf, "desc_cur": desc_cur, "map_xyz": Xw.float(), "kp_cur": kp_cur.float(),
            "pose_true": pose_true, "guess": guess}


# ---- reference-style containers around a synthetic BA problem (for the drop-in BundleAdjustment's Map interface) ----
class SynthFrame:
    """What LocalBA.py touches of the reference's Frame (src/v2/frame.py:51-125)."""

    def __init__(self, fid, pose, kf=False):
        self.ID, self.pose, self.kf, self.parents = fid, pose, kf, {}

    def GetPose(self): return self.pose
    def GetParentIDs(self): return self.parents.keys()
    def GetTransitionWithParentID(self, pid): return self.parents[pid]
    def IsKeyFrame(self): return self.kf
    def GetID(self): return self.ID


class SynthPoint:
    """... of Point (src/v2/point.py)."""

    def __init__(self, pid, xyz):
        self.ID, self.location_3d, self.frames = pid, xyz, {}

    def GetFrame(self, fid): return self.frames.get(fid)
    def Get3dPoint(self): return self.location_3d
    def GetID(self): return self.ID


class SynthMap:
    """... of Map (src/v2/map.py)."""

    def __init__(self): self.frames, self.points_3d = {}, {}
    def GetFrame(self, fid): return self.frames[fid]
    def GetPoint(self, pid): return self.points_3d[pid]
    def UpdatePose(self, new_pose, frame_id): self.frames[frame_id].pose = new_pose
    def UpdatePoint3D(self, new_point, point_id): self.points_3d[point_id].location_3d = new_point


class IclCamera:
    fx, fy, cx, cy = ICL_K


def map_from_problem(p):
    """A reference-style Map holding the BAProblem `p`: frame ids = camera rows, point ids = point rows, scale edges
    as parent links whose stored relative translation has the measured norm."""
    from .ba_problem import quat_to_R
    m = SynthMap()
    for c in range(p.n_cams):
        T = np.eye(4); T[:3, :3] = quat_to_R(p.cams[c, :4]); T[:3, 3] = p.cams[c, 4:]
        m.frames[c] = SynthFrame(c, T)
    for k in range(p.n_scale):
        a, b = int(p.scale_a[k]), int(p.scale_b[k])
        rel = np.eye(4); rel[:3, 3] = [p.scale_meas[k], 0, 0]
        m.frames[b].parents[a] = rel
    for j in range(p.n_pts):
        m.points_3d[j] = SynthPoint(j, p.pts[j].copy())
    for o in range(p.n_obs):
        m.points_3d[int(p.obs_pt[o])].frames[int(p.obs_cam[o])] = (m.frames[int(p.obs_cam[o])], p.obs_uv[o], None)
    return m
