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
