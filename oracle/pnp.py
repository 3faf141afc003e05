"""Oracle for the RANSAC PnP path.  TEST INFRASTRUCTURE ONLY.

Reference call: cv2.solvePnPRansac(objectPoints, imagePoints, cameraMatrix, distCoeffs=[], rvec,
tvec, useExtrinsicGuess=True) with every other argument defaulted (reference src/v2/main.py:196-197):
100 iterations, reprojectionError 8.0, confidence 0.99, SOLVEPNP_ITERATIVE.  OpenCV is a third-party
wheel (un-pinned in the reference; 4.13.0 here).  Its published algorithm (calib3d solvepnp.cpp /
ptsetreg.cpp), restated:

  Stage A  RANSAC: sample -> minimal solve -> err_i = |uv_i - proj_i|^2 computed in float32 from the
           float32-rounded projection; inlier iff err_i <= (float)(thr*thr); a model replaces the best
           iff its count is strictly larger (and >= sample size); the iteration budget shrinks with
           RANSACUpdateNumIters(confidence, outlier ratio, sample size, budget).
  Stage B  cv2.solvePnP(inliers, ITERATIVE, useExtrinsicGuess=True) from the caller's guess
           (verified bit-exact against solvePnPRansac's output in SURVEY.md 8a-2).

cv2 draws its samples from an internal RNG that Python cannot seed, so -- as BASELINE.json's north_star
prescribes -- parity is defined on a GIVEN hypothesis set: the harness draws the samples, and scoring,
selection, inlier sets and the refined pose must agree.  The minimal solver restated here is P3P on
three points with a fourth for disambiguation (what cv2 selects for SOLVEPNP_P3P); cv2's default
RANSAC kernel is EPnP on 5 points -- a documented deviation that does not change Stage B's optimum.
Pinned against live cv2 in tests/test_oracle_pnp.py (projectPoints, solvePnP(P3P), solvePnPRansac).
"""
import numpy as np


def rodrigues(rvec):
    r = np.asarray(rvec, dtype=np.float64).reshape(3)
    th = np.linalg.norm(r)
    if th < 1e-300:
        return np.eye(3)
    k = r / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)


def rodrigues_inv(R):
    """Rotation matrix -> rotation vector (angle < pi branch; near-pi handled through the axis)."""
    R = np.asarray(R, dtype=np.float64)
    c = np.clip((np.trace(R) - 1) / 2, -1, 1)
    th = np.arccos(c)
    v = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    s = np.linalg.norm(v) / 2
    if th < 1e-12:
        return v / 2
    if np.pi - th < 1e-6:
        A = (R + np.eye(3)) / 2
        ax = np.sqrt(np.maximum(np.diag(A), 0))
        i = int(np.argmax(ax))
        ax = A[:, i] / ax[i]
        if np.dot(ax, v) < 0:
            ax = -ax
        return ax / np.linalg.norm(ax) * th
    return v / (2 * s) * th


def project_f32(Rt, X, K):
    """cv2.projectPoints without distortion: double arithmetic, float32 result (PnPRansacCallback)."""
    R, t = Rt[:, :3], Rt[:, 3]
    Xc = X.astype(np.float64) @ R.T + t
    z = np.where(Xc[:, 2] != 0, 1.0 / Xc[:, 2], 1.0)
    u = Xc[:, 0] * z * K[0] + K[2]
    v = Xc[:, 1] * z * K[1] + K[3]
    return np.stack([u, v], axis=1).astype(np.float32)


def reproj_err_f32(Rt, X, uv, K):
    """err_i exactly as cv2 computes it: float32 difference, float32 sum of squares."""
    p = project_f32(Rt, X, K)
    d = uv.astype(np.float32) - p
    return (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32)


def score(hyps, valid, X, uv, K, thr=8.0):
    """Inlier masks [H,N] and counts [H] for every hypothesis (3x4 [R|t], world->camera)."""
    t2 = np.float32(thr * thr)
    masks = np.zeros((len(hyps), len(X)), dtype=bool)
    for h in range(len(hyps)):
        if valid[h]:
            masks[h] = reproj_err_f32(hyps[h], X, uv, K) <= t2
    return masks, masks.sum(axis=1).astype(np.int32)


def ransac_update_num_iters(p, ep, model_points, max_iters):
    """cv::RANSACUpdateNumIters (ptsetreg.cpp)."""
    p = min(max(p, 0.0), 1.0)
    ep = min(max(ep, 0.0), 1.0)
    num = max(1.0 - p, np.finfo(np.float64).tiny)
    denom = 1.0 - (1.0 - ep) ** model_points
    if denom < np.finfo(np.float64).tiny:
        return 0
    num = np.log(num)
    denom = np.log(denom)
    return max_iters if (denom >= 0 or -num >= max_iters * (-denom)) else int(round(num / denom))


def select(counts, valid, n_points, model_points=4, confidence=0.99):
    """RANSACPointSetRegistrator::run's loop over a fixed hypothesis list: returns best index or -1."""
    niters = len(counts)
    best, best_count = -1, 0
    it = 0
    while it < niters:
        if valid[it]:
            c = int(counts[it])
            if c > max(best_count, model_points - 1):
                best, best_count = it, c
                niters = ransac_update_num_iters(confidence, (n_points - c) / n_points, model_points, niters)
        it += 1
    return best, best_count


# ---- P3P (+1) ---------------------------------------------------------------------------------
def _polymul(a, b):
    return np.convolve(a, b)


def solve_quartic_real(c):
    """Real roots of c[0] x^4 + ... + c[4] (numpy companion matrix; the oracle may be slow)."""
    c = np.asarray(c, dtype=np.float64)
    if abs(c[0]) < 1e-14 * np.abs(c).max():
        r = np.roots(c[1:])
    else:
        r = np.roots(c)
    return np.array([x.real for x in r if abs(x.imag) <= 1e-9 * max(1.0, abs(x.real))])


def p3p(Xw, f):
    """All P3P solutions for 3 world points Xw [3,3] and unit bearings f [3,3]: list of 3x4 [R|t].
    Two-conic elimination: s2 = u s1, s3 = v s1;  u = N(v) / D(v);  quartic in v."""
    P1, P2, P3 = Xw
    a2 = np.sum((P2 - P3) ** 2); b2 = np.sum((P1 - P3) ** 2); c2 = np.sum((P1 - P2) ** 2)
    ca = f[1] @ f[2]; cb = f[0] @ f[2]; cg = f[0] @ f[1]
    # polynomials in v, highest power first
    q = np.array([1.0, -2 * cb, 1.0])                       # 1 + v^2 - 2 v cb
    N = (a2 - c2) * q - b2 * np.array([1.0, 0.0, -1.0])     # (a2-c2) q - b2 (v^2 - 1)
    D = 2 * b2 * np.array([-ca, cg])                        # 2 b2 (cg - v ca)
    D2 = _polymul(D, D)
    quart = b2 * (np.concatenate([[0, 0], D2]) + _polymul(N, N) - 2 * cg * np.concatenate([[0], _polymul(N, D)])) \
        - c2 * _polymul(q, D2)
    sols = []
    for v in solve_quartic_real(quart):
        Dv = np.polyval(D, v)
        if abs(Dv) < 1e-12:
            continue
        u = np.polyval(N, v) / Dv
        den = 1 + v * v - 2 * v * cb
        if den <= 0 or u <= 0 or v <= 0:
            continue
        s1 = np.sqrt(b2 / den)
        C = np.stack([s1 * f[0], u * s1 * f[1], v * s1 * f[2]])
        Rt = _align3(Xw, C)
        if Rt is not None:
            sols.append(Rt)
    return sols


def _frame(A):
    e1 = A[1] - A[0]
    n1 = np.linalg.norm(e1)
    e3 = np.cross(e1, A[2] - A[0])
    n3 = np.linalg.norm(e3)
    if n1 < 1e-300 or n3 < 1e-300:
        return None
    e1 = e1 / n1; e3 = e3 / n3
    return np.stack([e1, np.cross(e3, e1), e3], axis=1)


def _align3(Xw, Xc):
    Fw, Fc = _frame(Xw), _frame(Xc)
    if Fw is None or Fc is None:
        return None
    R = Fc @ Fw.T
    t = Xc[0] - R @ Xw[0]
    return np.concatenate([R, t[:, None]], axis=1)


def p3p_plus1(X4, uv4, K):
    """Minimal solve on 4 correspondences: P3P on the first three, the fourth picks the solution with
    the smallest reprojection error (cv2 solvePnP(SOLVEPNP_P3P) behaviour).  Returns 3x4 or None."""
    X4 = np.asarray(X4, dtype=np.float64); uv4 = np.asarray(uv4, dtype=np.float64)
    xn = np.stack([(uv4[:, 0] - K[2]) / K[0], (uv4[:, 1] - K[3]) / K[1], np.ones(4)], axis=1)
    f = xn / np.linalg.norm(xn, axis=1, keepdims=True)
    best, best_e = None, np.inf
    for Rt in p3p(X4[:3], f[:3]):
        pc = Rt[:, :3] @ X4[3] + Rt[:, 3]
        if pc[2] <= 0:
            continue
        e = (K[0] * pc[0] / pc[2] + K[2] - uv4[3, 0]) ** 2 + (K[1] * pc[1] / pc[2] + K[3] - uv4[3, 1]) ** 2
        if e < best_e:
            best, best_e = Rt, e
    return best


def draw_samples(words, n, k=4):
    """Harness sampling rule shared with the CUDA kernel: index j = (word_j * n) >> 32, duplicates
    resolved by stepping to the next free index modulo n.  words uint32 [H,k] -> int32 [H,k]."""
    words = np.asarray(words, dtype=np.uint64)
    out = np.zeros(words.shape, dtype=np.int32)
    for h in range(words.shape[0]):
        for j in range(k):
            idx = int((int(words[h, j]) * n) >> 32)
            while any(idx == out[h, i] for i in range(j)):
                idx = (idx + 1) % n
            out[h, j] = idx
    return out


def refine_cv2(X, uv, K, inl_idx, rvec0, tvec0):
    """Stage B with the reference's own function."""
    import cv2
    Kmat = np.array([[K[0], 0, K[2]], [0, K[1], K[3]], [0, 0, 1.0]])
    ok, rvec, tvec = cv2.solvePnP(X[inl_idx].astype(np.float64).reshape(-1, 1, 3),
                                  uv[inl_idx].astype(np.float64).reshape(-1, 1, 2), Kmat, None,
                                  rvec=np.asarray(rvec0, np.float64).reshape(3, 1).copy(),
                                  tvec=np.asarray(tvec0, np.float64).reshape(3, 1).copy(),
                                  useExtrinsicGuess=True, flags=cv2.SOLVEPNP_ITERATIVE)
    return ok, rvec.reshape(3), tvec.reshape(3)


def refine_numpy(X, uv, K, inl_idx, rvec0, tvec0, iters=50):
    """Independent Gauss-Newton/LM on the same cost (sum of squared reprojection errors), used to
    cross-check that cv2's Stage B sits at the optimum this restatement defines."""
    R = rodrigues(rvec0); t = np.asarray(tvec0, np.float64).reshape(3).copy()
    Xi = X[inl_idx].astype(np.float64); ui = uv[inl_idx].astype(np.float64)
    lam = 1e-3

    def cost_and_sys(R, t):
        pc = Xi @ R.T + t
        iz = 1.0 / pc[:, 2]
        r = np.stack([K[0] * pc[:, 0] * iz + K[2] - ui[:, 0], K[1] * pc[:, 1] * iz + K[3] - ui[:, 1]], axis=1)
        J = np.zeros((len(Xi), 2, 6))
        # d pc = dw x pc + dt ; columns 0..2 rotation (left perturbation), 3..5 translation
        dp = np.zeros((len(Xi), 3, 6))
        dp[:, 0, 1] = pc[:, 2]; dp[:, 0, 2] = -pc[:, 1]
        dp[:, 1, 0] = -pc[:, 2]; dp[:, 1, 2] = pc[:, 0]
        dp[:, 2, 0] = pc[:, 1]; dp[:, 2, 1] = -pc[:, 0]
        dp[:, 0, 3] = 1; dp[:, 1, 4] = 1; dp[:, 2, 5] = 1
        J[:, 0] = K[0] * (dp[:, 0] * iz[:, None] - pc[:, 0, None] * iz[:, None] ** 2 * dp[:, 2])
        J[:, 1] = K[1] * (dp[:, 1] * iz[:, None] - pc[:, 1, None] * iz[:, None] ** 2 * dp[:, 2])
        H = np.einsum("nki,nkj->ij", J, J); g = np.einsum("nki,nk->i", J, r)
        return float(np.sum(r * r)), H, g

    c, H, g = cost_and_sys(R, t)
    for _ in range(iters):
        d = np.linalg.solve(H + lam * np.diag(np.diag(H)), -g)
        dR = rodrigues(d[:3]); Rn = dR @ R; tn = dR @ t + d[3:]
        cn, Hn, gn = cost_and_sys(Rn, tn)
        if cn <= c:
            R, t, c, H, g = Rn, tn, cn, Hn, gn
            lam = max(lam / 10, 1e-12)
            if np.linalg.norm(d) < 1e-14:
                break
        else:
            lam *= 10
    return rodrigues_inv(R), t


def solve_pnp_ransac(X, uv, K, samples, rvec0, tvec0, thr=8.0, confidence=0.99):
    """Full restated path on a given sample list: minimal solves, scoring, selection, Stage B."""
    hyps = np.zeros((len(samples), 3, 4)); valid = np.zeros(len(samples), bool)
    for h, s in enumerate(samples):
        Rt = p3p_plus1(X[s], uv[s], K)
        if Rt is not None and np.all(np.isfinite(Rt)):
            hyps[h], valid[h] = Rt, True
    masks, counts = score(hyps, valid, X, uv, K, thr)
    best, cnt = select(counts, valid, len(X), 4, confidence)
    if best < 0:
        return {"ok": False, "hyps": hyps, "valid": valid, "counts": counts, "best": best}
    inl = np.nonzero(masks[best])[0]
    ok, rvec, tvec = refine_cv2(X, uv, K, inl, rvec0, tvec0)
    return {"ok": bool(ok), "rvec": rvec, "tvec": tvec, "inliers": inl.astype(np.int32), "best": best,
            "hyps": hyps, "valid": valid, "counts": counts}
