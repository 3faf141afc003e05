"""Oracle for the two-view (essential matrix) path.  TEST INFRASTRUCTURE ONLY.

Reference calls (all arithmetic lives in the un-vendored OpenCV wheel; 4.13.0 in this image, un-pinned
by the reference):
  * src/v2/helper_functions.py:47-70  estimateEssential: cv2.undistortPoints (x2) ->
    cv2.findEssentialMat(method=RANSAC, prob=0.999, threshold=essTh) -> epiline score.
  * src/v2/helper_functions.py:164-195 estimateRelativePose: cv2.recoverPose(E, p1, p2, cameraMatrix=K,
    distanceThresh=50) + mask filter.

Published algorithm restated here (calib3d five-point.cpp / ptsetreg.cpp / triangulate.cpp):
  five-point  Nister: null space of the 5x9 epipolar system, the ten cubic constraints
              det(E)=0, 2 E E^T E - tr(E E^T) E = 0 expanded in (x,y,z), Gauss-Jordan on the ten
              leading monomials, 3x3 polynomial matrix B(z), det B = degree-10 polynomial, real
              roots (|imag| <= 1e-10), (x,y) from the null vector of B(z), E = xE1+yE2+zE3+E4,
              normalised to unit Frobenius norm.
  scoring     Sampson error in double, stored as float32; inlier iff err <= (float)(thr*thr).
  RANSAC      model replaces the best iff its count is strictly larger (and >= 5); the budget
              shrinks with RANSACUpdateNumIters(0.999, outlier ratio, 5, budget); max 1000.
  recoverPose SVD of E -> (R1,R2,+-t); DLT triangulation (smallest right singular vector of the
              4x4 system) of every point for each of the four; mask = z1*w>0, z1<d, z2>0, z2<d;
              the candidate with the most survivors wins.
cv2 draws its samples from an internal RNG Python cannot seed, so parity is defined on a GIVEN sample
list (north_star); everything downstream of the samples is pinned against live cv2 in
tests/test_oracle_essential.py (5-point model sets, masks, recoverPose, undistortPoints, golden G1).
"""
import numpy as np

from .pnp import ransac_update_num_iters

MODEL_POINTS = 5


# ---- K2: cv2.undistortPoints without distortion ---------------------------------------------------
def undistort_points_f32(pts, K):
    """float32 pixels -> float32 normalised coordinates: ((double)u - cx) * (1/fx), rounded to f32."""
    fx, fy, cx, cy = [float(v) for v in K]
    p = np.asarray(pts, dtype=np.float32).astype(np.float64)
    out = np.stack([(p[:, 0] - cx) * (1.0 / fx), (p[:, 1] - cy) * (1.0 / fy)], axis=1)
    return out.astype(np.float32)


# ---- trivariate polynomials of total degree <= 3, coefficient array c[i,j,k] of x^i y^j z^k -------
def _pmul(a, b):
    out = np.zeros((4, 4, 4))
    ia = np.argwhere(a != 0)
    ib = np.argwhere(b != 0)
    for i, j, k in ia:
        for p, q, r in ib:
            out[i + p, j + q, k + r] += a[i, j, k] * b[p, q, r]
    return out


def _lin(cx, cy, cz, c1):
    p = np.zeros((4, 4, 4))
    p[1, 0, 0], p[0, 1, 0], p[0, 0, 1], p[0, 0, 0] = cx, cy, cz, c1
    return p


# Nister's monomial order: ten leading monomials, then [x z^2, x z, x, y z^2, y z, y, z^3, z^2, z, 1]
_MONO = [(3, 0, 0), (0, 3, 0), (2, 1, 0), (1, 2, 0), (2, 0, 1), (2, 0, 0), (0, 2, 1), (0, 2, 0),
         (1, 1, 1), (1, 1, 0),
         (1, 0, 2), (1, 0, 1), (1, 0, 0), (0, 1, 2), (0, 1, 1), (0, 1, 0), (0, 0, 3), (0, 0, 2),
         (0, 0, 1), (0, 0, 0)]


def _constraints(basis):
    """basis [4,9] (E1..E4 row-major) -> 10x20 coefficient matrix of the cubic constraints."""
    E = [[_lin(basis[0, 3 * r + c], basis[1, 3 * r + c], basis[2, 3 * r + c], basis[3, 3 * r + c])
          for c in range(3)] for r in range(3)]
    det = (_pmul(_pmul(E[0][0], E[1][1]), E[2][2]) + _pmul(_pmul(E[0][1], E[1][2]), E[2][0])
           + _pmul(_pmul(E[0][2], E[1][0]), E[2][1]) - _pmul(_pmul(E[0][2], E[1][1]), E[2][0])
           - _pmul(_pmul(E[0][1], E[1][0]), E[2][2]) - _pmul(_pmul(E[0][0], E[1][2]), E[2][1]))
    EEt = [[sum(_pmul(E[r][k], E[c][k]) for k in range(3)) for c in range(3)] for r in range(3)]
    tr = EEt[0][0] + EEt[1][1] + EEt[2][2]
    rows = [det]
    for r in range(3):
        for c in range(3):
            acc = np.zeros((4, 4, 4))
            for k in range(3):
                lam = EEt[r][k] - (0.5 * tr if r == k else 0.0)
                acc = acc + _pmul(lam, E[k][c])
            rows.append(acc)
    A = np.zeros((10, 20))
    for i, p in enumerate(rows):
        for j, m in enumerate(_MONO):
            A[i, j] = p[m]
    return A


def _constraint_residual(E):
    return np.concatenate([[np.linalg.det(E)], (E @ E.T @ E - 0.5 * np.trace(E @ E.T) * E).ravel()])


def _polish(basis, xyz, iters=4):
    """Gauss-Newton on the ten cubic constraints in (x, y, z): the degree-10 polynomial route loses up
    to ~1e-6 on ill-conditioned samples (cv2's own models show residuals up to 1e-3 there); the exact
    solution set is the parity target, so both the oracle and the CUDA solver polish their roots."""
    B = basis.reshape(4, 3, 3)
    E_of = lambda v: v[0] * B[0] + v[1] * B[1] + v[2] * B[2] + B[3]
    r0 = _constraint_residual(E_of(xyz))
    for _ in range(iters):
        E = E_of(xyz)
        J = np.zeros((10, 3))
        cof = np.linalg.det(E) * np.linalg.inv(E).T if abs(np.linalg.det(E)) > 1e-300 else None
        for k in range(3):
            D = B[k]
            if cof is not None:
                J[0, k] = np.sum(cof * D)
            else:   # det == 0 exactly: adjugate by finite difference is not needed, use the cofactors
                c = np.array([[np.linalg.det(np.delete(np.delete(E, i, 0), j, 1)) * (-1) ** (i + j)
                               for j in range(3)] for i in range(3)])
                J[0, k] = np.sum(c * D)
            dT = D @ E.T @ E + E @ D.T @ E + E @ E.T @ D - np.trace(E @ D.T) * E - 0.5 * np.trace(E @ E.T) * D
            J[1:, k] = dT.ravel()
        d = np.linalg.lstsq(J, -r0, rcond=None)[0]
        trial = xyz + d
        r1 = _constraint_residual(E_of(trial))
        if not (r1 @ r1 < r0 @ r0):
            break
        xyz, r0 = trial, r1
    return xyz


_G1 = np.array([0.9, -0.5, 0.3, 0.7, -0.2, 0.6, -0.8, 0.4, 0.1])
_G2 = np.array([0.2, 0.7, -0.6, -0.1, 0.9, 0.3, 0.5, -0.4, 0.8])


def canonical_models(Es):
    """Canonical sign (<E,g1> > 0) and order (ascending <E,g2>) of a model set: neither cv2's nor
    numpy's root order is canonical, and selection among equal counts depends on the order."""
    Es = np.asarray(Es, dtype=np.float64).reshape(-1, 3, 3).copy()
    for i in range(len(Es)):
        if Es[i].ravel() @ _G1 < 0:
            Es[i] = -Es[i]
    order = np.argsort([e.ravel() @ _G2 for e in Es], kind="stable")
    return Es[order]


def five_point(x1, x2, polish=False):
    """All essential matrices (unit Frobenius norm) with x2^T E x1 = 0 for 5 correspondences.
    x1, x2 [5,2] float64 normalised coordinates.  Returns array [k,3,3], k <= 10.
    polish=False is cv2's published algorithm; polish=True adds the root refinement and the
    canonical sign/order that the CUDA solver is checked against."""
    x1 = np.asarray(x1, dtype=np.float64)
    x2 = np.asarray(x2, dtype=np.float64)
    Q = np.stack([x2[:, 0] * x1[:, 0], x2[:, 0] * x1[:, 1], x2[:, 0],
                  x2[:, 1] * x1[:, 0], x2[:, 1] * x1[:, 1], x2[:, 1],
                  x1[:, 0], x1[:, 1], np.ones(len(x1))], axis=1)   # row-major vec(E)
    _, _, Vt = np.linalg.svd(Q, full_matrices=True)
    basis = Vt[5:9]
    A = _constraints(basis)
    try:
        G = np.linalg.solve(A[:, :10], A[:, 10:])
    except np.linalg.LinAlgError:
        return np.zeros((0, 3, 3))
    # rows of [I | G] are labelled by their leading monomial; k = <x^2 z> - z <x^2> and so on
    B = [[None] * 3 for _ in range(3)]
    for i, (ra, rb) in enumerate(((4, 5), (6, 7), (8, 9))):
        a, b = G[ra], G[rb]
        # columns: 0..2 = x*(z^2,z,1), 3..5 = y*(z^2,z,1), 6..9 = (z^3,z^2,z,1); polys highest power first
        B[i][0] = np.concatenate([[0.0], a[0:3]]) - np.concatenate([b[0:3], [0.0]])
        B[i][1] = np.concatenate([[0.0], a[3:6]]) - np.concatenate([b[3:6], [0.0]])
        B[i][2] = np.concatenate([[0.0], a[6:10]]) - np.concatenate([b[6:10], [0.0]])
    pm = np.polymul
    det = (pm(pm(B[0][0], B[1][1]), B[2][2]) + pm(pm(B[0][1], B[1][2]), B[2][0])
           + pm(pm(B[0][2], B[1][0]), B[2][1]) - pm(pm(B[0][2], B[1][1]), B[2][0])
           - pm(pm(B[0][1], B[1][0]), B[2][2]) - pm(pm(B[0][0], B[1][2]), B[2][1]))
    det = det[-11:] if len(det) >= 11 else np.concatenate([np.zeros(11 - len(det)), det])
    if not np.all(np.isfinite(det)) or np.all(det == 0):
        return np.zeros((0, 3, 3))
    roots = np.roots(det)
    out = []
    for r in roots:
        if abs(r.imag) > 1e-10:
            continue
        z = r.real
        Bz = np.array([[np.polyval(B[i][j], z) for j in range(3)] for i in range(3)])
        xy1 = np.linalg.svd(Bz)[2][2]
        if abs(xy1[2]) < 1e-10:
            continue
        x, y = xy1[0] / xy1[2], xy1[1] / xy1[2]
        if polish:
            x, y, z = _polish(basis, np.array([x, y, z]))
        e = x * basis[0] + y * basis[1] + z * basis[2] + basis[3]
        out.append((e / np.linalg.norm(e)).reshape(3, 3))
    out = np.array(out).reshape(-1, 3, 3)
    return canonical_models(out) if polish else out


# ---- scoring / RANSAC (EMEstimatorCallback::computeError, RANSACPointSetRegistrator::run) ---------
def sampson_err_f32(E, x1, x2):
    """err_i exactly as cv2 stores it: double arithmetic, float32 result."""
    x1 = np.asarray(x1, dtype=np.float64)
    x2 = np.asarray(x2, dtype=np.float64)
    E = np.asarray(E, dtype=np.float64).reshape(3, 3)
    Ex1_0 = E[0, 0] * x1[:, 0] + E[0, 1] * x1[:, 1] + E[0, 2]
    Ex1_1 = E[1, 0] * x1[:, 0] + E[1, 1] * x1[:, 1] + E[1, 2]
    Ex1_2 = E[2, 0] * x1[:, 0] + E[2, 1] * x1[:, 1] + E[2, 2]
    Etx2_0 = E[0, 0] * x2[:, 0] + E[1, 0] * x2[:, 1] + E[2, 0]
    Etx2_1 = E[0, 1] * x2[:, 0] + E[1, 1] * x2[:, 1] + E[2, 1]
    x2tEx1 = x2[:, 0] * Ex1_0 + x2[:, 1] * Ex1_1 + Ex1_2
    with np.errstate(divide="ignore", invalid="ignore"):
        err = x2tEx1 * x2tEx1 / (Ex1_0 * Ex1_0 + Ex1_1 * Ex1_1 + Etx2_0 * Etx2_0 + Etx2_1 * Etx2_1)
    return err.astype(np.float32)


def inlier_mask(E, x1, x2, thr):
    return sampson_err_f32(E, x1, x2) <= np.float32(thr * thr)


def score(models, n_models, x1, x2, thr):
    """models [H,10,3,3], n_models [H] -> counts [H,10] (0 for absent models)."""
    counts = np.zeros(models.shape[:2], dtype=np.int32)
    for h in range(models.shape[0]):
        for m in range(int(n_models[h])):
            counts[h, m] = int(inlier_mask(models[h, m], x1, x2, thr).sum())
    return counts


def select(counts, n_models, n_points, confidence=0.999, max_iters=None):
    """The registrator's loop over a fixed hypothesis list: returns (sample index, model index) or
    (-1, -1).  Models of one sample are visited in solver order."""
    niters = len(counts) if max_iters is None else min(max_iters, len(counts))
    best, best_count = (-1, -1), 0
    it = 0
    while it < niters:
        for m in range(int(n_models[it])):
            c = int(counts[it, m])
            if c > max(best_count, MODEL_POINTS - 1):
                best, best_count = (it, m), c
                niters = ransac_update_num_iters(confidence, (n_points - c) / n_points, MODEL_POINTS, niters)
        it += 1
    return best, best_count


def draw_samples(words, n, k=MODEL_POINTS):
    from .pnp import draw_samples as _d
    return _d(words, n, k)


def find_essential_ransac(x1, x2, samples, thr, confidence=0.999):
    """Restated cv2.findEssentialMat(RANSAC) on a given sample list.  x1, x2 float32/64 [N,2]."""
    x1 = np.asarray(x1).astype(np.float64)
    x2 = np.asarray(x2).astype(np.float64)
    H = len(samples)
    models = np.zeros((H, 10, 3, 3))
    n_models = np.zeros(H, dtype=np.int32)
    for h, s in enumerate(samples):
        Es = five_point(x1[s], x2[s], polish=True)
        n_models[h] = len(Es)
        models[h, :len(Es)] = Es
    counts = score(models, n_models, x1, x2, thr)
    (bh, bm), cnt = select(counts, n_models, len(x1), confidence)
    if bh < 0:
        return {"ok": False, "models": models, "n_models": n_models, "counts": counts, "best": (-1, -1)}
    E = models[bh, bm]
    return {"ok": True, "E": E, "mask": inlier_mask(E, x1, x2, thr), "best": (bh, bm), "models": models,
            "n_models": n_models, "counts": counts}


# ---- K4: epiline score of helper_functions.py:54-68 ------------------------------------------------
def _epilines_image2_f32(pts_f32, E):
    """cv2.computeCorrespondEpilines(pts, 2, E) for float32 points: l = E^T [x,y,1], scaled so that
    a^2+b^2 = 1, double arithmetic, float32 result."""
    p = np.asarray(pts_f32, dtype=np.float32).astype(np.float64)
    Et = np.asarray(E, dtype=np.float64).reshape(3, 3).T
    a = Et[0, 0] * p[:, 0] + Et[0, 1] * p[:, 1] + Et[0, 2]
    b = Et[1, 0] * p[:, 0] + Et[1, 1] * p[:, 1] + Et[1, 2]
    c = Et[2, 0] * p[:, 0] + Et[2, 1] * p[:, 1] + Et[2, 2]
    nu = a * a + b * b
    nu = np.where(nu != 0, 1.0 / np.sqrt(np.where(nu != 0, nu, 1.0)), 1.0)
    return np.stack([a * nu, b * nu, c * nu], axis=1).astype(np.float32)


def epiline_score(E, x1_f32, x2_f32, mask):
    """score of estimateEssential (helper_functions.py:54-68): both line sets use whichImage=2."""
    m = np.asarray(mask).astype(bool).ravel()
    p1 = np.asarray(x1_f32, dtype=np.float32)[m]
    p2 = np.asarray(x2_f32, dtype=np.float32)[m]
    loc1 = np.concatenate([p1.astype(np.float64), np.ones((len(p1), 1))], axis=1)
    loc2 = np.concatenate([p2.astype(np.float64), np.ones((len(p2), 1))], axis=1)
    l1 = _epilines_image2_f32(p2, E).astype(np.float64)
    e2in1 = np.sum(loc1 * l1, axis=1) ** 2 / np.sum(l1 ** 2, axis=1)
    l2 = _epilines_image2_f32(p1, E).astype(np.float64)
    e1in2 = np.sum(loc2 * l2, axis=1) ** 2 / np.sum(l2 ** 2, axis=1)
    return float(np.sum(np.maximum(4 - e1in2, 0)) + np.sum(np.maximum(4 - e2in1, 0)))


# ---- K5: cv2.recoverPose -----------------------------------------------------------------------------
def decompose_essential(E):
    U, _, Vt = np.linalg.svd(np.asarray(E, dtype=np.float64).reshape(3, 3))
    if np.linalg.det(U) < 0:
        U = -U
    if np.linalg.det(Vt) < 0:
        Vt = -Vt
    W = np.array([[0.0, 1, 0], [-1, 0, 0], [0, 0, 1]])
    return U @ W @ Vt, U @ W.T @ Vt, U[:, 2].copy()


def triangulate_dlt(P0, P1, x1, x2):
    """cv2.triangulatePoints: per point the right singular vector of the smallest singular value of
    [x P[2]-P[0]; y P[2]-P[1]] stacked for both views.  Returns [4,N] (unit columns, sign free)."""
    n = len(x1)
    out = np.zeros((4, n))
    for i in range(n):
        A = np.stack([x1[i, 0] * P0[2] - P0[0], x1[i, 1] * P0[2] - P0[1],
                      x2[i, 0] * P1[2] - P1[0], x2[i, 1] * P1[2] - P1[1]])
        out[:, i] = np.linalg.svd(A)[2][3]
    return out


def _cheirality(P1, Q, dist):
    with np.errstate(divide="ignore", invalid="ignore"):
        mask = (Q[2] * Q[3]) > 0
        Qn = Q / Q[3]
        mask &= Qn[2] < dist
        Q2 = P1 @ Qn
        mask &= (Q2[2] > 0) & (Q2[2] < dist)
    return mask


def recover_pose(E, p1, p2, K, distance_thresh=50.0):
    """Returns (n_good, R, t[3], mask bool[N], tri [4,N]) like cv2.recoverPose(E, p1, p2, K, dist)."""
    fx, fy, cx, cy = [float(v) for v in K]
    a = np.asarray(p1).astype(np.float64)
    b = np.asarray(p2).astype(np.float64)
    x1 = np.stack([(a[:, 0] - cx) / fx, (a[:, 1] - cy) / fy], axis=1)
    x2 = np.stack([(b[:, 0] - cx) / fx, (b[:, 1] - cy) / fy], axis=1)
    R1, R2, t = decompose_essential(E)
    P0 = np.eye(3, 4)
    cands = [(R1, t), (R2, t), (R1, -t), (R2, -t)]
    best = None
    for R, tt in cands:
        P1 = np.concatenate([R, tt[:, None]], axis=1)
        Q = triangulate_dlt(P0, P1, x1, x2)
        mask = _cheirality(P1, Q, distance_thresh)
        if best is None or int(mask.sum()) > best[0]:     # cv2 takes the first maximum in this order
            best = (int(mask.sum()), R, tt, mask, Q)
    return best
