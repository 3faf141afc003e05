"""CPU restatement of the reference optimiser for bundle adjustment.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the reference optimiser is g2o reached through the `g2o` (g2opy) wheel
(reference src/v2/LocalBA.py:20-131), which is not vendored in /root/reference, carries no
pinned version, and is not installable here.  The reference holds no golden vector for BA.  This
file restates the published g2o algorithm that LocalBA.py's call sites select:

  graph semantics      LocalBA.py:56-65 (add_pose: VertexCam / SBACam, camera->world),
                       :68-77 (add_point, marginalised), :79-94 (add_edge: EdgeProjectP2MC,
                       information I2, Huber sqrt(5.991)), :115-131 (AddScalingEdge: EdgeSBAScale,
                       DCS), :39-42 (optimize: 10 Levenberg iterations)
  arithmetic           g2o types/sba (SBACam::update, EdgeProjectP2MC::computeError /
                       linearizeOplus), core/robust_kernel_impl (Huber, DCS),
                       core/block_solver (Schur), core/optimization_algorithm_levenberg
                       -- restated in SURVEY.md 8a-3.

It is validated by internal consistency only (tests/test_oracle_ba.py): analytic Jacobians vs
central differences through the g2o pose increment, and the converged robust cost vs an
independent scipy.optimize.least_squares run on the same cost.  One deliberate deviation: the
EdgeSBAScale Jacobian is analytic here (g2o uses central differences with delta 1e-9).
"""
import numpy as np
import scipy.linalg
import scipy.sparse as sp

HUBER_DELTA = np.sqrt(5.991)  # LocalBA.py:82
DCS_PHI = 1.0                 # g2o RobustKernelDCS default delta, LocalBA.py:115


def quat_to_R(q):
    """q = (x, y, z, w) (Eigen coefficient order) -> rotation matrices [...,3,3]."""
    q = np.asarray(q, dtype=np.float64)
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    R = np.empty(q.shape[:-1] + (3, 3))
    R[..., 0, 0] = 1 - 2 * (y * y + z * z); R[..., 0, 1] = 2 * (x * y - z * w); R[..., 0, 2] = 2 * (x * z + y * w)
    R[..., 1, 0] = 2 * (x * y + z * w); R[..., 1, 1] = 1 - 2 * (x * x + z * z); R[..., 1, 2] = 2 * (y * z - x * w)
    R[..., 2, 0] = 2 * (x * z - y * w); R[..., 2, 1] = 2 * (y * z + x * w); R[..., 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def quat_mul(a, b):
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], axis=-1)


def apply_cam_update(cams, delta):
    """SBACam::update: t += d[0:3]; q <- normalize(q * (sqrt(1-|v|^2), v)), v = d[3:6]."""
    out = cams.copy()
    out[:, 4:7] += delta[:, 0:3]
    v = delta[:, 3:6]
    w = np.sqrt(1.0 - np.sum(v * v, axis=1))
    qn = quat_mul(cams[:, 0:4], np.concatenate([v, w[:, None]], axis=1))
    out[:, 0:4] = qn / np.linalg.norm(qn, axis=1, keepdims=True)
    return out


_A = np.array([[[0, 0, 0], [0, 0, 2], [0, -2, 0]],
               [[0, 0, -2], [0, 0, 0], [2, 0, 0]],
               [[0, 2, 0], [-2, 0, 0], [0, 0, 0]]], dtype=np.float64)  # dRidx, dRidy, dRidz


def residuals(prob, cams=None, pts=None):
    """EdgeProjectP2MC::computeError for every observation: projected - measured, [N,2]."""
    cams = prob.cams if cams is None else cams
    pts = prob.pts if pts is None else pts
    Rt = np.swapaxes(quat_to_R(cams[:, 0:4]), 1, 2)  # w2n rotation = R^T
    X = pts[prob.obs_pt]
    pc = np.einsum("nij,nj->ni", Rt[prob.obs_cam], X - cams[prob.obs_cam, 4:7])
    fx, fy, cx, cy = prob.K
    proj = np.stack([fx * pc[:, 0] / pc[:, 2] + cx, fy * pc[:, 1] / pc[:, 2] + cy], axis=1)
    return proj - prob.obs_uv, pc


def jacobians(prob, cams=None, pts=None):
    """EdgeProjectP2MC::linearizeOplus: Jc [N,2,6] (3 translation, 3 rotation), Jp [N,2,3]."""
    cams = prob.cams if cams is None else cams
    pts = prob.pts if pts is None else pts
    fx, fy, _, _ = prob.K
    Rt = np.swapaxes(quat_to_R(cams[:, 0:4]), 1, 2)[prob.obs_cam]  # [N,3,3]
    pwt = pts[prob.obs_pt] - cams[prob.obs_cam, 4:7]
    pc = np.einsum("nij,nj->ni", Rt, pwt)
    px, py, pz = pc[:, 0], pc[:, 1], pc[:, 2]
    ipz2 = 1.0 / (pz * pz)

    def g(dp):  # dp [N,3,k] -> [N,2,k]
        return np.stack([(pz[:, None] * dp[:, 0] - px[:, None] * dp[:, 2]) * (ipz2 * fx)[:, None],
                         (pz[:, None] * dp[:, 1] - py[:, None] * dp[:, 2]) * (ipz2 * fy)[:, None]], axis=1)

    Jp = g(Rt)                        # d/dX: +w2n columns
    Jt = g(-Rt)                       # d/dt: -w2n columns
    dR = np.einsum("kij,njl->nkil", _A, Rt)            # dRdk = A_k R^T
    drot = np.einsum("nkil,nl->nik", dR, pwt)          # [N,3,k]
    Jr = g(drot)
    return np.concatenate([Jt, Jr], axis=2), Jp


def huber(chi2, delta=HUBER_DELTA):
    d2 = delta * delta
    s = np.sqrt(np.maximum(chi2, 1e-300))
    rho = np.where(chi2 <= d2, chi2, 2 * s * delta - d2)
    w = np.where(chi2 <= d2, 1.0, delta / s)
    return rho, w


def dcs(chi2, phi=DCS_PHI):
    s = 2.0 * phi / (phi + chi2)
    rho = np.where(s >= 1.0, chi2, s * s * chi2)
    w = np.where(s >= 1.0, 1.0, s * s)
    return rho, w


def scale_residuals(prob, cams=None):
    """EdgeSBAScale::computeError: meas - |t_b - t_a|."""
    cams = prob.cams if cams is None else cams
    if prob.n_scale == 0:
        return np.zeros(0), np.zeros((0, 3))
    d = cams[prob.scale_b, 4:7] - cams[prob.scale_a, 4:7]
    n = np.linalg.norm(d, axis=1)
    return prob.scale_meas - n, d / n[:, None]


def cost(prob, cams=None, pts=None):
    """activeRobustChi2: sum of rho over all edges."""
    e, _ = residuals(prob, cams, pts)
    rho, _ = huber(np.sum(e * e, axis=1))
    es, _ = scale_residuals(prob, cams)
    rs, _ = dcs(es * es)
    return float(np.sum(rho) + np.sum(rs))


class Linearization:
    pass


def linearize(prob):
    """buildSystem(): Hpp (dense 6Cf x 6Cf incl. scale-edge off-diagonals), Hll [Pf,3,3],
    Hpl as a sparse 6Cf x 3Pf matrix, bp, bl and the robust cost.  Fixed vertices get no column."""
    L = Linearization()
    cam_col = prob.cam_col   # [C] -> free index or -1
    pt_col = prob.pt_col     # [P] -> free index or -1
    Cf, Pf = prob.n_free_cams, prob.n_free_pts
    e, _ = residuals(prob)
    Jc, Jp = jacobians(prob)
    rho, w = huber(np.sum(e * e, axis=1))
    L.cost = float(np.sum(rho))
    cc = cam_col[prob.obs_cam]
    pc = pt_col[prob.obs_pt]
    Hpp = np.zeros((Cf, 6, 6)); bp = np.zeros((Cf, 6))
    m = cc >= 0
    np.add.at(Hpp, cc[m], np.einsum("n,nki,nkj->nij", w[m], Jc[m], Jc[m]))
    np.add.at(bp, cc[m], -np.einsum("n,nki,nk->ni", w[m], Jc[m], e[m]))
    Hll = np.zeros((Pf, 3, 3)); bl = np.zeros((Pf, 3))
    mp = pc >= 0
    np.add.at(Hll, pc[mp], np.einsum("n,nki,nkj->nij", w[mp], Jp[mp], Jp[mp]))
    np.add.at(bl, pc[mp], -np.einsum("n,nki,nk->ni", w[mp], Jp[mp], e[mp]))
    both = m & mp
    Hpl_blocks = np.einsum("n,nki,nkj->nij", w[both], Jc[both], Jp[both])  # [Nb,6,3]
    L.Hpl_blocks, L.Hpl_cam, L.Hpl_pt = Hpl_blocks, cc[both], pc[both]
    Hpp_dense = np.zeros((6 * Cf, 6 * Cf))
    for c in range(Cf):
        Hpp_dense[6 * c:6 * c + 6, 6 * c:6 * c + 6] = Hpp[c]
    # scale edges (EdgeSBAScale + DCS): 1-D residual, Jacobian only on the translations
    es, u = scale_residuals(prob)
    rs, ws = dcs(es * es)
    L.cost += float(np.sum(rs))
    for k in range(prob.n_scale):
        a, b = cam_col[prob.scale_a[k]], cam_col[prob.scale_b[k]]
        Ja = np.zeros(6); Jb = np.zeros(6)
        Ja[0:3] = u[k]      # d(meas - |tb - ta|)/d ta = +u
        Jb[0:3] = -u[k]
        for (i, Ji) in ((a, Ja), (b, Jb)):
            if i < 0:
                continue
            bp[i] += -ws[k] * Ji * es[k]
            for (j, Jj) in ((a, Ja), (b, Jb)):
                if j < 0:
                    continue
                Hpp_dense[6 * i:6 * i + 6, 6 * j:6 * j + 6] += ws[k] * np.outer(Ji, Jj)
    L.Hpp, L.Hll, L.bp, L.bl = Hpp_dense, Hll, bp.reshape(-1), bl
    return L


def lambda_init(L, tau=1e-5):
    """computeLambdaInit: tau * max |diag(H)| over all free vertices."""
    mx = 0.0
    if L.Hpp.size:
        mx = max(mx, float(np.max(np.abs(np.diag(L.Hpp)))))
    if L.Hll.size:
        mx = max(mx, float(np.max(np.abs(np.einsum("nii->ni", L.Hll)))))
    return tau * mx


def schur_system(L, lam):
    """(S, bhat, W, Hinv): S = Hpp + lam I - Hpl (Hll + lam I)^-1 Hpl^T, bhat = bp - Hpl Hll^-1 bl."""
    n = L.Hpp.shape[0]
    Pf = L.Hll.shape[0]
    Hpp = L.Hpp + lam * np.eye(n)
    if Pf == 0:
        return Hpp, L.bp, None, None
    Hinv = np.linalg.inv(L.Hll + lam * np.eye(3)[None])
    nb = L.Hpl_blocks.shape[0]
    rows = (6 * L.Hpl_cam[:, None, None] + np.arange(6)[None, :, None]) + np.zeros((1, 1, 3), dtype=np.int64)
    cols = (3 * L.Hpl_pt[:, None, None] + np.arange(3)[None, None, :]) + np.zeros((1, 6, 1), dtype=np.int64)
    W = sp.csr_matrix((L.Hpl_blocks.reshape(-1), (rows.reshape(-1), cols.reshape(-1))),
                      shape=(n, 3 * Pf)) if nb else sp.csr_matrix((n, 3 * Pf))
    WD = W @ _blockdiag(Hinv)
    S = Hpp - (WD @ W.T).toarray()
    bh = L.bp - WD @ L.bl.reshape(-1)
    return S, bh, W, Hinv


def solve(L, lam):
    """BlockSolver::solve with lambda on every diagonal entry.  Returns (ok, dp [6Cf], dl [Pf,3])."""
    n = L.Hpp.shape[0]
    Pf = L.Hll.shape[0]
    S, bh, W, Hinv = schur_system(L, lam)
    ok = True
    if n == 0:
        dp = np.zeros(0)
    else:
        try:
            c = scipy.linalg.cho_factor(S, lower=True)   # CHOLMOD stand-in: fails if not SPD
            dp = scipy.linalg.cho_solve(c, bh)
        except np.linalg.LinAlgError:
            ok, dp = False, np.zeros(n)
    if Pf == 0:
        dl = np.zeros((0, 3))
    else:
        r = L.bl.reshape(-1) - W.T @ dp
        dl = np.einsum("nij,nj->ni", Hinv, r.reshape(-1, 3))
    return ok, dp, dl


def _blockdiag(blocks):
    P = blocks.shape[0]
    r = (3 * np.arange(P)[:, None, None] + np.arange(3)[None, :, None]) + np.zeros((1, 1, 3), dtype=np.int64)
    c = (3 * np.arange(P)[:, None, None] + np.arange(3)[None, None, :]) + np.zeros((1, 3, 1), dtype=np.int64)
    return sp.csr_matrix((blocks.reshape(-1), (r.reshape(-1), c.reshape(-1))), shape=(3 * P, 3 * P))


def apply_update(prob, dp, dl):
    cams = prob.cams.copy()
    free = np.nonzero(prob.cam_col >= 0)[0]
    if len(free):
        cams[free] = apply_cam_update(prob.cams[free], dp.reshape(-1, 6))
    pts = prob.pts.copy()
    fp = np.nonzero(prob.pt_col >= 0)[0]
    if len(fp):
        pts[fp] += dl
    return cams, pts


def optimize(prob, max_iterations=10, tau=1e-5, max_trials=10, verbose=False):
    """OptimizationAlgorithmLevenberg::solve driven by SparseOptimizer::optimize (LocalBA.py:39-42).
    Mutates prob.cams / prob.pts.  Returns a dict with the cost trajectory."""
    lam, ni = 0.0, 2.0
    hist = {"cost": [cost(prob)], "lambda": [], "trials": 0, "iterations": 0}
    for it in range(max_iterations):
        L = linearize(prob)
        cur = L.cost
        if it == 0:
            lam, ni = lambda_init(L, tau), 2.0
        rho, q = 0.0, 0
        while True:
            ok, dp, dl = solve(L, lam)
            cams, pts = apply_update(prob, dp, dl)
            tmp = cost(prob, cams, pts)
            if not ok:
                tmp = np.finfo(np.float64).max
            x = np.concatenate([dp, dl.reshape(-1)])
            b = np.concatenate([L.bp, L.bl.reshape(-1)])
            scale = float(np.sum(x * (lam * x + b))) + 1e-3
            rho = (cur - tmp) / scale
            hist["trials"] += 1
            if rho > 0 and np.isfinite(tmp):
                alpha = min(1.0 - (2 * rho - 1) ** 3, 2.0 / 3.0)
                lam *= max(1.0 / 3.0, alpha)
                ni = 2.0
                cur = tmp
                prob.cams, prob.pts = cams, pts
            else:
                lam *= ni
                ni *= 2
                if not np.isfinite(lam):
                    break
            q += 1
            if not (rho < 0 and q < max_trials):
                break
        hist["cost"].append(cur); hist["lambda"].append(lam); hist["iterations"] += 1
        if verbose:
            print("it %d cost %.9g lambda %.3g trials %d" % (it, cur, lam, q))
        if q == max_trials or rho == 0 or not np.isfinite(lam):
            break
    hist["final_cost"] = cost(prob)
    e, _ = residuals(prob)
    hist["rms_px"] = float(np.sqrt(np.mean(np.sum(e * e, axis=1))))
    return hist
