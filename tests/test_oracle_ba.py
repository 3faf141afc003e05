"""CPU: internal-consistency checks of the BA restatement (parity is UNPINNED: g2o is absent, see
oracle/ba.py).  Analytic Jacobians vs central differences through the g2o pose increment; converged
robust cost vs an independent scipy least_squares; LM bookkeeping; fixture F1 (debug.txt)."""
import os

import numpy as np
import pytest
import scipy.linalg
import scipy.optimize

from oracle import ba
from visual_slam_b200.ba_problem import BAProblem, R_to_quat, quat_to_R, synthetic_ba


def small(seed=0, **kw):
    return synthetic_ba(4, 60, 3.0, seed=seed, **kw)


def test_quaternion_roundtrip():
    rng = np.random.default_rng(0)
    for _ in range(20):
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        R = quat_to_R(q)
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-12)
        q2 = R_to_quat(R)
        assert np.allclose(quat_to_R(q2), R, atol=1e-12)
        assert np.allclose(ba.quat_to_R(q[None])[0], R, atol=1e-15)


def test_jacobians_match_central_differences_through_g2o_increment():
    p = small(1)
    Jc, Jp = ba.jacobians(p)
    e0, _ = ba.residuals(p)
    h = 1e-6
    for k in range(6):
        d = np.zeros((p.n_cams, 6)); d[:, k] = h
        ep, _ = ba.residuals(p, cams=ba.apply_cam_update(p.cams, d))
        em, _ = ba.residuals(p, cams=ba.apply_cam_update(p.cams, -d))
        num = (ep - em) / (2 * h)
        assert np.allclose(num, Jc[:, :, k], rtol=2e-5, atol=2e-4), k
    for k in range(3):
        d = np.zeros_like(p.pts); d[:, k] = h
        ep, _ = ba.residuals(p, pts=p.pts + d)
        em, _ = ba.residuals(p, pts=p.pts - d)
        assert np.allclose((ep - em) / (2 * h), Jp[:, :, k], rtol=2e-5, atol=2e-4)


def test_gradient_of_cost_is_minus_2b_where_huber_is_smooth():
    p = small(2)
    L = ba.linearize(p)
    h = 1e-6
    free = np.nonzero(~p.cam_fixed)[0]
    g = np.zeros(6 * len(free))
    for i, c in enumerate(free):
        for k in range(6):
            d = np.zeros((p.n_cams, 6)); d[c, k] = h
            g[6 * i + k] = (ba.cost(p, cams=ba.apply_cam_update(p.cams, d)) -
                            ba.cost(p, cams=ba.apply_cam_update(p.cams, -d))) / (2 * h)
    assert np.allclose(g, -2 * L.bp, rtol=1e-4, atol=1e-2 * np.abs(L.bp).max() * 1e-2)


def test_robust_kernels():
    d2 = 5.991
    rho, w = ba.huber(np.array([0.0, d2, 4 * d2]))
    assert np.allclose(rho, [0, d2, 2 * np.sqrt(4 * d2) * np.sqrt(d2) - d2]) and np.allclose(w, [1, 1, 0.5])
    rho, w = ba.dcs(np.array([0.5, 1.0, 3.0]))
    assert np.allclose(rho, [0.5, 1.0, 0.25 * 3.0]) and np.allclose(w, [1, 1, 0.25])


def test_lm_reduces_cost_and_matches_independent_optimum():
    p = synthetic_ba(4, 40, 3.0, seed=3, outlier_frac=0.0, with_scale=False)
    # a second fixed camera at its true pose removes the monocular scale gauge
    p.cams[1] = synthetic_ba(4, 40, 3.0, seed=3, pose_noise=(0, 0)).cams[1]
    p.cam_fixed[1] = True
    q = p.copy()
    h = ba.optimize(q, max_iterations=100)
    assert h["final_cost"] < 0.1 * h["cost"][0] and all(np.diff(h["cost"]) <= 1e-9)
    # independent: MINPACK LM (numeric Jacobian) on the same Huber cost and state parametrisation
    free_c = np.nonzero(~p.cam_fixed)[0]

    def fun(x):
        cams = p.cams.copy()
        cams[free_c] = ba.apply_cam_update(p.cams[free_c], x[:6 * len(free_c)].reshape(-1, 6))
        e, _ = ba.residuals(p, cams, p.pts + x[6 * len(free_c):].reshape(-1, 3))
        chi2 = np.sum(e * e, axis=1)
        rho, _ = ba.huber(chi2)
        return (e * np.sqrt(rho / np.maximum(chi2, 1e-300))[:, None]).reshape(-1)

    x0 = np.zeros(6 * len(free_c) + 3 * p.n_pts)
    r = scipy.optimize.least_squares(fun, x0, method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15,
                                     max_nfev=20000)
    assert abs(2 * r.cost - h["final_cost"]) <= 1e-6 * h["final_cost"]


def test_ten_iteration_budget_and_lambda_schedule():
    p = small(4)
    h = ba.optimize(p, max_iterations=10)
    assert h["iterations"] <= 10 and len(h["lambda"]) == h["iterations"]
    assert h["trials"] >= h["iterations"]


def test_motion_only_is_block_diagonal():
    p = synthetic_ba(5, 80, 4.0, seed=5, motion_only=True)
    assert p.n_free_pts == 0
    L = ba.linearize(p)
    n = L.Hpp.shape[0]
    mask = np.kron(np.eye(n // 6), np.ones((6, 6))).astype(bool)
    assert np.all(L.Hpp[~mask] == 0)
    h = ba.optimize(p, 10)
    assert h["final_cost"] < h["cost"][0]


def test_fixture_f1_reduced_system_is_rejected(golden_dir):
    S = np.load(os.path.join(golden_dir, "ba_debug_reduced_system.npz"))["S"]
    assert S.shape == (90, 90) and np.array_equal(S, S.T)
    w = np.linalg.eigvalsh(S)
    assert (w < 0).sum() == 1
    with pytest.raises(np.linalg.LinAlgError):
        scipy.linalg.cho_factor(S, lower=True)
    L = ba.Linearization()
    L.Hpp, L.Hll, L.bp, L.bl = S, np.zeros((0, 3, 3)), np.ones(90), np.zeros((0, 3))
    ok, dp, dl = ba.solve(L, 0.0)
    assert not ok
