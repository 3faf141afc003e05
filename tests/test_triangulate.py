"""Two-view triangulation: oracle vs the reference's own outputs (CPU), CUDA kernel vs oracle (GPU)."""
import os

import numpy as np
import pytest

from oracle import triangulate as otri

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "triangulate_ref.npz")


def _cases():
    g = np.load(GOLD)
    for c in range(3):
        yield {k: g["%s_%d" % (k, c)] for k in ("P1", "P2", "T1", "T2", "x1", "x2", "tri")}


def _dehom(X):
    return X[:, :3] / X[:, 3:]


def test_oracle_reproduces_reference_outputs():
    for c in _cases():
        mine = otri.triangulate(c["P1"], c["P2"], c["x1"], c["x2"])
        # same LAPACK underneath: identical up to the sign convention of the singular vector
        s = np.sign(np.sum(mine * c["tri"], axis=1))[:, None]
        assert np.abs(mine * s - c["tri"]).max() < 1e-12


def test_filter_matches_reference_expression():
    c = list(_cases())[2]
    X3, depth, good = otri.dehomogenise_and_filter(c["tri"], c["T1"], c["T2"])
    X = c["tri"] / c["tri"][:, 3:]
    proj1 = np.vstack([c["T1"], [0, 0, 0, 1]]) @ X.T
    proj2 = np.vstack([c["T2"], [0, 0, 0, 1]]) @ X.T
    ref = np.where((proj1[2] > 0) & (proj2[2] > 0) & (proj2[2] < 1) & (proj1[2] < 1))[0]   # main.py:306
    assert np.array_equal(np.nonzero(good)[0], ref)
    assert 0 < len(ref) < len(X)


@pytest.mark.gpu
def test_kernel_matches_reference_and_oracle():
    from visual_slam_b200.triangulate import triangulate, triangulate_and_filter
    for c in _cases():
        Xh = triangulate(c["P1"], c["P2"], c["x1"], c["x2"])
        assert Xh.shape == c["tri"].shape
        assert np.abs(np.linalg.norm(Xh, axis=1) - 1).max() < 1e-12
        ref = _dehom(c["tri"])
        assert np.abs(_dehom(Xh) - ref).max() <= 1e-9 * max(1.0, np.abs(ref).max())
        X3, idx = triangulate_and_filter(c["P1"], c["P2"], c["T1"], c["T2"], c["x1"], c["x2"])
        _, _, good = otri.dehomogenise_and_filter(c["tri"], c["T1"], c["T2"])
        assert np.array_equal(idx, np.nonzero(good)[0])
        assert np.abs(X3 - ref).max() <= 1e-9 * max(1.0, np.abs(ref).max())


@pytest.mark.gpu
def test_batched_ragged_and_empty():
    import torch
    from visual_slam_b200.triangulate import triangulate_batch
    cs = list(_cases())
    n_max = max(len(c["x1"]) for c in cs)
    B = len(cs) + 1                      # last pair is empty
    proj = np.zeros((B, 2, 3, 4)); extr = np.zeros((B, 2, 3, 4))
    x1 = np.zeros((B, n_max, 2)); x2 = np.zeros((B, n_max, 2)); n = np.zeros(B, dtype=np.int32)
    for b, c in enumerate(cs):
        k = len(c["x1"])
        proj[b, 0], proj[b, 1], extr[b, 0], extr[b, 1] = c["P1"], c["P2"], c["T1"], c["T2"]
        x1[b, :k], x2[b, :k], n[b] = c["x1"], c["x2"], k
    proj[-1] = proj[0]; extr[-1] = extr[0]
    d = lambda a: torch.from_numpy(a).cuda()
    Xh, X3, depth, good = triangulate_batch(d(proj), d(x1), d(x2), d(n), d(extr), 1.0)
    for b, c in enumerate(cs):
        k = len(c["x1"])
        ref = _dehom(c["tri"])
        assert np.abs(X3[b, :k].cpu().numpy() - ref).max() <= 1e-9 * max(1.0, np.abs(ref).max())
        _, dep, g = otri.dehomogenise_and_filter(c["tri"], c["T1"], c["T2"])
        assert np.array_equal(good[b, :k].cpu().numpy().astype(bool), g)
        assert np.abs(depth[b, :k].cpu().numpy() - dep).max() < 1e-8
        assert not good[b, k:].any()       # padding untouched
    assert not Xh[-1].any() and not good[-1].any()


@pytest.mark.gpu
def test_no_correspondences_returns_an_empty_array_like_the_reference():
    """helper_functions.py:281-291 on N = 0 returns np.array([]) reshaped by its caller; the drop-in gives (0, 4)."""
    from visual_slam_b200.triangulate import triangulate, triangulate_and_filter
    P = np.hstack([np.eye(3), np.zeros((3, 1))])
    out = triangulate(P, P, np.zeros((0, 2)), np.zeros((0, 2)))
    assert out.shape == (0, 4)
    X3, good = triangulate_and_filter(P, P, P, P, np.zeros((0, 2)), np.zeros((0, 2)))
    assert X3.shape == (0, 3) and len(good) == 0


def test_shape_errors_raise_before_any_device_work():
    from visual_slam_b200.triangulate import triangulate
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(Exception):
            triangulate(np.eye(3, 4), np.eye(3, 4), np.zeros((4, 2)), np.zeros((4, 2)))
    with pytest.raises(ValueError):
        from visual_slam_b200.triangulate import _one
        _one(np.eye(4, 4)[:2], np.eye(3, 4), np.zeros((4, 2)), np.zeros((4, 2)), None, 1.0, "cpu")


# The reference's only known-answer script for triangulation (src/v1/testing.py:46-97): hard-coded projection
# matrices and pixels, recorded stdout "[-0.035 -0.006 2.022]" for cv2.triangulatePoints on the first pair and the
# two-point set x1h2 / x2h2.  The numbers below are the script's literals.
_V1_P1 = np.array([[5.010e+03, 0.000e+00, 3.600e+02, 0.000e+00],
                   [0.000e+00, 5.010e+03, 6.400e+02, 0.000e+00],
                   [0.000e+00, 0.000e+00, 1.000e+00, 0.000e+00]])
_V1_P2 = np.array([[5.037e+03, -9.611e+01, -1.756e+03, 4.284e+03],
                   [2.148e+02, 5.354e+03, 1.918e+02, 8.945e+02],
                   [3.925e-01, 7.092e-02, 9.169e-01, 4.930e-01]])
_V1_X1 = np.array([[274.128, 624.409], [274.129, 624.409], [322.527, 624.869]])
_V1_X2 = np.array([[239.571, 533.568], [239.572, 533.568], [284.507, 534.572]])


def test_oracle_reproduces_the_v1_known_answer():
    import cv2
    X = _dehom(otri.triangulate(_V1_P1, _V1_P2, _V1_X1, _V1_X2))
    assert np.allclose(np.round(X[0], 3), [-0.035, -0.006, 2.022])          # the recorded stdout
    ref = cv2.triangulatePoints(_V1_P1, _V1_P2, _V1_X1.T, _V1_X2.T)          # what the script calls
    ref = (ref[:3] / ref[3]).T
    # the reference's DLT (helper_functions.py:281-291) and cv2's differ in row scaling only: same point to 1e-4 m
    assert np.abs(X - ref).max() < 1e-4


@pytest.mark.gpu
def test_kernel_reproduces_the_v1_known_answer():
    from visual_slam_b200.triangulate import triangulate
    X = _dehom(triangulate(_V1_P1, _V1_P2, _V1_X1, _V1_X2))
    assert np.allclose(np.round(X[0], 3), [-0.035, -0.006, 2.022])
    assert np.abs(X - _dehom(otri.triangulate(_V1_P1, _V1_P2, _V1_X1, _V1_X2))).max() < 1e-9
