"""Pose plumbing (SURVEY 8a-2 last row; reference src/v2/main.py:191-204, src/v2/helper_functions.py:269-278): the
device kernels against cv2.Rodrigues (the reference's own call) and the restated inversion.  The recorded outputs
of the reference's functions over the ICL-NUIM replay are checked in tests/test_c1_replay.py."""
import cv2
import numpy as np
import pytest

from oracle import poses as opo


def _rvecs():
    rng = np.random.default_rng(5)
    v = [rng.normal(size=3) * s for s in (1e-300, 1e-17, 1e-9, 1e-3, 0.1, 1.0, 2.0, 3.0) for _ in range(6)]
    ax = rng.normal(size=(12, 3)); ax /= np.linalg.norm(ax, axis=1, keepdims=True)
    v += [a * (np.pi - d) for a, d in zip(ax, (0, 1e-12, 1e-9, 1e-7, 1e-6, 1e-5, 1e-4, 1e-3, 1e-2, 0.1, 0.5, 1.0))]
    v += [np.zeros(3), np.array([np.pi, 0, 0]), np.array([0, np.pi, 0]), np.array([0, 0, -np.pi])]
    return np.array(v)


def test_oracle_is_the_reference_composition():
    """oracle/poses.py restates helper_functions.py:269-278 / main.py:199-203 on top of cv2.Rodrigues."""
    for rv in _rvecs()[::5]:
        T = opo.transformMatrix(rv, [0.1, -0.2, 0.3])
        assert np.array_equal(T[:3, :3], cv2.Rodrigues(rv.reshape(3, 1))[0])
        W = opo.camera_to_world(rv, [0.1, -0.2, 0.3])
        assert np.abs(W @ T - np.eye(4)).max() < 1e-14
        q = opo.sbacam(W)
        assert abs(np.linalg.norm(q[:4]) - 1) < 1e-15 and q[3] >= 0


@pytest.mark.gpu
def test_rodrigues_both_ways_match_cv2():
    import torch
    from visual_slam_b200 import poses
    rv = _rvecs()
    R = poses.rodrigues_batch(torch.from_numpy(rv).cuda()).cpu().numpy()
    Rc = np.array([cv2.Rodrigues(r.reshape(3, 1))[0] for r in rv])
    assert np.abs(R - Rc).max() < 1e-15
    # matrix -> vector on cv2's own matrices and on perturbed (non-orthonormal) ones: cv2 projects on SO(3) first
    rng = np.random.default_rng(6)
    mats = np.concatenate([Rc, Rc + rng.normal(0, 1e-3, Rc.shape), Rc * 1.01])
    back = poses.rodrigues_inverse_batch(torch.from_numpy(mats).cuda()).cpu().numpy()
    ref = np.array([cv2.Rodrigues(m)[0].ravel() for m in mats])
    th = np.linalg.norm(ref, axis=1)
    regular = th < np.pi - 1e-4                      # away from the antipodal singularity the axis is well conditioned
    assert np.abs(back[regular] - ref[regular]).max() < 1e-9
    for a, b in zip(back[~regular], ref[~regular]):  # near pi: r and -r describe (almost) the same rotation
        assert min(np.abs(a - b).max(), np.abs(a + b).max()) < 1e-6 or abs(np.linalg.norm(a) - np.linalg.norm(b)) < 1e-6
        assert np.abs(cv2.Rodrigues(a.reshape(3, 1))[0] - cv2.Rodrigues(b.reshape(3, 1))[0]).max() < 1e-6


@pytest.mark.gpu
def test_camera_to_world_guess_and_ba_state():
    import torch
    from visual_slam_b200 import poses
    rng = np.random.default_rng(7)
    rv = _rvecs()
    pose6 = np.concatenate([rv, rng.normal(size=(len(rv), 3))], axis=1)
    o = poses.camera_to_world_batch(torch.from_numpy(pose6).cuda(), want_w2c=True, want_cam7=True)
    c2w, w2c, cam7 = (o[k].cpu().numpy() for k in ("c2w", "w2c", "cam7"))
    for p, W, T, q in zip(pose6, c2w, w2c, cam7):
        assert np.abs(T - opo.transformMatrix(p[:3], p[3:])).max() < 1e-15
        assert np.abs(W - opo.camera_to_world(p[:3], p[3:])).max() < 1e-14
        assert np.abs(q - opo.sbacam(W)).max() < 1e-14
    g = poses.guess_batch(o["c2w"]).cpu().numpy()
    for W, gi in zip(c2w, g):
        r, t = opo.pose_guess(W)
        if np.linalg.norm(r) < np.pi - 1e-4:
            assert np.abs(gi[:3] - r.ravel()).max() < 1e-9
        assert np.array_equal(gi[3:], W[:3, 3])
    # drop-ins with the reference's names and shapes; main.py:199 hands transformMatrix a (3, 1, 1) tvec on OpenCV 4.13
    T = poses.transformMatrix(rv[20].reshape(3, 1), np.array([1.0, 2.0, 3.0]).reshape(3, 1)[:, np.newaxis])
    assert isinstance(T, np.matrix) and T.shape == (4, 4) and np.array_equal(np.asarray(T)[:3, 3], [1.0, 2.0, 3.0])
    r = poses.Rtorvec(np.asarray(T)[:3, :3])
    assert r.shape == (3, 1) and np.abs(r.ravel() - rv[20]).max() < 1e-12
    assert poses.rodrigues_batch(torch.zeros((0, 3), dtype=torch.float64, device="cuda")).shape == (0, 3, 3)
