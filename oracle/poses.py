"""Oracle for the pose plumbing of the tracking loop.  TEST INFRASTRUCTURE -- never imported by the product.

Restates reference src/v2/helper_functions.py:269-278 (transformMatrix, Rtorvec: both are thin calls of
cv2.Rodrigues) and src/v2/main.py:191-194, 199-203 (guess from the previous camera->world pose; inversion of the
PnP result to camera->world).  cv2 is the reference's own third-party call (4.13.0 here, un-pinned there).
Pinned by tests/golden/c1_replay.npz, which holds what the reference's own functions returned for every tracked
frame of the ICL-NUIM replay (tests/golden/make_c1_replay_golden.py).
"""
import cv2
import numpy as np


def transformMatrix(rvec, tvec):
    """helper_functions.py:269-274."""
    R, _ = cv2.Rodrigues(np.asarray(rvec, dtype=np.float64).reshape(3, 1))
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = np.asarray(tvec, dtype=np.float64).reshape(3)
    return T


def Rtorvec(R):
    """helper_functions.py:276-278."""
    r, _ = cv2.Rodrigues(np.asarray(R, dtype=np.float64).reshape(3, 3))
    return r


def camera_to_world(rvec, tvec):
    """main.py:201-203: inverse of [R | t]."""
    T = transformMatrix(rvec, tvec)
    out = np.eye(4)
    out[:3, :3] = T[:3, :3].T
    out[:3, 3] = -T[:3, :3].T @ T[:3, 3]
    return out


def pose_guess(W_T_prev):
    """main.py:193-194."""
    W = np.asarray(W_T_prev, dtype=np.float64)
    return Rtorvec(W[:3, :3]), W[:3, 3].copy()


def sbacam(T):
    """(qx, qy, qz, qw, tx, ty, tz) of a camera->world 4x4: Eigen's matrix -> quaternion, then the SBACam
    constructor's normalisation with w >= 0 (LocalBA.py:57-58)."""
    from .g2o_shim import Isometry3d, SBACam
    iso = Isometry3d(np.asarray(T, dtype=np.float64))
    return SBACam(iso.orientation(), iso.position()).as7()
