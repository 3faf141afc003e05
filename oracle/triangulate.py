"""CPU oracle for two-view triangulation.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Restates reference src/v2/helper_functions.py:281-291 (`triangulate`) and the de-homogenisation +
cheirality/depth filter of src/v2/main.py:284-306.  Pinned by tests/golden/triangulate_ref.npz, which
holds outputs of the reference's own function imported from /root/reference (generating script:
tests/golden/make_triangulate_golden.py).
"""
import numpy as np


def triangulate(pose1, pose2, pts1, pts2):
    """helper_functions.py:281-291: rows u*P[2]-P[0], v*P[2]-P[1] of both views, SVD, vt[3].
    pose1/pose2: 3x4 projection matrices; pts: (N, >=2).  Returns (N, 4), LAPACK's sign."""
    pose1 = np.asarray(pose1, dtype=np.float64)
    pose2 = np.asarray(pose2, dtype=np.float64)
    ret = np.zeros((len(pts1), 4))
    for i, (p, q) in enumerate(zip(np.asarray(pts1, dtype=np.float64), np.asarray(pts2, dtype=np.float64))):
        A = np.zeros((4, 4))
        A[0] = p[0] * pose1[2] - pose1[0]
        A[1] = p[1] * pose1[2] - pose1[1]
        A[2] = q[0] * pose2[2] - pose2[0]
        A[3] = q[1] * pose2[2] - pose2[1]
        _, _, vt = np.linalg.svd(A)
        ret[i] = vt[3]
    return ret


def dehomogenise_and_filter(Xh, T1, T2, z_max=1.0):
    """main.py:285 (divide by w), :293-294 (depth in both cameras through the world->camera poses),
    :306 (keep 0 < z < z_max in both).  Returns (X3 (N,3), depth (N,2), good (N,) bool)."""
    X = Xh / Xh[:, 3:]
    z1 = (np.asarray(T1, dtype=np.float64)[:3] @ X.T)[2]
    z2 = (np.asarray(T2, dtype=np.float64)[:3] @ X.T)[2]
    good = (z1 > 0) & (z2 > 0) & (z2 < z_max) & (z1 < z_max)
    return X[:, :3], np.stack([z1, z2], axis=1), good
