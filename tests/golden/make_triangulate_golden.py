"""Generates tests/golden/triangulate_ref.npz by running the REFERENCE's own `triangulate`
(/root/reference/src/v2/helper_functions.py:281-291) on seeded two-view scenes.  Run in the build
container only (the GPU box has no /root/reference); the .npz is committed."""
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference/src/v2")
import helper_functions as ref  # noqa: E402

rng = np.random.default_rng(7)
K = np.array([[481.2, 0, 319.5], [0, 480.0, 239.5], [0, 0, 1.0]])
out = {}
for case, (n, noise, scale) in enumerate([(64, 0.0, 1.0), (257, 0.5, 1.0), (33, 1.0, 0.2)]):
    # world points in front of both cameras; `scale` shrinks the scene so that some depths fall in (0, 1)
    X = np.stack([rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(2, 8, n)], axis=1) * scale
    ang = 0.05
    R2 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    t2 = np.array([-0.3, 0.02, 0.05]) * scale
    T1 = np.hstack([np.eye(3), np.zeros((3, 1))])
    T2 = np.hstack([R2, t2[:, None]])
    P1, P2 = K @ T1, K @ T2
    Xh = np.hstack([X, np.ones((n, 1))])
    x1 = (P1 @ Xh.T).T; x1 = x1[:, :2] / x1[:, 2:]
    x2 = (P2 @ Xh.T).T; x2 = x2[:, :2] / x2[:, 2:]
    x1 += rng.normal(0, noise, x1.shape); x2 += rng.normal(0, noise, x2.shape)
    x1h = np.hstack([x1, np.ones((n, 1))]); x2h = np.hstack([x2, np.ones((n, 1))])   # MakeHomogeneous (main.py:271-272)
    tri = ref.triangulate(pose1=P1, pose2=P2, pts1=x1h, pts2=x2h)
    out["P1_%d" % case], out["P2_%d" % case] = P1, P2
    out["T1_%d" % case], out["T2_%d" % case] = T1, T2
    out["x1_%d" % case], out["x2_%d" % case] = x1, x2
    out["tri_%d" % case] = tri
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "triangulate_ref.npz"), **out)
print("written", {k: v.shape for k, v in out.items() if k.startswith("tri")})
