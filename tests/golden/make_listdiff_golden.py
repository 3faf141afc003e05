"""Generates tests/golden/listdiff_ref.npz by running the REFERENCE's own `GetListDiff`
(/root/reference/src/v2/helper_functions.py:316-326) on seeded key-point sets.  Run in the build
container only (the GPU box has no /root/reference); the .npz is committed."""
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference/src/v2")
import helper_functions as ref  # noqa: E402

rng = np.random.default_rng(11)
out = {}
cases = [(40, 12, "grid"), (300, 120, "grid"), (257, 64, "subpixel"), (50, 0, "grid"), (64, 64, "same")]
for c, (n1, n2, kind) in enumerate(cases):
    if kind == "grid":      # integer corner coordinates (GFTT without sub-pixel refinement): coordinates repeat
        kp1 = np.stack([rng.integers(0, 640, n1), rng.integers(0, 480, n1)], axis=1).astype(np.float32)
    else:
        kp1 = np.stack([rng.uniform(0, 640, n1), rng.uniform(0, 480, n1)], axis=1).astype(np.float32)
    if kind == "same":
        kp2 = kp1.copy()
    elif n2:
        kp2 = kp1[rng.choice(n1, n2, replace=False)].copy()       # the matched (mapped) subset, shuffled
    else:
        kp2 = np.zeros((1, 2), np.float32) - 1.0                   # nothing mapped yet (one off-image row)
    out["kp1_%d" % c], out["kp2_%d" % c] = kp1, kp2
    out["idx_%d" % c] = np.asarray(ref.GetListDiff(kp1, kp2), dtype=np.int64)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "listdiff_ref.npz"), **out)
print({k: v.shape for k, v in out.items() if k.startswith("idx")})
