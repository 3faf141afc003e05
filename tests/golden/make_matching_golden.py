"""Generates tests/golden/matching_icl_orb.npz by RUNNING THE REFERENCE's FeatureMatcher
(/root/reference/src/v2/frame.py:16-49, imported unmodified) with its matcher attribute set to
cv2.BFMatcher(cv2.NORM_HAMMING) -- the ORB variant the reference leaves commented at frame.py:8
-- on real ORB descriptors of ICL-NUIM frames 1 and 32.  Only runs in the build container
(needs /root/reference); the .npz travels.  Also stores a low-entropy tie-heavy case.
"""
import os
import sys

import cv2
import numpy as np

REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "src", "v2"))
import frame as ref_frame  # noqa: E402

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from visual_slam_b200 import synth  # noqa: E402


def run_ref(kp1, d1, kp2, d2, ratio=0.8):
    fm = ref_frame.FeatureMatcher()
    fm.matcher = cv2.BFMatcher(cv2.NORM_HAMMING)
    matches, p1, f1, p2, f2 = fm.match_features(kp1, d1, kp2, d2, ratio)
    raw = fm.matcher.knnMatch(d1, d2, k=2)
    idx = np.array([[m.trainIdx for m in r] for r in raw], dtype=np.int32)
    dist = np.array([[int(m.distance) for m in r] for r in raw], dtype=np.int32)
    kq = np.array([m[0].queryIdx for m in matches], dtype=np.int32)
    kt = np.array([m[0].trainIdx for m in matches], dtype=np.int32)
    cc = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).knnMatch(d1, d2, k=1)
    mutual = np.array(sorted((m.queryIdx, m.trainIdx) for r in cc for m in r), dtype=np.int32)
    return dict(idx=idx, dist=dist, keep_q=kq, keep_t=kt, mutual=mutual.reshape(-1, 2),
                pts1=p1, pts2=p2)


def main():
    orb = cv2.ORB_create(1000)
    out = {}
    imgs = [cv2.imread(os.path.join(REF, "data/ICL_NUIM/rgb/%d.png" % i)) for i in (1, 32)]
    (k1, d1), (k2, d2) = [orb.detectAndCompute(im, None) for im in imgs]
    kp1, kp2 = cv2.KeyPoint_convert(k1), cv2.KeyPoint_convert(k2)
    out.update({"icl_" + k: v for k, v in run_ref(kp1, d1, kp2, d2).items()})
    out.update(icl_kp1=kp1, icl_kp2=kp2, icl_desc1=d1, icl_desc2=d2)
    q, t = synth.low_entropy_pair(300, 275, seed=7)
    kpq = np.arange(600, dtype=np.float32).reshape(300, 2)
    kpt = np.arange(550, dtype=np.float32).reshape(275, 2)
    out.update({"tie_" + k: v for k, v in run_ref(kpq, q, kpt, t, ratio=0.95).items()})
    out.update(tie_kp1=kpq, tie_kp2=kpt, tie_desc1=q, tie_desc2=t)
    path = os.path.join(os.path.dirname(__file__), "matching_icl_orb.npz")
    np.savez_compressed(path, **out)
    print(path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
