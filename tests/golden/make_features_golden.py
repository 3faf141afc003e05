"""Generates tests/golden/features_icl.npz by RUNNING THE REFERENCE's own FeatureExtractor
(/root/reference/src/v2/frame.py:5-14) on ICL-NUIM frames, with its extractor set to the ORB variant the
file itself names (frame.py:8, `self.extractor = cv2.ORB_create()`).  Also stores what cv2 returns for the
intermediate goodFeaturesToTrack call of frame.py:11.  Run in the build container only (the GPU box has no
/root/reference); the .npz (images included, so the test needs nothing else) is committed.

Run from the repo root:  python tests/golden/make_features_golden.py
"""
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, "/root/reference/src/v2")
import frame as ref_frame  # noqa: E402

out = {"cv2_version": np.array(cv2.__version__)}
fx = ref_frame.FeatureExtractor()
fx.extractor = cv2.ORB_create()          # frame.py:8
for name, path, crop in [("icl1", "/root/reference/data/ICL_NUIM/rgb/1.png", None),
                         ("icl32_crop", "/root/reference/data/ICL_NUIM/rgb/32.png", (40, 401, 100, 547))]:
    img = cv2.imread(path)
    if crop:
        img = np.ascontiguousarray(img[crop[0]:crop[1], crop[2]:crop[3]])   # odd sizes: 361 x 447
    kp, des = fx.compute_features(img)
    corners = cv2.goodFeaturesToTrack(np.mean(img, axis=2).astype(np.uint8), 3000, qualityLevel=0.01, minDistance=7)
    out[name + "_img"] = img
    out[name + "_kp"] = kp
    out[name + "_des"] = des
    out[name + "_corners"] = corners.reshape(-1, 2)
    print(name, img.shape, "corners", corners.shape[0], "kp", kp.shape, "des", des.shape)
# C1 front end on the frame pair 1 -> 32: the reference's FeatureExtractor (ORB) and its FeatureMatcher with the
# Hamming norm the ORB variant needs (frame.py:18 -> cv2.BFMatcher(cv2.NORM_HAMMING)), both classes unmodified
img32 = cv2.imread("/root/reference/data/ICL_NUIM/rgb/32.png")
kp1, des1 = out["icl1_kp"], out["icl1_des"]
kp32, des32 = fx.compute_features(img32)
fm = ref_frame.FeatureMatcher()
fm.matcher = cv2.BFMatcher(cv2.NORM_HAMMING)
matches, p1, f1, p2, f2 = fm.match_features(kp1, des1, kp32, des32)
out["icl32_img"], out["icl32_kp"], out["icl32_des"] = img32, kp32, des32
out["pair_query"] = np.array([m[0].queryIdx for m in matches], np.int32)
out["pair_train"] = np.array([m[0].trainIdx for m in matches], np.int32)
out["pair_dist"] = np.array([m[0].distance for m in matches], np.float32)
out["pair_pts1"], out["pair_pts2"], out["pair_ft1"], out["pair_ft2"] = p1, p2, f1, f2
print("pair 1->32: kp", kp1.shape[0], kp32.shape[0], "ratio matches", len(matches))
# ... and the reference's own map-initialisation calls on those matches (main.py:103,111 with the ICL intrinsics)
import helper_functions as ref_hf  # noqa: E402
K = np.matrix([[481.20, 0, 319.5], [0, 480.0, 239.5], [0, 0, 1]])
E, inl, score = ref_hf.estimateEssential(p1, p2, K, essTh=3.0 / K[0, 0])
sel = inl.ravel() == 1
R, t, frac, tri, ip1, ip2, if1, if2 = ref_hf.estimateRelativePose(E, p1[sel], p2[sel], f1[sel], f2[sel], K, "Essential")
out["pair_E"], out["pair_inliers"], out["pair_R"], out["pair_t"] = E, inl, R, t
out["pair_valid_fraction"] = np.array(frac)
print("essential inliers", int(sel.sum()), "of", len(sel), "valid fraction", frac)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "features_icl.npz")
np.savez_compressed(path, **out)
print("written", path, os.path.getsize(path))
