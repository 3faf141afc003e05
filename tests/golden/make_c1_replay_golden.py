"""Generates tests/golden/c1_replay.npz: the reference's tracking loop over ICL-NUIM frames, stage by stage.

Runs in the build container only (needs /root/reference).  The loop of /root/reference/src/v2/main.py:73-348
(re-created in tests/c1_loop.py, see there for the forced deviations) is driven with the reference's OWN modules,
imported from /root/reference/src/v2 and unmodified:

  frame.FeatureExtractor (extractor = cv2.ORB_create(), the variant frame.py:8 names), frame.FeatureMatcher
  (matcher = cv2.BFMatcher(cv2.NORM_HAMMING), which ORB needs), frame.Frame, map.Map, point.Point,
  helper_functions.estimateEssential / estimateRelativePose / Rtorvec / transformMatrix / GetListDiff /
  triangulate, cv2.solvePnPRansac exactly as called at main.py:196-197, and LocalBA.BundleAdjustment.

LocalBA.py needs the g2o wheel, which cannot be had here; it runs over oracle/g2o_shim.py, i.e. the graph is
built by the reference's code and optimised by the CPU restatement oracle/ba.py.  BA numbers in this file are
therefore "restatement, not g2o" (PARITY UNPINNED for the BA arithmetic); everything else is the reference's
and OpenCV's own output.

Run from the repo root:  python tests/golden/make_c1_replay_golden.py [last_index]
"""
import contextlib
import copy
import io
import os
import sys
import zlib

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference/src/v2")

from oracle import g2o_shim  # noqa: E402

g2o_shim.install()
import frame as ref_frame  # noqa: E402
import helper_functions as ref_hf  # noqa: E402
import LocalBA as ref_ba  # noqa: E402
import map as ref_map  # noqa: E402
import point as ref_point  # noqa: E402

import c1_loop  # noqa: E402

RGB = "/root/reference/data/ICL_NUIM/rgb/%d.png"
DEPTH = "/root/reference/data/ICL_NUIM/depth/%d.png"
FIRST = 1


def crc(*arrays):
    c = 0
    for a in arrays:
        a = np.ascontiguousarray(a)
        c = zlib.crc32(a.tobytes(), zlib.crc32(str((a.shape, a.dtype.str)).encode(), c))
    return np.uint32(c)


class Backend:
    Map, Point = ref_map.Map, ref_point.Point
    deepcopy = staticmethod(copy.deepcopy)
    estimateEssential = staticmethod(ref_hf.estimateEssential)
    estimateRelativePose = staticmethod(ref_hf.estimateRelativePose)
    solvePnPRansac = staticmethod(cv2.solvePnPRansac)
    GetListDiff = staticmethod(ref_hf.GetListDiff)
    triangulate = staticmethod(ref_hf.triangulate)
    BundleAdjustment = ref_ba.BundleAdjustment

    def __init__(self):
        self.feature_extractor = ref_frame.FeatureExtractor()
        self.feature_extractor.extractor = cv2.ORB_create()            # frame.py:8
        self.feature_matcher = ref_frame.FeatureMatcher()
        self.feature_matcher.matcher = cv2.BFMatcher(cv2.NORM_HAMMING)
        self.frames = {}

    def make_frame(self, i, id):
        fr = ref_frame.Frame(RGB % i, DEPTH % i, id=id)
        self.frames[i] = fr
        return fr

    @staticmethod
    def pose_guess(W_T_prev):
        """main.py:193-194."""
        return ref_hf.Rtorvec(W_T_prev[0:3, 0:3]), W_T_prev[0:3, 3]

    @staticmethod
    def camera_to_world(rvec, tvec):
        """main.py:201-203."""
        T = ref_hf.transformMatrix(rvec, tvec)
        r, t = T[:3, :3], np.asarray(T[:3, -1]).squeeze()
        return c1_loop.Isometry3d(R=r, t=t).inverse().matrix()

    @staticmethod
    def ba_stats(BA):
        h = BA.history
        return {"initial_cost": h["cost"][0], "final_cost": h["final_cost"], "iterations": h["iterations"],
                "trials": h["trials"], "costs": np.array(h["cost"]), "lambdas": np.array(h["lambda"]),
                "n_cams": BA.problem.n_cams, "n_pts": BA.problem.n_pts, "n_obs": BA.problem.n_obs,
                "n_scale": BA.problem.n_scale, "n_free_cams": BA.problem.n_free_cams,
                "n_free_pts": BA.problem.n_free_pts,
                "obs_crc": crc(*[a[np.lexsort((BA.problem.obs_cam, BA.problem.obs_pt))]
                                 for a in (BA.problem.obs_pt, BA.problem.obs_cam, BA.problem.obs_uv)])}


def main():
    last = int(sys.argv[1]) if len(sys.argv) > 1 else 80
    be = Backend()
    out = {"cv2_version": np.array(cv2.__version__), "first_index": np.array(FIRST), "last_index": np.array(last),
           "K": c1_loop.ICL_K}
    log = []

    def hook(stage, i, fn, inputs):
        with contextlib.redirect_stdout(io.StringIO()):      # LocalBA.get_point prints every point (:138)
            r = fn()
        key = "%s_%d_" % (stage, i)
        log.append((stage, i))
        if stage.startswith("match"):
            m = r[0]
            out[key + "q"] = np.array([x[0].queryIdx for x in m], np.int32)
            out[key + "t"] = np.array([x[0].trainIdx for x in m], np.int32)
            out[key + "d"] = np.array([x[0].distance for x in m], np.float32)
            out[key + "in_crc"] = crc(inputs["kp1"], inputs["desc1"], inputs["kp2"], inputs["desc2"])
            if stage != "match_init":
                out[key + "kp1"], out[key + "desc1"] = inputs["kp1"], inputs["desc1"]
        elif stage == "essential":
            out[key + "E"], out[key + "inliers"], out[key + "score"] = r[0], r[1], np.array(r[2])
        elif stage == "relpose":
            out[key + "R"], out[key + "t"], out[key + "vf"], out[key + "tri"] = r[0], r[1], np.array(r[2]), r[3]
            p1 = np.asarray(inputs["pts1"])
            keep = [int(np.nonzero((p1 == row).all(axis=1))[0][0]) for row in r[4]]   # GFTT corners are unique
            assert all(np.array_equal(p1[k], row) for k, row in zip(keep, r[4])) and keep == sorted(keep)
            out[key + "keep"] = np.array(keep, np.int32)
        elif stage == "pnp":
            out[key + "obj"], out[key + "img"] = inputs["objectPoints"], inputs["imagePoints"]
            out[key + "rvec_guess"], out[key + "tvec_guess"] = inputs["rvec"], np.asarray(inputs["tvec"])
            out[key + "retval"] = np.array(bool(r[0]))
            out[key + "rvec"], out[key + "tvec"] = r[1], r[2]
            out[key + "inliers"] = (np.zeros((0, 1), np.int32) if r[3] is None else r[3])
        elif stage == "pose_plumbing":
            out[key + "W_T_curr"] = np.asarray(r, dtype=np.float64)
        elif stage == "guess":
            out[key + "W_T_prev"] = np.asarray(inputs["W_T_prev"], dtype=np.float64)
        elif stage == "listdiff":
            out[key + "idx"] = np.array(r, np.int32)
            out[key + "kp1"], out[key + "kp2"] = inputs["kp1"], inputs["kp2"]
        elif stage == "triangulate":
            out[key + "X"], out[key + "good"] = r[0], np.asarray(r[1], np.int32)
            for k in ("Proj1", "Proj2", "p1", "p2", "x1", "x2"):
                out[key + k] = np.asarray(inputs[k], dtype=np.float64)
        elif stage.startswith("ba_"):
            for k in ("frame_ids", "poses", "point_ids", "points"):
                out[key + k] = r[k]
            for k, v in r["stats"].items():
                out[key + "stat_" + k] = np.asarray(v)
        return r

    summary = c1_loop.run(be, FIRST, last, hook)
    print("summary:", {k: v for k, v in summary.items() if k != "tracked"})
    print("tracked (index, ratio matches, PnP inliers):", summary["tracked"])
    for i, fr in sorted(be.frames.items()):
        out["kp_%d" % i], out["desc_%d" % i] = fr.GetKeyPoints(), fr.GetFeatures()
    out["stages"] = np.array(["%s %d" % s for s in log])
    out["keyframe_indices"] = np.array(summary["keyframe_indices"], np.int32)
    out["init_index"] = np.array(summary["init_index"])
    path = os.path.join(HERE, "c1_replay.npz")
    np.savez_compressed(path, **out)
    print("written", path, os.path.getsize(path), "bytes,", len(log), "stage records")


if __name__ == "__main__":
    main()
