"""C1 replay: the control flow of the reference driver, re-created once and shared by the golden generator and
the parity tests (SURVEY.md 7 step 0(v), 8d "C1").

The reference's driver, /root/reference/src/v2/main.py, is a script that cannot run (it imports `imp`, pangolin,
OpenGL, graphslam, g2o and reads /home/juuso/... paths), so its two loops are restated here, statement
by statement, over a *backend*: the objects the script gets from its imports.  The generator
(tests/golden/make_c1_replay_golden.py) passes the reference's OWN modules (frame.py, map.py, point.py,
helper_functions.py, LocalBA.py over oracle/g2o_shim.py, cv2.solvePnPRansac); the tests pass the drop-ins of
visual_slam_b200.  Every hot-path call goes through `hook(stage, i, fn, inputs)`, which lets the generator record
each stage's inputs / outputs and lets a test compare its own results and, if it wishes, continue from the
recorded ones (teacher forcing), so that one stage's permitted differences (RANSAC sampling) do not leak into
the next stage's comparison.

Deviations from main.py, all forced and stated: ICL-NUIM intrinsics instead of the TUM ones hard-coded at
main.py:57-58 (the bundled data is ICL); ORB descriptors + Hamming norm (frame.py:8, north_star);
`tvec[:, np.newaxis]` (main.py:199) only when tvec.ndim == 1 (OpenCV 4.13 already returns (3, 1), App. A);
viewer calls and prints left out; the loop ends at `last_index`.
"""
import numpy as np

ICL_K = np.array([[481.20, 0, 319.5], [0, 480.0, 239.5], [0, 0, 1]])


class Camera:                      # main.py:24-30
    def __init__(self, fx, fy, cx, cy, baseline=1):
        self.fx, self.fy, self.cx, self.cy, self.baseline = fx, fy, cx, cy, baseline


class Isometry3d:                  # main.py:32-51
    def __init__(self, R, t):
        self.R, self.t = R, t

    def matrix(self):
        m = np.identity(4)
        m[:3, :3] = self.R
        m[:3, 3] = self.t
        return m

    def inverse(self):
        return Isometry3d(self.R.T, -self.R.T @ self.t)


def MakeHomogeneous(x):            # helper_functions.py (main.py:279-280)
    return np.concatenate([x, np.ones((x.shape[0], 1))], axis=1)


def CameraProjectionMatrix2(Pose, K):   # helper_functions.py: K @ Pose[:3] (main.py:275-276)
    return np.asarray(K) @ np.asarray(Pose)[:3, :]


def snapshot(m):
    """Poses and points of a map after a BA call (what LocalBA.py:185-190 / :226-229 wrote back)."""
    fids = list(m.frames.keys())
    pids = list(m.points_3d.keys())
    return {"frame_ids": np.array(fids, np.int64),
            "poses": np.array([np.asarray(m.GetFrame(f).GetPose(), dtype=np.float64) for f in fids]).reshape(-1, 4, 4),
            "point_ids": np.array(pids, np.int64),
            "points": np.array([np.asarray(m.GetPoint(p).Get3dPoint(), dtype=np.float64) for p in pids]).reshape(-1, 3)}


def apply_snapshot(m, s):
    for f, T in zip(s["frame_ids"], s["poses"]):
        m.UpdatePose(new_pose=np.array(T), frame_id=int(f))
    for p, X in zip(s["point_ids"], s["points"]):
        m.UpdatePoint3D(new_point=np.array(X), point_id=int(p))


def map_arrays(m):
    """The BA inputs a map holds, as arrays (for the record): observations (point id, frame id, uv), key-frame
    flags, parent links with the stored relative translation norm (LocalBA.py:126)."""
    s = snapshot(m)
    obs = [(p, f, c[1][0], c[1][1]) for p, pt in m.points_3d.items() for f, c in pt.frames.items() if f in m.frames]
    s["obs"] = np.array(obs, np.float64).reshape(-1, 4)
    s["is_kf"] = np.array([bool(m.GetFrame(int(f)).IsKeyFrame()) for f in s["frame_ids"]])
    par = [(f, p, np.linalg.norm(np.asarray(m.GetFrame(f).GetTransitionWithParentID(p))[:3, 3]))
           for f in m.frames for p in m.GetFrame(f).GetParentIDs() if p is not None]
    s["parents"] = np.array(par, np.float64).reshape(-1, 3)
    return s


def run(be, first_index, last_index, hook, K=ICL_K, max_keyframes=None):
    """be: backend namespace with Map, Point, make_frame(i, id) -> Frame, feature_matcher, estimateEssential,
    estimateRelativePose, solvePnPRansac, BundleAdjustment, GetListDiff, triangulate, pose_guess (main.py:191-194),
    camera_to_world (main.py:199-203), ba_stats, deepcopy, feature_extractor.  Returns a summary dict."""
    K = np.matrix(K)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    feature_matcher = be.feature_matcher
    gmap = be.Map()
    id_frame, id_point = 0, 1
    # ---- first frame (main.py:73-86)
    cur_frame = be.make_frame(first_index, id_frame)
    cur_frame.AddPose(init_pose=np.eye(4))
    cur_frame.SetAsKeyFrame()
    cur_frame.AddParent(None, None)
    kp_prev, features_prev, _ = cur_frame.process_frame(feature_extractor=be.feature_extractor)
    gmap.AddFrame(frame_id=id_frame, frame=cur_frame)
    id_frame += 1
    # ---- map initialisation (main.py:88-138): always against the first frame, stops at the first success
    i = first_index
    initialised = False
    for i in range(first_index, last_index + 1):
        prev_frame = gmap.GetFrame(id_frame - 1)
        cur_frame = be.make_frame(i, id_frame)
        kp_cur, features_cur, _ = cur_frame.process_frame(feature_extractor=be.feature_extractor)
        matches, pre_pts, pre_fts, cur_pts, cur_fts = hook(
            "match_init", i, lambda: feature_matcher.match_features(kp_prev, features_prev, kp_cur, features_cur),
            dict(kp1=kp_prev, desc1=features_prev, kp2=kp_cur, desc2=features_cur))
        if len(matches) < 100:
            continue
        E, inliers, score = hook("essential", i, lambda: be.estimateEssential(pre_pts, cur_pts, K, essTh=3.0 / K[0, 0]),
                                 dict(pts1=pre_pts, pts2=cur_pts))
        sel = inliers[:, 0] == 1
        in_pre, in_pre_f, in_cur, in_cur_f = pre_pts[sel, :], pre_fts[sel, :], cur_pts[sel, :], cur_fts[sel, :]
        R, t, validFraction, tri, in_pre, in_cur, in_pre_f, in_cur_f = hook(
            "relpose", i, lambda: be.estimateRelativePose(E, in_pre, in_cur, in_pre_f, in_cur_f, K, "Essential"),
            dict(E=E, pts1=in_pre, pts2=in_cur, fts1=in_pre_f, fts2=in_cur_f))
        if validFraction < 0.9:
            continue
        rel = Isometry3d(R=np.asarray(R), t=np.squeeze(np.asarray(t))).inverse().matrix()
        pose = rel @ prev_frame.GetPose()
        gmap.AddParentAndPose(parent_id=id_frame - 1, frame_id=id_frame, frame_obj=cur_frame, rel_pose_trans=rel,
                              pose=pose)
        id_frame += 1
        pts_objs = (np.linalg.inv(prev_frame.GetPose()) @ tri).T
        pts_objs = pts_objs[:, :3] / np.asarray(pts_objs[:, -1]).reshape(-1, 1)
        for pt, uv1, uv2, ft1, ft2 in zip(pts_objs, in_pre, in_cur, in_pre_f, in_cur_f):
            pt_object = be.Point(location=pt, id=id_point)
            pt_object.AddFrame(frame=prev_frame, uv=uv1, descriptor=ft1)
            pt_object.AddFrame(frame=cur_frame, uv=uv2, descriptor=ft2)
            gmap.AddPoint3D(point_id=id_point, point_3d=pt_object)
            id_point += 1
        cur_frame.SetAsKeyFrame()
        initialised = True
        break
    summary = {"init_index": i, "initialised": initialised, "keyframe_indices": [first_index, i], "tracked": []}
    if not initialised:
        return summary
    # ---- BA of the initial map, rescaled to unit median depth (main.py:144-148)
    camera = Camera(fx, fy, cx, cy)

    def ba_init():
        BA = be.BundleAdjustment(camera)
        BA.localBundleAdjustement(gmap, scale=True)
        return dict(snapshot(gmap), stats=be.ba_stats(BA))
    apply_snapshot(gmap, hook("ba_init", i, ba_init, dict(map=gmap)))
    # ---- tracking (main.py:153-348)
    last_keyframe = be.deepcopy(gmap.GetFrame(frame_id=id_frame - 1))
    local_map = be.Map()
    local_map.AddFrame(last_keyframe.GetID(), last_keyframe)
    id_frame_local = id_frame
    local_map.Store3DPoints(gmap.GetCopyOfPointObjects(last_keyframe.GetID()))
    loop_idx = i
    n_kf = 0
    for i in range(loop_idx + 1, last_index + 1):
        cur_frame = be.make_frame(i, id_frame_local)
        kp_cur, features_cur, _ = cur_frame.process_frame(feature_extractor=be.feature_extractor)
        kp_prev, features_prev, known_3d, point_IDs = local_map.GetImagePointsWithFrameID(last_keyframe.GetID())
        matches, pre_pts, pre_fts, cur_pts, cur_fts = hook(
            "match_track", i, lambda: feature_matcher.match_features(kp_prev, features_prev, kp_cur, features_cur),
            dict(kp1=kp_prev, desc1=features_prev, kp2=kp_cur, desc2=features_cur))
        known_3d_matched = np.array([known_3d[m[0].queryIdx] for m in matches])
        known_3d_matched_ids = [point_IDs[m[0].queryIdx] for m in matches]
        W_T_prev = local_map.GetFrame(id_frame_local - 1).GetPose()
        prev_T_W = Isometry3d(R=W_T_prev[0:3, 0:3], t=np.asarray(W_T_prev[:3, -1]).squeeze()).inverse().matrix()
        rvec_guess, tvec_guess = hook("guess", i, lambda: be.pose_guess(W_T_prev), dict(W_T_prev=W_T_prev))
        obj = known_3d_matched[:, np.newaxis, :].astype(np.float32)
        img = cur_pts[:, np.newaxis, :].astype(np.float32)
        retval, rvec, tvec, inliers = hook(
            "pnp", i, lambda: be.solvePnPRansac(objectPoints=obj, imagePoints=img, cameraMatrix=K,
                                                distCoeffs=np.array([]), rvec=rvec_guess.copy(),
                                                tvec=tvec_guess.copy(), useExtrinsicGuess=True),
            dict(objectPoints=obj, imagePoints=img, rvec=rvec_guess, tvec=tvec_guess))
        tvec = np.asarray(tvec)
        if tvec.ndim == 1:
            tvec = tvec[:, np.newaxis]                                  # main.py:199, see the module docstring
        W_T_curr = hook("pose_plumbing", i, lambda: be.camera_to_world(rvec, tvec), dict(rvec=rvec, tvec=tvec))
        rel = prev_T_W @ W_T_curr
        local_map.AddParentAndPose(parent_id=id_frame_local - 1, frame_id=id_frame_local, frame_obj=cur_frame,
                                   rel_pose_trans=rel, pose=W_T_curr)
        local_map.AddPointToFrameCorrespondences(point_ids=[point_IDs[m[0].queryIdx] for m in matches],
                                                 image_points=cur_pts, descriptors=cur_fts, frame_obj=cur_frame)

        def ba_motion():
            localBA = be.BundleAdjustment(camera)
            localBA.motionOnlyBundleAdjustement(local_map, scale=False, save=True)
            return dict(snapshot(local_map), stats=be.ba_stats(localBA))
        apply_snapshot(local_map, hook("ba_motion", i, ba_motion, dict(map=local_map)))
        summary["tracked"].append((i, len(matches), 0 if inliers is None else len(inliers)))
        if (i - loop_idx > 20 or len(cur_pts) < 80) and (len(cur_pts) < 0.9 * len(known_3d)):
            # ---- key-frame insertion (main.py:222-345)
            loop_idx = i
            cur_frame.SetAsKeyFrame()
            W_T_prev_key = gmap.GetFrame(id_frame - 1).GetPose()
            prev_key_T_W = Isometry3d(R=W_T_prev_key[0:3, 0:3],
                                      t=np.asarray(W_T_prev_key[:3, -1]).squeeze()).inverse().matrix()
            W_T_cur_key = local_map.GetFrame(id_frame_local).GetPose()
            cur_frame.ClearParent()
            gmap.AddParentAndPose(parent_id=id_frame - 1, frame_id=id_frame, frame_obj=cur_frame,
                                  rel_pose_trans=prev_key_T_W @ W_T_cur_key, pose=W_T_cur_key)
            gmap.AddPointToFrameCorrespondences(point_ids=known_3d_matched_ids, image_points=cur_pts,
                                                descriptors=cur_fts, frame_obj=cur_frame)
            if id_frame >= 6 and id_frame % 4 == 0:
                gmap.DiscardOutlierMapPoints(n_visible_frames=3)
            in_map = gmap.GetImagePointsWithFrameID(id_frame - 1)[0]
            kp1 = gmap.GetFrame(id_frame - 1).GetKeyPoints()
            desc1 = gmap.GetFrame(id_frame - 1).GetFeatures()
            idx = hook("listdiff", i, lambda: be.GetListDiff(kp1, in_map), dict(kp1=kp1, kp2=in_map))
            kp1, desc1 = kp1[idx], desc1[idx]
            kp2, desc2 = gmap.GetFrame(id_frame).GetKeyPoints(), gmap.GetFrame(id_frame).GetFeatures()
            matches, last_pts, last_fts, new_pts, new_fts = hook(
                "match_kf", i, lambda: feature_matcher.match_features(kp1=kp1, desc1=desc1, kp2=kp2, desc2=desc2),
                dict(kp1=kp1, desc1=desc1, kp2=kp2, desc2=desc2))
            T1, T2 = gmap.GetFrame(id_frame - 1).GetPose(), gmap.GetFrame(id_frame).GetPose()
            p1 = Isometry3d(R=T1[0:3, 0:3], t=np.asarray(T1[:3, -1]).squeeze()).inverse().matrix()
            p2 = Isometry3d(R=T2[0:3, 0:3], t=np.asarray(T2[:3, -1]).squeeze()).inverse().matrix()
            Proj1, Proj2 = CameraProjectionMatrix2(Pose=p1, K=K), CameraProjectionMatrix2(Pose=p2, K=K)
            if len(matches):
                x1, x2 = MakeHomogeneous(last_pts), MakeHomogeneous(new_pts)

                def tri_and_filter():
                    X = be.triangulate(pose1=Proj1, pose2=Proj2, pts1=x1, pts2=x2)
                    X = X / X[:, 3:]
                    proj1, proj2 = p1 @ X.T, p2 @ X.T
                    good = np.where((proj1[2] > 0) & (proj2[2] > 0) & (proj2[2] < 1) & (proj1[2] < 1))[0]
                    return X[:, :3], good
                new_X, good = hook("triangulate", i, tri_and_filter,
                                   dict(Proj1=Proj1, Proj2=Proj2, p1=p1, p2=p2, x1=x1, x2=x2))
                for pt, uv1, uv2, ft1, ft2 in zip(new_X[good], last_pts[good], new_pts[good], last_fts[good],
                                                  new_fts[good]):
                    pt_object = be.Point(location=pt, id=id_point)
                    pt_object.AddFrame(frame=gmap.GetFrame(id_frame - 1), uv=uv1, descriptor=ft1)
                    pt_object.AddFrame(frame=gmap.GetFrame(id_frame), uv=uv2, descriptor=ft2)
                    gmap.AddPoint3D(point_id=id_point, point_3d=pt_object)
                    id_point += 1

            def ba_full():
                BA = be.BundleAdjustment(camera)
                BA.localBundleAdjustement(gmap)
                return dict(snapshot(gmap), stats=be.ba_stats(BA))
            apply_snapshot(gmap, hook("ba_full", i, ba_full, dict(map=gmap)))
            last_keyframe = be.deepcopy(gmap.GetFrame(frame_id=id_frame))
            local_map = be.Map()
            local_map.AddFrame(last_keyframe.GetID(), last_keyframe)
            id_frame += 1
            id_frame_local = id_frame
            local_map.Store3DPoints(gmap.GetCopyOfPointObjects(last_keyframe.GetID()))
            summary["keyframe_indices"].append(i)
            n_kf += 1
            if max_keyframes is not None and n_kf >= max_keyframes:
                break
        else:
            id_frame_local += 1
    summary["last_index"] = i
    summary["n_points"] = len(gmap.points_3d)
    summary["n_frames"] = len(gmap.frames)
    return summary
