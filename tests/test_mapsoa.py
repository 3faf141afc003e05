"""MapSoA: packing equals the drop-in's graph; BA through it equals the drop-in's BA."""
import numpy as np
import pytest

from visual_slam_b200.ba_problem import synthetic_ba
from map_helpers import _Cam, _map_from_problem


def _canon(prob):
    order = np.lexsort((prob.obs_pt, prob.obs_cam))
    return prob.obs_cam[order], prob.obs_pt[order], prob.obs_uv[order]


def test_packing_matches_the_drop_in_graph_builder():
    from visual_slam_b200.ba import BundleAdjustment
    from visual_slam_b200.mapsoa import MapSoA
    p = synthetic_ba(6, 200, 4.0, seed=5)
    m = _map_from_problem(p)
    ba = BundleAdjustment.__new__(BundleAdjustment)          # graph construction only: no device needed
    ba.fx, ba.fy, ba.cx, ba.cy = _Cam.fx, _Cam.fy, _Cam.cx, _Cam.cy
    ba._poses, ba._points, ba._scale, ba._result = {}, {}, [], None
    ba._e_pt, ba._e_pose, ba._e_uv = [], [], []
    frame_ids = list(m.frames.keys())
    for fid in frame_ids:
        fr = m.GetFrame(fid)
        ba.add_pose(fid, fr.GetPose(), fixed=(fid == 0))
        if fid != 0:
            for par in fr.GetParentIDs():
                ba.AddScalingEdge(par, fid, fr.GetTransitionWithParentID(par))
    ba._add_points_and_edges(m, frame_ids, list(m.points_3d.keys()), fixed=False)
    ref, _, _ = ba._problem()
    soa = MapSoA.from_map(m)
    got = soa.problem((_Cam.fx, _Cam.fy, _Cam.cx, _Cam.cy))
    assert soa.n_frames == p.n_cams and soa.n_points == p.n_pts and soa.n_observations == p.n_obs
    assert np.allclose(got.cams, ref.cams, atol=1e-15) and np.array_equal(got.pts, ref.pts)
    for a, b in zip(_canon(got), _canon(ref)):
        assert np.array_equal(a, b)
    assert np.array_equal(got.cam_fixed, ref.cam_fixed) and np.array_equal(got.pt_fixed, ref.pt_fixed)
    assert sorted(zip(got.scale_a, got.scale_b, got.scale_meas)) == sorted(zip(ref.scale_a, ref.scale_b, ref.scale_meas))
    mo = soa.problem((_Cam.fx, _Cam.fy, _Cam.cx, _Cam.cy), motion_only=True)
    assert mo.pt_fixed.all() and mo.n_scale == 0


def test_incremental_api_and_error_conventions():
    from visual_slam_b200.mapsoa import MapSoA
    s = MapSoA()
    s.AddFrame(0, np.eye(4), keyframe=True)
    with pytest.raises(Exception, match="Duplicate frame warning"):
        s.AddFrame(0, np.eye(4))
    with pytest.raises(Exception, match="No point yet added"):
        s.AddObservation(3, 0, (1.0, 2.0))
    s.AddPoint3D(3, [0, 0, 2.0])
    with pytest.raises(Exception, match="No frame yet added"):
        s.AddObservation(3, 9, (1.0, 2.0))
    for k in range(100):                                   # growth past the initial capacity
        s.AddPoint3D(10 + k, [k, 0, 3.0])
        s.AddObservation(10 + k, 0, (k, k))
    assert s.n_points == 101 and s.n_observations == 100
    T = np.eye(4); T[:3, 3] = [1, 2, 3]
    s.UpdatePose(T, 0)
    assert np.allclose(s.GetPose(0), T)
    s.UpdatePoint3D([7, 8, 9], 3)
    assert np.array_equal(s.GetAll3DPoints()[0], [7, 8, 9])


def test_repeated_observation_replaces_the_first_like_point_frames_dict():
    """Point.frames is a dict keyed by frame id (reference src/v2/point.py:14-46): the last observation of a
    (point, frame) wins and BA sees ONE edge."""
    from visual_slam_b200.mapsoa import MapSoA
    s = MapSoA()
    s.AddFrame(0, np.eye(4), keyframe=True); s.AddFrame(1, np.eye(4))
    s.AddPoint3D(5, [0, 0, 2.0]); s.AddPoint3D(6, [1, 0, 2.0])
    s.AddObservation(5, 0, (1.0, 2.0)); s.AddObservation(6, 0, (9.0, 9.0)); s.AddObservation(5, 1, (3.0, 4.0))
    v = s._version
    s.AddObservation(5, 0, (7.0, 8.0))                      # same (point, frame): replaces, no structural change
    assert s.n_observations == 3 and s._version == v
    prob = s.problem((1, 1, 0, 0))
    assert prob.n_obs == 3
    row = np.nonzero((prob.obs_pt == 0) & (prob.obs_cam == 0))[0]
    assert len(row) == 1 and np.array_equal(prob.obs_uv[row[0]], [7.0, 8.0])
    uv, _, _, ids = s.GetImagePointsWithFrameID(0)
    assert list(ids) == [5, 6] and np.array_equal(uv[0], [7.0, 8.0])
    assert np.array_equal(s.GetNVisibleFrames(), [2, 1])
    # the (point, frame) index survives culling: point 6 (one frame) goes, point 5 keeps its rows
    assert s.DiscardOutlierMapPoints(2) == 1
    s.AddObservation(5, 1, (30.0, 40.0))
    assert s.n_observations == 2 and np.array_equal(s.problem((1, 1, 0, 0)).obs_uv[1], [30.0, 40.0])


def test_R_to_quat_normalises_like_the_sbacam_constructor():
    from visual_slam_b200.ba_problem import R_to_quat, quat_to_R
    rng = np.random.default_rng(3)
    for _ in range(20):
        a = rng.normal(size=3); th = np.linalg.norm(a); k = a / th
        Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        R = np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx
        q = R_to_quat(R * (1 + 1e-3) + rng.normal(0, 1e-4, (3, 3)))   # scaled, slightly non-orthonormal
        assert abs(np.linalg.norm(q) - 1) < 1e-15 and q[3] >= 0
        assert np.abs(quat_to_R(q) - R).max() < 2e-3
        assert np.abs(quat_to_R(R_to_quat(R)) - R).max() < 1e-14


@pytest.mark.gpu
def test_ba_through_mapsoa_equals_the_drop_in():
    from visual_slam_b200.ba import BundleAdjustment
    from visual_slam_b200.mapsoa import MapSoA
    p = synthetic_ba(5, 300, 4.0, seed=13)
    m = _map_from_problem(p)
    soa = MapSoA.from_map(m)
    ba = BundleAdjustment(_Cam())
    ba.localBundleAdjustement(m)                            # mutates m
    sink = _map_from_problem(p)
    st = soa.localBundleAdjustement(_Cam(), sink=sink)
    assert st["final_cost"] == ba.stats["final_cost"]       # same arrays, same kernels: bit-identical
    for c in range(p.n_cams):
        assert np.allclose(sink.frames[c].pose, m.frames[c].pose, atol=1e-12)
        assert np.allclose(soa.GetPose(c), m.frames[c].pose, atol=1e-12)
    assert np.allclose(soa.GetAll3DPoints(), np.array([m.points_3d[j].location_3d for j in range(p.n_pts)]), atol=1e-12)
    # ... and both equal the CPU restatement run on the arrays MapSoA.problem() hands out (not on the drop-in's)
    from oracle import ba as oba
    ref = MapSoA.from_map(_map_from_problem(p)).problem((_Cam.fx, _Cam.fy, _Cam.cx, _Cam.cy))
    h = oba.optimize(ref, 10)
    assert abs(st["final_cost"] - h["final_cost"]) <= 1e-9 * h["final_cost"]
    assert (int(st["iterations"]), int(st["trials"])) == (h["iterations"], h["trials"])
    assert np.abs(soa.problem((_Cam.fx, _Cam.fy, _Cam.cx, _Cam.cy)).cams - ref.cams).max() < 1e-8
    assert np.abs(soa.GetAll3DPoints() - ref.pts).max() < 1e-7
    # unchanged structure: the solver object is reused (state upload only) and converges further
    eng = soa._engine[1]
    st2 = soa.localBundleAdjustement(_Cam())
    assert soa._engine[1] is eng and st2["initial_cost"] <= st["final_cost"] * (1 + 1e-9)
    # a structural change keeps the engine object (device buffers re-used) and gives it the new structure
    soa.AddPoint3D(10 ** 6, [0.0, 0.0, 3.0])
    soa.AddObservation(10 ** 6, 1, (300.0, 200.0)); soa.AddObservation(10 ** 6, 2, (310.0, 205.0))
    st3 = soa.localBundleAdjustement(_Cam())
    assert soa._engine[1] is eng and eng.prob.n_pts == p.n_pts + 1
    ref3 = soa.problem((_Cam.fx, _Cam.fy, _Cam.cx, _Cam.cy))
    assert abs(oba.cost(ref3) - st3["final_cost"]) <= 1e-9 * st3["final_cost"]
    # motion only
    p2 = synthetic_ba(5, 300, 4.0, seed=14, motion_only=True)
    m2 = _map_from_problem(p2); m2.frames[0].kf = True
    soa2 = MapSoA.from_map(m2)
    ba2 = BundleAdjustment(_Cam()); ba2.motionOnlyBundleAdjustement(m2)
    ref2 = soa2.problem((_Cam.fx, _Cam.fy, _Cam.cx, _Cam.cy), motion_only=True)
    h2 = oba.optimize(ref2, 10)
    st4 = soa2.motionOnlyBundleAdjustement(_Cam())
    assert st4["final_cost"] == ba2.stats["final_cost"]
    assert abs(st4["final_cost"] - h2["final_cost"]) <= 1e-9 * h2["final_cost"]
    soa.close(); soa2.close()
