"""Key-frame control logic: oracle vs the reference's own GetListDiff outputs (CPU),
CUDA list diff vs oracle (GPU), MapSoA culling vs the reference Map semantics."""
import os

import numpy as np
import pytest

from oracle import keyframe as okf

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "listdiff_ref.npz")


def _cases():
    g = np.load(GOLD)
    for c in range(5):
        yield g["kp1_%d" % c], g["kp2_%d" % c], g["idx_%d" % c].tolist()


def test_oracle_reproduces_reference_list_diff():
    n = 0
    for kp1, kp2, want in _cases():
        assert okf.get_list_diff(kp1, kp2) == want
        n += 1
    assert n == 5


def test_reference_semantics_differ_from_rowwise_on_repeated_coordinates():
    kp1, kp2, want = list(_cases())[1]      # integer corners: coordinates repeat across key points
    rows = okf.get_list_diff_rows(kp1, kp2)
    assert len(rows) == len(kp1) - len(kp2) and set(want) < set(rows)


def test_keyframe_predicates_match_reference_expression():
    from visual_slam_b200 import keyframe as kf
    for i, loop_idx, n_tr, n_known in [(25, 0, 100, 200), (25, 0, 190, 200), (10, 0, 79, 200), (10, 0, 79, 80),
                                       (21, 0, 50, 100), (20, 0, 80, 1000)]:
        want = (i - loop_idx > 20 or n_tr < 80) and (n_tr < 0.9 * n_known)      # main.py:221
        assert kf.is_keyframe(i, loop_idx, n_tr, n_known) == want == okf.is_keyframe(i, loop_idx, n_tr, n_known)
    assert [f for f in range(14) if kf.cull_due(f)] == [8, 12] == [f for f in range(14) if okf.cull_due(f)]


def _soa(rng, n_frames=6, n_pts=50):
    from visual_slam_b200.mapsoa import MapSoA
    m = MapSoA()
    for f in range(n_frames):
        m.AddFrame(f, np.eye(4), keyframe=(f % 2 == 0))
    ref = {}                                            # point id -> {frame id: uv}  (the reference's Point.frames)
    for p in range(n_pts):
        pid = 100 + p
        m.AddPoint3D(pid, rng.normal(size=3))
        ref[pid] = {}
        for f in rng.choice(n_frames, rng.integers(1, n_frames + 1), replace=False):
            uv = rng.uniform(0, 480, 2)
            m.AddObservation(pid, int(f), uv)
            ref[pid][int(f)] = uv
    return m, ref


def test_mapsoa_discard_outlier_map_points_and_frame_query():
    rng = np.random.default_rng(3)
    m, ref = _soa(rng)
    pts_before = {pid: m.GetAll3DPoints()[r].copy() for r, pid in enumerate(m._pids)}
    want_ids = [pid for pid, fr in ref.items() if len(fr) >= 3]            # map.py:124-131
    assert want_ids == okf.discard_outlier_map_points(list(ref), m._op.view(), 3)
    removed = m.DiscardOutlierMapPoints(n_visible_frames=3)
    assert removed == len(ref) - len(want_ids) and m._pids == want_ids and m.n_points == len(want_ids)
    assert all(np.array_equal(m.GetAll3DPoints()[r], pts_before[pid]) for r, pid in enumerate(m._pids))
    assert m.n_observations == sum(len(ref[pid]) for pid in want_ids)
    # map.py:28-44 on the culled map
    for f in range(6):
        uv, desc, X, ids = m.GetImagePointsWithFrameID(f)
        want = [pid for pid in want_ids if f in ref[pid]]
        assert ids.tolist() == want and desc is None
        assert all(np.array_equal(uv[k], ref[pid][f]) for k, pid in enumerate(want))
        assert all(np.array_equal(X[k], pts_before[pid]) for k, pid in enumerate(want))
    assert m.DiscardOutlierMapPoints(3) == 0


@pytest.mark.gpu
def test_cuda_list_diff_matches_reference_and_oracle():
    from visual_slam_b200.keyframe import GetListDiff
    for kp1, kp2, want in _cases():
        assert GetListDiff(kp1, kp2) == want
        assert GetListDiff(kp1, kp2, rowwise=True) == okf.get_list_diff_rows(kp1, kp2)
    rng = np.random.default_rng(5)
    # frame-sized: 3000 corners (frame.py:7), 1100 of them mapped; more than one shared-memory tile
    kp1 = np.stack([rng.integers(0, 640, 3000), rng.integers(0, 480, 3000)], axis=1).astype(np.float64)
    kp2 = kp1[rng.choice(3000, 1100, replace=False)]
    assert GetListDiff(kp1, kp2) == okf.get_list_diff(kp1, kp2)
    assert GetListDiff(kp1, kp2, rowwise=True) == okf.get_list_diff_rows(kp1, kp2)
    # edge cases: nothing mapped, nothing to select from
    assert GetListDiff(kp1[:7], np.zeros((0, 2))) == list(range(7))
    assert GetListDiff(np.zeros((0, 2)), kp2) == []
