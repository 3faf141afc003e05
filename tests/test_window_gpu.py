"""GPU: the device-resident tracking window (csrc/ba_window.cuh, visual_slam_b200/window.py).

"Add a tracked frame and its observations, then motion-only BA" -- what the reference does after every frame
(src/v2/main.py:208-214, src/v2/LocalBA.py:195-229) -- replayed for whole key-frame windows and checked, frame by
frame, against the CPU restatement: on the ICL-NUIM record (tests/golden/c1_replay.npz, where the reference's own
LocalBA.py built each graph) and on a synthetic window through oracle.ba.optimize run live."""
import os

import numpy as np
import pytest
import torch

from oracle import ba as oba
from oracle import poses as opo
from visual_slam_b200.ba_problem import BAProblem, R_to_quat, quat_to_R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c1_replay.npz")
K4 = (481.2, 480.0, 319.5, 239.5)


def _T(c):
    T = np.eye(4)
    T[:3, :3] = quat_to_R(c[:4])
    T[:3, 3] = c[4:]
    return T


@pytest.mark.parametrize("kf_index,first,last", [(18, 19, 39), (39, 40, 60), (60, 61, 81)])
def test_window_replays_every_tracked_frame_of_the_icl_record(kf_index, first, last):
    from visual_slam_b200.window import TrackingWindow
    g = np.load(GOLD)
    pts = g["ba_motion_%d_points" % first]
    win = TrackingWindow(max_frames=32, slot=512, max_points=512)
    win.reset(pts, g["match_track_%d_kp1" % first], g["ba_motion_%d_poses" % first][0], K4)
    worst_pose, worst_cost, same = 0.0, 0.0, 0
    for i in range(first, last + 1):
        q, t = g["match_track_%d_q" % i], g["match_track_%d_t" % i]
        win.append(g["pose_plumbing_%d_W_T_curr" % i], q, t, np.array([len(q)], np.int32), g["kp_%d" % i])
        st = win.motion_only_ba(10)
        ref = g["ba_motion_%d_poses" % i]
        assert win.n_frames == len(ref) == i - first + 2
        cams = win.cams().cpu().numpy()
        d = max(np.abs(_T(c) - T).max() for c, T in zip(cams, ref))
        fc = float(g["ba_motion_%d_stat_final_cost" % i])
        worst_pose, worst_cost = max(worst_pose, d), max(worst_cost, abs(st["final_cost"] - fc) / fc)
        assert d < 1e-6 and abs(st["final_cost"] - fc) <= 1e-4 * fc, (i, d, st["final_cost"], fc)
        assert abs(st["initial_cost"] - float(g["ba_motion_%d_stat_initial_cost" % i])) <= 1e-8 * st["initial_cost"]
        same += (int(st["iterations"]) == int(g["ba_motion_%d_stat_iterations" % i]) and
                 int(st["trials"]) == int(g["ba_motion_%d_stat_trials" % i]))
    n = last - first + 1
    print("window %d: %d frames, worst pose diff %.2e, worst rel. cost diff %.2e, same LM path %d/%d"
          % (kf_index, n, worst_pose, worst_cost, same, n))
    # (iterations, trials) differ only where the cost has stagnated and rho is rounding noise (measured: 15 of 21 equal
    # in the first window, poses 4.9e-10, cost 1.9e-15 rel.)
    assert same >= 0.5 * n
    win.close()


def _synthetic_window(seed, F=12, P=300, drop=0.3):
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = K4
    pts = np.stack([rng.uniform(-2, 2, P), rng.uniform(-1.5, 1.5, P), rng.uniform(3, 7, P)], axis=1)
    cams_true, cams = [], []
    for f in range(F):
        r = rng.normal(size=3) * 0.02 * f
        T = opo.camera_to_world(r, rng.normal(size=3) * 0.03 * f)
        cams_true.append(T)
        Tn = opo.camera_to_world(r + rng.normal(size=3) * 0.01 * (f > 0), T[:3, :3].T @ -T[:3, 3] * 0 + rng.normal(size=3) * 0.02 * (f > 0))
        Tn[:3, 3] = T[:3, 3] + rng.normal(size=3) * 0.02 * (f > 0)
        cams.append(Tn)
    obs = []
    for f in range(F):
        T = np.linalg.inv(cams_true[f])
        pc = pts @ T[:3, :3].T + T[:3, 3]
        uv = np.stack([fx * pc[:, 0] / pc[:, 2] + cx, fy * pc[:, 1] / pc[:, 2] + cy], axis=1) + rng.normal(0, 0.7, (P, 2))
        keep = np.nonzero(rng.random(P) > (0.0 if f == 0 else drop))[0]
        bad = rng.random(len(keep)) < 0.1
        uv[keep[bad]] += rng.uniform(-40, 40, (int(bad.sum()), 2))
        obs.append((keep.astype(np.int32), uv[keep].astype(np.float32)))
    return pts, cams, obs


def test_window_equals_live_restatement_on_a_synthetic_window():
    from visual_slam_b200.window import TrackingWindow
    pts, cams, obs = _synthetic_window(3)
    win = TrackingWindow(max_frames=16, slot=300, max_points=300)
    win.reset(pts, obs[0][1], cams[0], K4)
    cur = [np.concatenate([R_to_quat(cams[0][:3, :3]), cams[0][:3, 3]])]
    same = 0
    for f in range(1, len(cams)):
        rows, uv = obs[f]
        # the frame's observations arrive as (kept query rows, kept train rows) into the frame's key-point array
        perm = np.random.default_rng(f).permutation(len(rows)).astype(np.int32)
        kp_cur = np.zeros((len(rows), 2), np.float32); kp_cur[perm] = uv
        win.append(cams[f], rows, perm, np.array([len(rows)], np.int32), kp_cur)
        st = win.motion_only_ba(10)
        cur.append(np.concatenate([R_to_quat(cams[f][:3, :3]), cams[f][:3, 3]]))
        oc = np.concatenate([np.full(len(obs[k][0]), k, np.int32) for k in range(f + 1)])
        op = np.concatenate([obs[k][0] for k in range(f + 1)])
        ouv = np.concatenate([obs[k][1] for k in range(f + 1)]).astype(np.float64)
        fixed = np.zeros(f + 1, bool); fixed[0] = True
        prob = BAProblem(np.array(cur), pts, oc, op, ouv, K4, fixed, np.ones(len(pts), bool))
        h = oba.optimize(prob, 10)
        got = win.cams().cpu().numpy()
        assert abs(st["final_cost"] - h["final_cost"]) <= 1e-9 * h["final_cost"], (f, st, h["final_cost"])
        assert abs(st["initial_cost"] - h["cost"][0]) <= 1e-11 * h["cost"][0]
        assert np.abs(got - prob.cams).max() < 1e-8, (f, np.abs(got - prob.cams).max())
        same += (int(st["iterations"]), int(st["trials"])) == (h["iterations"], h["trials"])
        cur = [c.copy() for c in prob.cams]          # both sides continue from the optimised window (main.py:213-214)
    # (iterations, trials) agree only while rho is not rounding noise: on this window the optimum is reached after a few
    # iterations and the stopping test rho == 0 fires at different trials; cost (1e-9) and poses (1e-8) above are the check
    print("synthetic window: same (iterations, trials) in %d of %d frames" % (same, len(cams) - 1))
    # no iterations: the cost of the window as it stands; a second run is bit-identical (fixed-order reductions)
    c0 = win.motion_only_ba(0)
    assert c0["initial_cost"] == c0["final_cost"] and c0["iterations"] == 0
    a = win.cams().clone(); s1 = win.motion_only_ba(3); b = win.cams().clone()
    win2 = TrackingWindow(max_frames=16, slot=300, max_points=300)
    win2.reset(pts, obs[0][1], a[0].cpu().numpy(), K4)
    for f in range(1, len(cams)):
        rows, uv = obs[f]
        win2.append(a[f], rows, np.arange(len(rows), dtype=np.int32), np.array([len(rows)], np.int32), uv)
    s2 = win2.motion_only_ba(3)
    assert s1 == s2 and torch.equal(win2.cams(), b)
    win.close(); win2.close()


def test_window_limits_and_errors():
    from visual_slam_b200 import VslamError
    from visual_slam_b200.window import TrackingWindow
    pts, cams, obs = _synthetic_window(5, F=4, P=50)
    win = TrackingWindow(max_frames=3, slot=40, max_points=64)
    with pytest.raises(VslamError):
        win.append(cams[1], obs[1][0], obs[1][0], np.array([3], np.int32), obs[1][1])       # before reset
    win.reset(pts, obs[0][1], cams[0], K4)
    rows = np.arange(50, dtype=np.int32)
    kp = np.tile(obs[0][1], (2, 1))
    win.append(cams[1], rows, rows, np.array([50], np.int32), kp)          # 50 matches into a 40-observation slot: clamped
    win.append(cams[2], np.array([7, 999, -1, 3], np.int32), np.array([0, 1, 2, 3], np.int32), np.array([4], np.int32), kp)
    st = win.motion_only_ba(5)                                            # rows outside the point list are ignored
    assert np.isfinite(st["final_cost"]) and st["final_cost"] <= st["initial_cost"]
    with pytest.raises(VslamError):
        win.append(cams[3], rows, rows, np.array([5], np.int32), kp)      # window full
    with pytest.raises(ValueError):
        win.reset(np.zeros((100, 3)), np.zeros((100, 2), np.float32), cams[0], K4)   # more points than capacity
    with pytest.raises(VslamError):
        TrackingWindow(max_frames=10 ** 6)                                 # one co-resident CTA per frame
    win.close()


def test_track_runs_the_whole_per_frame_step_on_the_device():
    """window.track = match -> gather -> RANSAC PnP -> camera->world -> append -> motion-only BA without a host round
    trip; equal to the same stages called one by one."""
    from visual_slam_b200 import synth
    from visual_slam_b200.matching import match_batch
    from visual_slam_b200.window import TrackingWindow
    d = synth.c5_pairs(2, 0, 6, 512, device="cuda")
    fx, fy, cx, cy = K4
    X = d["map_xyz"][0].double()
    uv0 = torch.stack([fx * X[:, 0] / X[:, 2] + cx, fy * X[:, 1] / X[:, 2] + cy], dim=1).float()   # key frame = identity
    win = TrackingWindow(max_frames=8, slot=512, max_points=512)
    win.reset(X, uv0, np.eye(4), K4)
    win.set_keyframe_features(d["desc_kf"][0])
    words = torch.from_numpy(np.random.default_rng(4).integers(0, 2 ** 32, size=(100, 4), dtype=np.uint64)
                             .astype(np.uint32).view(np.int32)).cuda()
    # frame 1 of the sequence sees the SAME world points (pair 0 of the generator): track it twice, pipeline vs stages
    o = win.track(d["desc_cur"][0], d["kp_cur"][0], words, want_stats=True)
    m = match_batch(d["desc_kf"][0:1], d["desc_cur"][0:1], 0.8, False)
    assert torch.equal(o["n_keep"], m["n_keep"]) and torch.equal(o["keep_q"], m["keep_q"])
    assert win.n_frames == 2 and o["stats"]["final_cost"] <= o["stats"]["initial_cost"]
    pose = o["pose"][0].cpu().numpy()
    assert np.abs(pose - d["pose_true"][0].cpu().numpy()).max() < 5e-3
    W = opo.camera_to_world(pose[:3], pose[3:])
    assert np.abs(o["c2w"][0].cpu().numpy() - W).max() < 1e-14
    after = _T(win.cams()[1].cpu().numpy())
    assert np.abs(after - W).max() < 2e-2                                   # BA moved the pose, but not far
    win.close()
