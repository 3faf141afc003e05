"""GPU: the fused tracking step (match -> gather -> RANSAC PnP) equals its stages' oracles chained."""
import numpy as np
import pytest
import torch

from oracle import matching as om
from oracle import pnp as op
from visual_slam_b200 import synth
from visual_slam_b200.pnp import unpack_inliers
from visual_slam_b200.tracking import track_batch

pytestmark = pytest.mark.gpu
K = synth.ICL_K


def test_tracking_step_matches_chained_oracles():
    B, n, H = 4, 512, 100
    data = [synth.tracking_frame_pair(s, n) for s in range(B)]
    dk, dc, X, kp = (np.stack([d[i] for d in data]) for i in range(4))
    words = np.random.default_rng(5).integers(0, 2 ** 32, size=(B, H, 4), dtype=np.uint64).astype(np.uint32)
    guess = np.zeros((B, 6))
    dev = "cuda"
    out = track_batch(torch.from_numpy(dk).to(dev), torch.from_numpy(dc).to(dev), torch.from_numpy(X).to(dev),
                      torch.from_numpy(kp).to(dev), torch.from_numpy(words.view(np.int32)).to(dev), K,
                      torch.from_numpy(guess).to(dev))
    torch.cuda.synchronize()
    g = {k: v.cpu().numpy() for k, v in out.items()}
    for b in range(B):
        idx, dist = om.knn2_numpy(dk[b], dc[b])
        kq, kt = om.filter_numpy(idx, dist, 0.8)
        nb = int(g["n_keep"][b])
        assert nb == len(kq) and np.array_equal(g["keep_q"][b, :nb], kq) and np.array_equal(g["keep_t"][b, :nb], kt)
        assert np.array_equal(g["X"][b, :nb], X[b][kq]) and np.array_equal(g["uv"][b, :nb], kp[b][kt])
        assert not g["X"][b, nb:].any() and not g["uv"][b, nb:].any()
        hyps = g["hyps"][b].reshape(-1, 3, 4); valid = g["valid"][b].astype(bool)
        masks, counts = op.score(hyps, valid, X[b][kq], kp[b][kt], K, 8.0)
        assert np.array_equal(counts, g["counts"][b])
        best, _ = op.select(counts, valid, nb, 4, 0.99)
        assert best == int(g["best"][b]) and best >= 0
        inl = np.nonzero(masks[best])[0]
        assert np.array_equal(unpack_inliers(g["inlier_bits"][b], nb), inl)
        ok, rvec, tvec = op.refine_cv2(X[b][kq], kp[b][kt], K, inl, guess[b, :3], guess[b, 3:])
        assert ok and np.abs(g["pose"][b, :3] - rvec).max() < 1e-6 and np.abs(g["pose"][b, 3:] - tvec).max() < 1e-6
        assert np.abs(g["pose"][b, :3] - data[b][4]).max() < 3e-3 and np.abs(g["pose"][b, 3:] - data[b][5]).max() < 1e-2


def test_c5_generator_is_device_independent_and_tracks():
    """C5 data (per-pair seed hash(seq, i), random-walk poses, previous pose as the guess): identical bytes
    on the CPU and on the device, any pair addressable on its own; the fused step recovers the true pose and its
    camera->world outputs equal the reference's plumbing (main.py:199-203) applied to its own PnP result."""
    from oracle import poses as opo
    cpu = synth.c5_pairs(5, 40, 46, 512, device="cpu")
    dev = synth.c5_pairs(5, 40, 46, 512, device="cuda")
    for k in cpu:
        assert torch.equal(cpu[k], dev[k].cpu()), k
    one = synth.c5_pairs(5, 43, 44, 512, device="cuda")
    assert all(torch.equal(one[k][0], dev[k][3]) for k in one)
    step = np.linalg.norm((cpu["pose_true"] - cpu["guess"]).numpy()[:, :3], axis=1)
    assert (np.rad2deg(step) <= 2.0 + 1e-9).all() and (np.rad2deg(step) > 0.1).all()
    B, H = 6, 100
    words = torch.from_numpy(np.random.default_rng(8).integers(0, 2 ** 32, size=(B, H, 4), dtype=np.uint64)
                             .astype(np.uint32).view(np.int32)).cuda()
    out = track_batch(dev["desc_kf"], dev["desc_cur"], dev["map_xyz"], dev["kp_cur"], words, K, dev["guess"],
                      want_c2w=True)
    pose = out["pose"].cpu().numpy()
    err = np.abs(pose - cpu["pose_true"].numpy())
    assert err[:, :3].max() < 2e-3 and err[:, 3:].max() < 1e-2
    assert (out["n_inliers"].cpu().numpy() > 0.8 * out["n_keep"].cpu().numpy()).all()
    for b in range(B):
        W = opo.camera_to_world(pose[b, :3], pose[b, 3:])
        assert np.abs(out["c2w"][b].cpu().numpy() - W).max() < 1e-14
        assert np.abs(out["cam7"][b].cpu().numpy() - opo.sbacam(W)).max() < 1e-14
    # the reference's way of seeding Stage B: the previous camera->world pose, turned into (rvec, tvec) on the device
    prev = torch.from_numpy(np.stack([opo.camera_to_world(g[:3], g[3:]) for g in cpu["guess"].numpy()])).cuda()
    out2 = track_batch(dev["desc_kf"], dev["desc_cur"], dev["map_xyz"], dev["kp_cur"], words, K, prev_c2w=prev)
    for b in range(B):
        r, t = opo.pose_guess(prev[b].cpu().numpy())
        assert np.abs(out2["guess"][b].cpu().numpy() - np.concatenate([r.ravel(), t])).max() < 1e-9
