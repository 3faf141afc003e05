"""GPU: two-view path (undistortPoints, essential-matrix RANSAC, recoverPose) against the oracle
(oracle/essential.py, pinned to live cv2 and golden G1 in tests/test_oracle_essential.py)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from oracle import essential as oe  # noqa: E402
from visual_slam_b200 import synth, twoview  # noqa: E402

pytestmark = pytest.mark.gpu
K = synth.ICL_K
G1 = os.path.join(os.path.dirname(__file__), "golden", "essential_g1.npz")


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _words(seed, B, H):
    w = np.random.default_rng(seed).integers(0, 2 ** 32, size=(B, H, 5), dtype=np.uint64).astype(np.uint32)
    return w, _dev(w.view(np.int32))


def test_undistort_points_bit_exact():
    p1, p2, _, _, _ = synth.two_view_pair(3, 777)
    for p in (p1, p2):
        got = twoview.undistort_points(_dev(p), K).cpu().numpy()
        assert np.array_equal(got, oe.undistort_points_f32(p, K))


def test_five_point_models_match_oracle_on_the_same_samples():
    n, H = 300, 64
    p1, p2, _, _, _ = synth.two_view_pair(5, n)
    x1 = oe.undistort_points_f32(p1, K); x2 = oe.undistort_points_f32(p2, K)
    w, wd = _words(1, 1, H)
    models, n_models = twoview.essential_hypotheses(_dev(x1)[None], _dev(x2)[None],
                                                    torch.tensor([n], dtype=torch.int32, device="cuda"), wd)
    models = models[0].cpu().numpy().reshape(H, 10, 3, 3); n_models = n_models[0].cpu().numpy()
    samples = oe.draw_samples(w[0], n)
    same_count, worst = 0, 0.0
    for h in range(H):
        s = samples[h]
        Eo = oe.five_point(x1[s].astype(np.float64), x2[s].astype(np.float64), polish=True)
        Eg = models[h, :n_models[h]]
        # every device model is an exact solution: epipolar constraint on the sample, essential constraint
        for E in Eg:
            h1 = np.c_[x1[s].astype(np.float64), np.ones(5)]; h2 = np.c_[x2[s].astype(np.float64), np.ones(5)]
            assert np.abs(np.einsum("ni,ij,nj->n", h2, E, h1)).max() < 1e-12
            assert np.abs(2 * E @ E.T @ E - np.trace(E @ E.T) * E).max() < 1e-10
            assert abs(np.linalg.norm(E) - 1) < 1e-12
        assert np.all(models[h, n_models[h]:] == 0)
        if len(Eo) == len(Eg):   # a root on the |imag| = 1e-10 edge may be kept by one side only
            same_count += 1
            if len(Eg):
                worst = max(worst, np.abs(Eg - Eo).max())   # same canonical sign AND order
    assert same_count >= H - 2
    assert worst < 1e-8


def _ransac_case(seed, n, H, outlier_frac=0.25):
    p1, p2, R, t, inl = synth.two_view_pair(seed, n, outlier_frac=outlier_frac)
    x1 = oe.undistort_points_f32(p1, K); x2 = oe.undistort_points_f32(p2, K)
    w, wd = _words(seed + 100, 1, H)
    out = twoview.essential_ransac_batch(_dev(x1)[None], _dev(x2)[None],
                                         torch.tensor([n], dtype=torch.int32, device="cuda"), wd, 3.0 / K[0])
    return p1, p2, x1, x2, R, t, inl, w, out


@pytest.mark.parametrize("seed,n,H", [(0, 400, 200), (1, 2048, 1000), (2, 37, 50)])
def test_scoring_selection_and_inlier_sets_identical_given_the_hypotheses(seed, n, H):
    p1, p2, x1, x2, R, t, inl, w, out = _ransac_case(seed, n, H)
    thr = 3.0 / K[0]
    models = out["models"][0].cpu().numpy().reshape(H, 10, 3, 3)
    n_models = out["n_models"][0].cpu().numpy()
    counts = oe.score(models, n_models, x1, x2, thr)                 # oracle scorer on the device models
    assert np.array_equal(counts, out["counts"][0].cpu().numpy())
    (bh, bm), cnt = oe.select(counts, n_models, n)
    best = int(out["best"][0])
    assert best == (bh * 10 + bm if bh >= 0 else -1)
    if bh < 0:
        return
    E = out["E"][0].cpu().numpy()
    assert np.array_equal(E, models[bh, bm])
    mask = out["mask"][0].cpu().numpy().astype(bool)
    assert np.array_equal(mask, oe.inlier_mask(E, x1, x2, thr))      # identical inlier set
    assert int(out["n_inliers"][0]) == cnt == int(mask.sum())
    want = oe.epiline_score(E, x1, x2, mask)
    assert abs(float(out["score"][0]) - want) <= 1e-9 * max(1.0, abs(want))


def test_full_oracle_ransac_agrees_end_to_end():
    # oracle solver + scorer + selection on the same samples: same winner, same inlier set, E within 1e-8
    n, H = 500, 150
    p1, p2, x1, x2, R, t, inl, w, out = _ransac_case(4, n, H)
    r = oe.find_essential_ransac(x1, x2, oe.draw_samples(w[0], n), 3.0 / K[0])
    assert r["ok"]
    E = out["E"][0].cpu().numpy()
    assert np.abs(E - r["E"]).max() < 1e-8
    assert np.array_equal(out["mask"][0].cpu().numpy().astype(bool), r["mask"])
    assert (r["mask"] & inl).sum() >= 0.9 * inl.sum()


def test_recover_pose_matches_oracle_and_truth():
    for seed in range(4):
        n, H = 600, 200
        p1, p2, x1, x2, R, t, inl, w, out = _ransac_case(seed + 10, n, H)
        E = out["E"][0].cpu().numpy()
        mask = out["mask"][0].cpu().numpy().astype(bool)
        ip1, ip2 = p1[mask], p2[mask]
        m = len(ip1)
        got = twoview.recover_pose_batch(out["E"], _dev(ip1)[None], _dev(ip2)[None],
                                         torch.tensor([m], dtype=torch.int32, device="cuda"), K, 50.0)
        g, Ro, to, mo, Qo = oe.recover_pose(E, ip1, ip2, K, 50.0)
        assert int(got["good"][0, 0]) == g
        assert np.abs(got["R"][0].cpu().numpy() - Ro).max() < 1e-6       # north_star: 1e-6 rad
        assert np.abs(got["t"][0].cpu().numpy() - to).max() < 1e-6       # 1e-6 m (unit baseline)
        assert np.array_equal(got["mask"][0].cpu().numpy() == 255, mo)
        tri = got["tri"][0].cpu().numpy()
        d = np.minimum(np.abs(tri - Qo).max(axis=0), np.abs(tri + Qo).max(axis=0))
        assert d.max() < 1e-9
        assert np.abs(Ro - R).max() < 3e-2 and np.abs(to - t / np.linalg.norm(t)).max() < 0.2


def test_batched_pairs_with_ragged_sizes_equal_single_calls():
    B, n_max, H = 5, 300, 100
    ns = [300, 211, 5, 4, 120]
    x1 = np.zeros((B, n_max, 2), np.float32); x2 = np.zeros((B, n_max, 2), np.float32)
    for b in range(B):
        p1, p2, _, _, _ = synth.two_view_pair(20 + b, n_max)
        x1[b] = oe.undistort_points_f32(p1, K); x2[b] = oe.undistort_points_f32(p2, K)
    w, wd = _words(9, B, H)
    n_pts = torch.tensor(ns, dtype=torch.int32, device="cuda")
    out = twoview.essential_ransac_batch(_dev(x1), _dev(x2), n_pts, wd, 3.0 / K[0])
    assert int(out["best"][3]) == -1 and int(out["n_inliers"][3]) == 0      # 4 points: no model
    assert not out["mask"][3].any()
    for b in (0, 1, 4):
        single = twoview.essential_ransac_batch(_dev(x1[b:b + 1, :ns[b]]), _dev(x2[b:b + 1, :ns[b]]),
                                                n_pts[b:b + 1], wd[b:b + 1], 3.0 / K[0])
        assert torch.equal(single["E"][0], out["E"][b])
        assert torch.equal(single["mask"][0], out["mask"][b, :ns[b]])
        assert not out["mask"][b, ns[b]:].any()


def test_recover_pose_of_a_batch_equals_pair_by_pair():
    """A batch of pairs with ragged sizes against the same pairs one by one: same bits."""
    B, n_max = 40, 96
    rng = np.random.default_rng(5)
    ns = rng.integers(8, n_max + 1, size=B)
    p1 = np.zeros((B, n_max, 2), np.float32); p2 = np.zeros((B, n_max, 2), np.float32)
    E = np.zeros((B, 3, 3))
    for b in range(B):
        a, c, R, t, _ = synth.two_view_pair(100 + b, n_max)
        p1[b], p2[b] = a, c
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        E[b] = (tx @ R) / np.linalg.norm(tx @ R)
    n_pts = torch.from_numpy(ns.astype(np.int32)).cuda()
    Ed = torch.from_numpy(E).cuda()
    got = twoview.recover_pose_batch(Ed, _dev(p1), _dev(p2), n_pts, K, 50.0)
    for b in range(0, B, 7):
        one = twoview.recover_pose_batch(Ed[b:b + 1], _dev(p1[b:b + 1]), _dev(p2[b:b + 1]), n_pts[b:b + 1], K, 50.0)
        for key in ("R", "t", "mask", "good", "tri"):
            assert torch.equal(one[key][0], got[key][b]), (key, b)


def test_drop_in_signatures_and_golden_g1():
    g = np.load(G1)
    # estimateRelativePose on the notebook's E and inliers (normalised coordinates, identity camera)
    p1 = g["rp_p1"].reshape(-1, 2); p2 = g["rp_p2"].reshape(-1, 2)
    f = list(range(len(p1)))
    R, t, vf, tri, k1, k2, f1, f2 = twoview.estimateRelativePose(g["rp_E"], p1, p2, f, f, np.eye(3), "Essential")
    assert np.abs(R - g["g1_R"]).max() < 5e-9 and np.abs(t.ravel() - g["g1_t"]).max() < 5e-9   # G1 digits
    assert t.shape == (3, 1) and tri.shape[0] == 4 and tri.shape[1] == len(k1) == len(f1)
    assert abs(vf - int(g["rp_n"]) / len(p1)) < 1e-12
    # estimateEssential on the notebook's pixel matches: own samples, so compare geometry not bits
    Kn = g["K"]
    pix1 = g["undist_in"].reshape(-1, 2)
    # the second undistort call holds the other image's pixels only implicitly; rebuild from fe_p2
    k4 = (Kn[0, 0], Kn[1, 1], Kn[0, 2], Kn[1, 2])
    pix2 = (g["fe_p2"].reshape(-1, 2).astype(np.float64) * [k4[0], k4[1]] + [k4[2], k4[3]]).astype(np.float32)
    E, inl, score = twoview.estimateEssential(pix1, pix2, Kn, float(g["fe_threshold"]))
    assert E.shape == (3, 3) and inl.shape == (len(pix1), 1) and inl.dtype == np.uint8
    assert abs(np.linalg.norm(E) - 1) < 1e-12 and score > 0
    agree = (inl.ravel() == g["fe_mask"].ravel()).mean()
    assert agree > 0.9
    with pytest.raises(NotImplementedError):
        twoview.estimateRelativePose(np.eye(3), p1, p2, f, f, np.eye(3), "Homography")


def test_round_based_single_pair_ransac_equals_the_full_run():
    """essential_ransac_single solves the hypotheses in rounds and stops once cv2's iteration budget
    (RANSACUpdateNumIters, started at maxIters) is covered: same model, mask and support as solving all maxIters
    hypotheses, for easy pairs (stop after the first round), hard ones (60 % wrong matches) and pure noise."""
    import torch
    from visual_slam_b200 import synth, twoview
    for seed, n, outl in ((1, 300, 0.2), (2, 400, 0.6), (3, 120, 0.75), (4, 200, 1.0)):
        p1, p2, _, _, _ = synth.two_view_pair(seed, n, outlier_frac=outl)
        both = twoview.undistort_points(torch.from_numpy(np.stack([p1, p2])).cuda(), synth.ICL_K)
        words = torch.from_numpy(np.random.default_rng(seed).integers(0, 2 ** 32, size=(1, 1000, 5), dtype=np.uint64)
                                 .astype(np.uint32).view(np.int32)).cuda()
        n_pts = torch.tensor([n], dtype=torch.int32, device="cuda")
        thr = 3.0 / synth.ICL_K[0]
        full = twoview.essential_ransac_batch(both[0:1], both[1:2], n_pts, words, thr)
        part = twoview.essential_ransac_single(both[0:1], both[1:2], n, words, thr)
        assert int(full["best"][0]) == int(part["best"][0])
        assert torch.equal(full["E"], part["E"]) and torch.equal(full["mask"], part["mask"])
        assert int(full["n_inliers"][0]) == int(part["n_inliers"][0])
        if outl <= 0.2:
            assert part["hypotheses_solved"] == 48          # cv2 would have stopped after ~20 iterations as well
        if outl >= 0.75:
            assert part["hypotheses_solved"] == 1000
