"""Feature extraction.  CPU: the oracle against the reference's own FeatureExtractor outputs
(golden) and against live cv2, stage by stage.  GPU: the CUDA path through the C ABI against the oracle, the
golden vectors and live cv2 -- bit-exact key points (order included) and descriptors."""
import os

import cv2
import numpy as np
import pytest

from oracle import features as of

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "features_icl.npz")
NAMES = ["icl1", "icl32_crop"]


def _cv2_features(img, max_corners=3000, quality=0.01, min_distance=7):
    """frame.py:10-14 with cv2.ORB_create() (frame.py:8), straight cv2 calls."""
    pts = cv2.goodFeaturesToTrack(np.mean(img, axis=2).astype(np.uint8), max_corners, qualityLevel=quality,
                                  minDistance=min_distance)
    pts = np.zeros((0, 1, 2), np.float32) if pts is None else pts
    kps = [cv2.KeyPoint(x=float(f[0][0]), y=float(f[0][1]), size=20) for f in pts]
    kp, des = cv2.ORB_create().compute(img, kps)
    des = np.zeros((0, 32), np.uint8) if des is None else des
    return pts.reshape(-1, 2), np.asarray(cv2.KeyPoint_convert(kp), np.float32).reshape(-1, 2), des


def _images():
    rng = np.random.default_rng(17)
    g = np.load(GOLD)
    yield "icl1", g["icl1_img"]
    yield "noise", rng.integers(0, 256, (120, 200, 3), dtype=np.uint8)                     # > 3000 candidates / cap
    yield "smooth_odd", cv2.GaussianBlur(rng.integers(0, 256, (211, 333, 3), dtype=np.uint8), (5, 5), 1.2)
    flat = np.full((100, 140, 3), 77, np.uint8)
    yield "flat", flat                                                                       # no corners at all
    sq = flat.copy(); sq[40:60, 50:90] = 200
    yield "plateaus", sq                                                                     # exact ties
    chk = (np.indices((160, 224)).sum(0) // 16 % 2 * 255).astype(np.uint8)
    yield "checker", np.repeat(chk[:, :, None], 3, axis=2)                                   # many equal responses


def test_constants_match_cv2():
    assert np.array_equal(of.GAUSS7, cv2.getGaussianKernel(7, 2, cv2.CV_32F).ravel())
    from visual_slam_b200.orb_pattern import ORB_PATTERN
    assert np.array_equal(ORB_PATTERN, of.load_pattern())
    # the table against cv2 itself: a descriptor of a random patch at angle 0 read back pair by pair
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (101, 101), dtype=np.uint8)
    _, des = cv2.ORB_create().compute(img, [cv2.KeyPoint(x=50.0, y=50.0, size=20, angle=0.0)])
    bl = cv2.sepFilter2D(img, cv2.CV_8U, of.GAUSS7, of.GAUSS7, borderType=cv2.BORDER_REFLECT_101).astype(int)
    p = ORB_PATTERN.astype(int)
    bits = bl[50 + p[0::2, 1], 50 + p[0::2, 0]] < bl[50 + p[1::2, 1], 50 + p[1::2, 0]]
    assert np.array_equal(np.packbits(bits, bitorder="little"), des[0])


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_reference_feature_extractor(name):
    g = np.load(GOLD)
    img = g[name + "_img"]
    corners = of.good_features_to_track(of.gray_mean(img))
    assert np.array_equal(corners, g[name + "_corners"])
    kp, des = of.compute_features(img)
    assert kp.dtype == np.float32 and np.array_equal(kp, g[name + "_kp"])
    assert np.array_equal(des, g[name + "_des"])


def test_oracle_stages_match_live_cv2():
    for name, img in _images():
        gm = of.gray_mean(img)
        assert np.array_equal(gm, np.mean(img, axis=2).astype(np.uint8)), name
        assert np.array_equal(of.gray_bgr(img), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)), name
        eig = of.corner_min_eigen_val(gm)
        assert np.array_equal(eig, cv2.cornerMinEigenVal(gm, 3, ksize=3)), name
        gb = of.gray_bgr(img)
        assert np.array_equal(of.gaussian_blur7(gb), cv2.sepFilter2D(gb, cv2.CV_8U, of.GAUSS7, of.GAUSS7,
                                                                     borderType=cv2.BORDER_REFLECT_101)), name
        corners, kp, des = _cv2_features(img)
        assert np.array_equal(of.good_features_to_track(gm, eig=eig), corners), name
        okp, odes = of.compute_features(img)
        assert np.array_equal(okp, kp) and np.array_equal(odes, des), name


def test_oracle_other_arguments_and_angles():
    g = np.load(GOLD)
    img = g["icl32_crop_img"]
    gm = of.gray_mean(img)
    for mc, q, d in [(50, 0.01, 7), (3000, 0.1, 3), (3000, 0.001, 0), (200, 0.02, 12)]:
        want = cv2.goodFeaturesToTrack(gm, mc, qualityLevel=q, minDistance=d).reshape(-1, 2)
        assert np.array_equal(of.good_features_to_track(gm, mc, q, float(d)), want), (mc, q, d)
    pts = of.good_features_to_track(gm)
    pts = pts[of.border_keep(pts, img.shape[1], img.shape[0])]
    bl = of.gaussian_blur7(of.gray_bgr(img))
    for ang in (0.0, -1.0, 33.3, 180.0, 271.0):
        kps = [cv2.KeyPoint(x=float(p[0]), y=float(p[1]), size=20, angle=ang) for p in pts]
        _, des = cv2.ORB_create().compute(img, kps)
        assert np.array_equal(of.orb_descriptors(bl, pts, np.full(len(pts), ang, np.float32)), des), ang


# ------------------------------------------------------------------------------------------ GPU
def _gpu_run(imgs, **kw):
    import torch
    from visual_slam_b200.features import extract_features_batch
    o = extract_features_batch(torch.from_numpy(np.ascontiguousarray(imgs)).cuda(), **kw)
    torch.cuda.synchronize()
    assert int(o["status"].abs().sum()) == 0
    return {k: v.cpu().numpy() for k, v in o.items() if not k.startswith("_")}


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_gpu_matches_reference_golden(name):
    from visual_slam_b200.features import FeatureExtractor
    g = np.load(GOLD)
    kp, des = FeatureExtractor().compute_features(g[name + "_img"])
    assert kp.dtype == np.float32 and des.dtype == np.uint8
    assert np.array_equal(kp, g[name + "_kp"])
    assert np.array_equal(des, g[name + "_des"])


@pytest.mark.gpu
def test_gpu_stages_bit_exact_vs_oracle_and_cv2():
    for name, img in _images():
        o = _gpu_run(img[None])
        gm = of.gray_mean(img)
        eig = of.corner_min_eigen_val(gm)
        assert np.array_equal(o["eig"][0].view(np.uint32), eig.view(np.uint32)), name
        assert np.array_equal(o["blur"][0], of.gaussian_blur7(of.gray_bgr(img))), name
        corners, kp, des = _cv2_features(img)
        n = int(o["n_corners"][0])
        assert n == len(corners) and np.array_equal(o["corners"][0, :n], corners), name
        m = int(o["n_kp"][0])
        assert m == len(kp) and np.array_equal(o["kp"][0, :m], kp), name
        assert np.array_equal(o["desc"][0, :m], des), name
        okp, odes = of.compute_features(img)
        assert np.array_equal(o["kp"][0, :m], okp) and np.array_equal(o["desc"][0, :m], odes), name


@pytest.mark.gpu
def test_gpu_batch_equals_single_frames_and_other_arguments():
    rng = np.random.default_rng(5)
    imgs = np.stack([cv2.GaussianBlur(rng.integers(0, 256, (240, 320, 3), dtype=np.uint8), (5, 5), s)
                     for s in (0.8, 1.2, 2.0, 3.0, 1.0)])
    for kw in [dict(), dict(max_corners=100, quality=0.05, min_distance=3.0), dict(min_distance=0.0, max_corners=5000),
               dict(angle=33.3)]:
        o = _gpu_run(imgs, **kw)
        mc, q, d = kw.get("max_corners", 3000), kw.get("quality", 0.01), kw.get("min_distance", 7.0)
        for b, img in enumerate(imgs):
            gm = of.gray_mean(img)
            want = cv2.goodFeaturesToTrack(gm, mc, qualityLevel=q, minDistance=d).reshape(-1, 2)
            n = int(o["n_corners"][b])
            assert n == len(want) and np.array_equal(o["corners"][b, :n], want), (kw, b)
            pts = want[of.border_keep(want, img.shape[1], img.shape[0])]
            kps = [cv2.KeyPoint(x=float(p[0]), y=float(p[1]), size=20, angle=kw.get("angle", -1.0)) for p in pts]
            _, des = cv2.ORB_create().compute(img, kps)
            m = int(o["n_kp"][b])
            assert m == len(pts) and np.array_equal(o["kp"][b, :m], pts), (kw, b)
            assert np.array_equal(o["desc"][b, :m], des), (kw, b)


@pytest.mark.gpu
def test_gpu_full_size_batch_properties():
    """64 VGA frames: idempotence (same input -> same output), corners sorted by response, pairwise distance >= 7,
    every key point inside ORB's border, descriptors of the matching pipeline's shape."""
    rng = np.random.default_rng(9)
    base = cv2.GaussianBlur(rng.integers(0, 256, (480, 640, 3), dtype=np.uint8), (7, 7), 2.0)
    imgs = np.stack([np.roll(base, 3 * i, axis=1) for i in range(64)])
    o1, o2 = _gpu_run(imgs), _gpu_run(imgs)
    for k in ("n_corners", "n_kp", "eig", "blur"):
        assert np.array_equal(o1[k], o2[k])
    for b in (0, 31, 63):
        n, m = int(o1["n_corners"][b]), int(o1["n_kp"][b])
        assert np.array_equal(o1["corners"][b, :n], o2["corners"][b, :n])
        assert np.array_equal(o1["desc"][b, :m], o2["desc"][b, :m])
        c = o1["corners"][b, :n].astype(np.int64)
        resp = o1["eig"][b][c[:, 1], c[:, 0]]
        assert np.all(np.diff(resp) <= 0)
        d2 = ((c[:, None, :] - c[None, :, :]) ** 2).sum(-1) + np.eye(n, dtype=np.int64) * 10 ** 6
        assert d2.min() >= 49
        kp = o1["kp"][b, :m]
        assert kp[:, 0].min() >= 31 and kp[:, 0].max() < 640 - 31 and kp[:, 1].min() >= 31 and kp[:, 1].max() < 480 - 31
    corners, kp, des = _cv2_features(imgs[40])
    m = int(o1["n_kp"][40])
    assert m == len(kp) and np.array_equal(o1["kp"][40, :m], kp) and np.array_equal(o1["desc"][40, :m], des)


@pytest.mark.gpu
def test_gpu_rejects_bad_arguments():
    import torch
    from visual_slam_b200.features import FeatureExtractor, extract_features_batch
    with pytest.raises(ValueError):
        FeatureExtractor().compute_features(np.zeros((480, 640), np.uint8))
    with pytest.raises(ValueError):
        extract_features_batch(torch.zeros((1, 64, 64, 3), dtype=torch.uint8, device="cuda"), min_distance=7.5)
    with pytest.raises(ValueError):
        extract_features_batch(torch.zeros((1, 64, 64, 3), dtype=torch.float32, device="cuda"))


def test_oracle_front_end_pair_matches_reference():
    """C1 front end on ICL-NUIM frames 1 -> 32: oracle features + oracle matcher == the reference's own classes."""
    from oracle import matching as om
    g = np.load(GOLD)
    kp1, des1 = of.compute_features(g["icl1_img"])
    kp2, des2 = of.compute_features(g["icl32_img"])
    assert np.array_equal(kp2, g["icl32_kp"]) and np.array_equal(des2, g["icl32_des"])
    idx, dist = om.knn2_numpy(des1, des2)
    kq, kt = om.filter_numpy(idx, dist, 0.8, None)
    assert np.array_equal(kq, g["pair_query"]) and np.array_equal(kt, g["pair_train"])


@pytest.mark.gpu
def test_gpu_front_end_pair_matches_reference():
    """The drop-in chain FeatureExtractor -> FeatureMatcher on the device against the reference's own chain
    (tests/golden/make_features_golden.py): same key points, descriptors, matches and gathered arrays."""
    from visual_slam_b200 import FeatureMatcher
    from visual_slam_b200.features import FeatureExtractor
    g = np.load(GOLD)
    fx, fm = FeatureExtractor(), FeatureMatcher()
    kp1, des1 = fx.compute_features(g["icl1_img"])
    kp2, des2 = fx.compute_features(g["icl32_img"])
    assert np.array_equal(kp2, g["icl32_kp"]) and np.array_equal(des2, g["icl32_des"])
    matches, p1, f1, p2, f2 = fm.match_features(kp1, des1, kp2, des2)
    assert np.array_equal([m[0].queryIdx for m in matches], g["pair_query"])
    assert np.array_equal([m[0].trainIdx for m in matches], g["pair_train"])
    assert np.array_equal(np.array([m[0].distance for m in matches], np.float32), g["pair_dist"])
    assert np.array_equal(p1, g["pair_pts1"]) and np.array_equal(p2, g["pair_pts2"])
    assert np.array_equal(f1, g["pair_ft1"]) and np.array_equal(f2, g["pair_ft2"])


@pytest.mark.gpu
def test_gpu_host_extractor_equals_direct_call():
    """HostFeatureExtractor (pinned staging, chunks on two streams, reused buffers) == one direct batch call."""
    import torch
    from visual_slam_b200 import synth
    from visual_slam_b200.features import HostFeatureExtractor
    frames = synth.frames(7, 120, 160, seed=21)
    want = _gpu_run(frames)
    hx = HostFeatureExtractor(7, 120, 160, chunks=3)
    hx.load(frames)
    for _ in range(2):                      # second run reuses every buffer
        kp, desc, n_kp = hx.run()
        torch.cuda.synchronize()
        assert np.array_equal(n_kp.numpy(), want["n_kp"])
        for b in range(7):
            n = int(n_kp[b])
            assert np.array_equal(kp[b, :n].numpy(), want["kp"][b, :n])
            assert np.array_equal(desc[b, :n].numpy(), want["desc"][b, :n])
    kp.zero_(); desc.zero_(); n_kp.zero_()
    for _ in range(3):                      # pipelined calls (the bench's end-to-end loop), joined once
        kp, desc, n_kp = hx.run(join=False)
    hx.join()
    torch.cuda.synchronize()
    assert np.array_equal(n_kp.numpy(), want["n_kp"])
    for b in range(7):
        n = int(n_kp[b])
        assert np.array_equal(kp[b, :n].numpy(), want["kp"][b, :n])
        assert np.array_equal(desc[b, :n].numpy(), want["desc"][b, :n])


@pytest.mark.gpu
def test_gpu_large_batch_of_odd_frames_uses_throughput_variant():
    """100 frames of 75 x 203 (odd width: scalar-tail columns of OpenCV's row filters; enough strips that the
    one-thread-per-column response kernel is dispatched): response bit-exact vs the oracle, features vs cv2."""
    rng = np.random.default_rng(31)
    imgs = np.stack([cv2.GaussianBlur(rng.integers(0, 256, (75, 203, 3), dtype=np.uint8), (3, 3), 0.8)
                     for _ in range(100)])
    o = _gpu_run(imgs)
    for b in (0, 57, 99):
        eig = of.corner_min_eigen_val(of.gray_mean(imgs[b]))
        assert np.array_equal(o["eig"][b].view(np.uint32), eig.view(np.uint32)), b
        corners, kp, des = _cv2_features(imgs[b])
        n, m = int(o["n_corners"][b]), int(o["n_kp"][b])
        assert n == len(corners) and np.array_equal(o["corners"][b, :n], corners), b
        assert m == len(kp) and np.array_equal(o["kp"][b, :m], kp) and np.array_equal(o["desc"][b, :m], des), b


@pytest.mark.gpu
def test_gpu_both_response_kernels_are_bit_identical():
    """The warp-specialised and the one-thread-per-column response kernels forced on the same frames."""
    from visual_slam_b200 import _cabi
    rng = np.random.default_rng(41)
    imgs = np.stack([cv2.GaussianBlur(rng.integers(0, 256, (97, 203, 3), dtype=np.uint8), (3, 3), 0.9)
                     for _ in range(6)])
    lib = _cabi.lib()
    try:
        outs = []
        for variant in (1, 2):
            _cabi.check(lib.vslam_features_set_variant(variant))
            outs.append(_gpu_run(imgs))
    finally:
        _cabi.check(lib.vslam_features_set_variant(0))
    eig = of.corner_min_eigen_val(of.gray_mean(imgs[3]))
    for o in outs:
        assert np.array_equal(o["eig"][3].view(np.uint32), eig.view(np.uint32))
    assert np.array_equal(outs[0]["eig"].view(np.uint32), outs[1]["eig"].view(np.uint32))
    assert np.array_equal(outs[0]["n_kp"], outs[1]["n_kp"])


@pytest.mark.gpu
def test_gpu_tiny_and_large_frames():
    """A 40 x 50 frame (bands thinner than the suppression radius, no key point survives ORB's border), a
    1280 x 960 frame (status band > 48 KB: opt-in shared memory) and a frame too large for the band."""
    import torch
    from visual_slam_b200.features import extract_features_batch
    rng = np.random.default_rng(77)
    tiny = cv2.GaussianBlur(rng.integers(0, 256, (40, 50, 3), dtype=np.uint8), (3, 3), 0.8)
    o = _gpu_run(tiny[None])
    corners, kp, des = _cv2_features(tiny)
    n = int(o["n_corners"][0])
    assert n == len(corners) > 0 and np.array_equal(o["corners"][0, :n], corners)
    assert int(o["n_kp"][0]) == len(kp) == 0
    big = cv2.resize(synth_frame(), (1280, 960), interpolation=cv2.INTER_LINEAR)
    o = _gpu_run(big[None])
    corners, kp, des = _cv2_features(big)
    n, m = int(o["n_corners"][0]), int(o["n_kp"][0])
    assert n == len(corners) and np.array_equal(o["corners"][0, :n], corners)
    assert m == len(kp) and np.array_equal(o["kp"][0, :m], kp) and np.array_equal(o["desc"][0, :m], des)
    with pytest.raises(ValueError):
        extract_features_batch(torch.zeros((1, 1200, 1600, 3), dtype=torch.uint8, device="cuda"))


def synth_frame():
    from visual_slam_b200 import synth
    return synth.frames(1, 480, 640, seed=12)[0]


@pytest.mark.gpu
def test_gpu_map_initialisation_chain_on_icl_pair():
    """C1 slice, frames 1 -> 32: device features -> device matches -> estimateEssential -> estimateRelativePose,
    against the reference's own chain (tests/golden/make_features_golden.py).  The matches are identical; the
    essential-matrix RANSAC draws its own samples (cv2's are not seedable), so E is checked through its inlier
    support, and recoverPose is checked on the reference's E, where it is deterministic."""
    from visual_slam_b200 import FeatureMatcher
    from visual_slam_b200.features import FeatureExtractor
    from visual_slam_b200.twoview import estimateEssential, estimateRelativePose
    g = np.load(GOLD)
    K = np.matrix([[481.20, 0, 319.5], [0, 480.0, 239.5], [0, 0, 1]])
    fx, fm = FeatureExtractor(), FeatureMatcher()
    kp1, des1 = fx.compute_features(g["icl1_img"])
    kp2, des2 = fx.compute_features(g["icl32_img"])
    _, p1, f1, p2, f2 = fm.match_features(kp1, des1, kp2, des2)
    assert np.array_equal(p1, g["pair_pts1"]) and np.array_equal(p2, g["pair_pts2"])
    E, inl, _ = estimateEssential(p1, p2, K, essTh=3.0 / K[0, 0])
    n_ref = int(g["pair_inliers"].sum())
    assert E is not None and abs(np.linalg.norm(E) - 1.0) < 1e-9
    assert int(inl.sum()) >= 0.9 * n_ref, (int(inl.sum()), n_ref)
    sel = g["pair_inliers"].ravel() == 1
    R, t, frac, tri, ip1, ip2, if1, if2 = estimateRelativePose(g["pair_E"], p1[sel], p2[sel], f1[sel], f2[sel], K,
                                                               "Essential")
    assert np.abs(np.asarray(R) - g["pair_R"]).max() < 1e-6
    assert np.abs(np.asarray(t).ravel() - g["pair_t"].ravel()).max() < 1e-6
    assert abs(frac - float(g["pair_valid_fraction"])) < 1e-12


@pytest.mark.gpu
def test_gpu_frontend_batch_device_resident_chain():
    """features.frontend_batch (frames -> features of both -> ragged matching, nothing leaves the device) on the
    ICL-NUIM pair and on a second, different pair in the same batch, against the reference's chain / the oracle."""
    import torch
    from oracle import matching as om
    from visual_slam_b200 import synth
    from visual_slam_b200.features import frontend_batch
    g = np.load(GOLD)
    other = synth.frames(2, 480, 640, seed=55)
    a = np.stack([g["icl1_img"], other[0]])
    b = np.stack([g["icl32_img"], np.roll(other[0], 5, axis=1)])
    fa, fb, m = frontend_batch(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda())
    torch.cuda.synchronize()
    n = int(m["n_keep"][0])
    assert np.array_equal(m["keep_q"][0, :n].cpu().numpy(), g["pair_query"])
    assert np.array_equal(m["keep_t"][0, :n].cpu().numpy(), g["pair_train"])
    na, nb = int(fa["n_kp"][1]), int(fb["n_kp"][1])
    oi, od = om.knn2_numpy(fa["desc"][1, :na].cpu().numpy(), fb["desc"][1, :nb].cpu().numpy())
    kq, kt = om.filter_numpy(oi, od, 0.8, None)
    n1 = int(m["n_keep"][1])
    assert n1 == len(kq) > 100 and np.array_equal(m["keep_q"][1, :n1].cpu().numpy(), kq)
    assert np.array_equal(m["keep_t"][1, :n1].cpu().numpy(), kt)
