"""GPU parity: CUDA Hamming kNN-2 + ratio + cross-check vs the oracle, bit-exact."""
import os

import numpy as np
import pytest
import torch

from oracle import matching as om
from visual_slam_b200 import FeatureMatcher, match_batch, synth

pytestmark = pytest.mark.gpu


def run_gpu(q, t, ratio=0.8, cross=True):
    out = match_batch(torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda(), ratio, cross)
    torch.cuda.synchronize()
    return {k: (v.cpu().numpy() if v is not None else None) for k, v in out.items()}


def check_pair(q, t, ratio=0.8):
    for cross in (False, True):
        g = run_gpu(q[None], t[None], ratio, cross)
        idx, dist = om.knn2_numpy(q, t)
        assert np.array_equal(g["idx"][0], idx)
        assert np.array_equal(g["dist"][0], dist)
        rev_idx = None
        if cross and q.shape[0] and t.shape[0]:
            rev_idx, rev_d = om.rev_nearest_numpy(q, t)
            key = g["rev_key"][0].view(np.uint32)
            assert np.array_equal((key & ((1 << 22) - 1)).astype(np.int32), rev_idx)
            assert np.array_equal((key >> 22).astype(np.int32), rev_d)
        kq, kt = om.filter_numpy(idx, dist, ratio, rev_idx)
        n = int(g["n_keep"][0])
        assert n == len(kq)
        assert np.array_equal(g["keep_q"][0, :n], kq) and np.array_equal(g["keep_t"][0, :n], kt)
        assert (g["keep_q"][0, n:] == -1).all()


@pytest.mark.parametrize("nq,nt", [(1, 2), (2, 2), (31, 33), (33, 31), (2000, 2049), (2049, 2000),
                                   (513, 1025), (5, 1), (1, 1)])
def test_ragged_sizes_random_and_ties(nq, nt):
    q, t, _ = synth.descriptors_pair(nq, nt, seed=nq + nt)
    check_pair(q, t)
    q, t = synth.low_entropy_pair(nq, nt, seed=nq * 7 + nt)
    check_pair(q, t, ratio=0.95)


def test_golden_reference_vectors(golden_dir):
    gold = np.load(os.path.join(golden_dir, "matching_icl_orb.npz"))
    for case, ratio in (("icl", 0.8), ("tie", 0.95)):
        q, t = gold[case + "_desc1"], gold[case + "_desc2"]
        g = run_gpu(q[None], t[None], ratio, cross=False)
        assert np.array_equal(g["idx"][0], gold[case + "_idx"])
        assert np.array_equal(g["dist"][0], gold[case + "_dist"])
        n = int(g["n_keep"][0])
        assert np.array_equal(g["keep_q"][0, :n], gold[case + "_keep_q"])
        assert np.array_equal(g["keep_t"][0, :n], gold[case + "_keep_t"])
        g = run_gpu(q[None], t[None], 10.0, cross=True)  # ratio never rejects: pure mutual-NN
        n = int(g["n_keep"][0])
        pairs = np.stack([g["keep_q"][0, :n], g["keep_t"][0, :n]], axis=1)
        keep_all = gold[case + "_dist"][:, 1] > 0  # ratio 10 still rejects d1 == 0 (0 < 0 false)
        want = np.array([r for r in gold[case + "_mutual"] if keep_all[r[0]]]).reshape(-1, 2)
        assert np.array_equal(pairs, want)


def test_drop_in_feature_matcher_matches_reference_wrapper(golden_dir):
    gold = np.load(os.path.join(golden_dir, "matching_icl_orb.npz"))
    kp1, kp2, d1, d2 = (gold["icl_" + k] for k in ("kp1", "kp2", "desc1", "desc2"))
    matches, p1, f1, p2, f2 = FeatureMatcher().match_features(kp1, d1, kp2, d2)
    assert [m[0].queryIdx for m in matches] == gold["icl_keep_q"].tolist()
    assert [m[0].trainIdx for m in matches] == gold["icl_keep_t"].tolist()
    assert np.array_equal(p1, gold["icl_pts1"]) and np.array_equal(p2, gold["icl_pts2"])
    assert np.array_equal(f1, d1[gold["icl_keep_q"]]) and np.array_equal(f2, d2[gold["icl_keep_t"]])
    rm, rp1, rf1, rp2, rf2 = om.match_features_cv2(kp1, d1, kp2, d2)
    assert [m[0].distance for m in matches] == [m[0].distance for m in rm]
    with pytest.raises(ValueError):
        FeatureMatcher().match_features(kp1, d1, kp2[:1], d2[:1])


def test_c2_full_size_batch_properties():
    """2048 x 2048 x 64 pairs: planted matches recovered, batch entries independent, cv2 parity
    on one pair, and idempotence across two launches."""
    B = 64
    q, t = synth.descriptors_batch(B, seed=1)
    g1 = run_gpu(q, t, 0.8, True)
    g2 = run_gpu(q, t, 0.8, True)
    for k in ("idx", "dist", "keep_q", "keep_t", "n_keep", "rev_key"):
        assert np.array_equal(g1[k], g2[k])
    for b in (0, B - 1):
        idx, dist = om.knn2_cv2(q[b], t[b])
        assert np.array_equal(g1["idx"][b], idx) and np.array_equal(g1["dist"][b], dist)
    one = run_gpu(q[5:6], t[5:6], 0.8, True)
    assert np.array_equal(one["idx"][0], g1["idx"][5])
    assert np.array_equal(one["keep_q"][0], g1["keep_q"][5])
    assert (g1["n_keep"] > 0.6 * 2048).all() and (g1["n_keep"] <= 0.7 * 2048 + 1).all()
    d = g1["dist"]
    assert (d[..., 0] <= d[..., 1]).all() and (d >= 0).all() and (d <= 256).all()


@pytest.mark.gpu
def test_ragged_batch_matches_oracle_per_pair():
    """vslam_hamming_knn2_ragged: pairs padded to a common stride with per-pair row counts (what the feature
    extractor hands over).  Every pair must equal the oracle on its real rows; padding rows are -1 and never win."""
    import torch
    from oracle import matching as om
    from visual_slam_b200 import match_batch, synth
    stride_q, stride_t = 1100, 1030
    counts = [(1100, 1030), (513, 1030), (512, 511), (1, 2), (700, 1), (0, 50), (40, 0), (1024, 1029)]
    B = len(counts)
    rng = np.random.default_rng(12)
    q = rng.integers(0, 256, (B, stride_q, 32), dtype=np.uint8)      # padding rows hold arbitrary bytes
    t = rng.integers(0, 256, (B, stride_t, 32), dtype=np.uint8)
    for b, (nq, nt) in enumerate(counts):
        qq, tt, _ = synth.descriptors_pair(max(nq, 1), max(nt, 1), seed=100 + b)
        q[b, :nq] = qq[:nq]
        t[b, :nt] = tt[:nt]
    nq_d = torch.tensor([c[0] for c in counts], dtype=torch.int32, device="cuda")
    nt_d = torch.tensor([c[1] for c in counts], dtype=torch.int32, device="cuda")
    for cross in (False, True):
        out = match_batch(torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda(), 0.8, cross, n_q=nq_d, n_t=nt_d)
        torch.cuda.synchronize()
        idx, dist = out["idx"].cpu().numpy(), out["dist"].cpu().numpy()
        for b, (nq, nt) in enumerate(counts):
            assert (idx[b, nq:] == -1).all() and (dist[b, nq:] == -1).all(), (b, cross)
            n = int(out["n_keep"][b])
            if nq == 0 or nt == 0:
                assert n == 0 and (idx[b, :nq] == -1).all()
                continue
            oi, od = om.knn2_numpy(q[b, :nq], t[b, :nt])
            assert np.array_equal(idx[b, :nq], oi) and np.array_equal(dist[b, :nq], od), (b, cross)
            rev_idx = om.rev_nearest_numpy(q[b, :nq], t[b, :nt])[0] if cross else None
            kq, kt = om.filter_numpy(oi, od, 0.8, rev_idx)
            assert n == len(kq), (b, cross)
            assert np.array_equal(out["keep_q"][b, :n].cpu().numpy(), kq), (b, cross)
            assert np.array_equal(out["keep_t"][b, :n].cpu().numpy(), kt), (b, cross)


@pytest.mark.parametrize("batch,chunks", [(7, 3), (5, 8), (4, 1), (9, 4)])
def test_host_matcher_equals_device_call_and_oracle(batch, chunks):
    """HostMatcher (the object behind bench.py's e2e figure: pinned host buffers -> H2D -> kNN-2 + ratio + cross-check
    -> D2H, chunks on two streams) on batches that do not divide into the chunk count, against match_batch on the
    same data and against the oracle pair by pair; a second run() reuses the buffers."""
    from visual_slam_b200.matching import HostMatcher
    nq, nt = 300, 257
    qs, ts = synth.descriptors_batch(batch, nq, nt, seed=31 + batch)
    hm = HostMatcher(batch, nq, nt, 0.8, True, chunks=chunks)
    assert hm.chunks == min(chunks, batch)
    for rep in range(2):
        hm.load(qs, ts)
        kq, kt, nk = hm.run()
        torch.cuda.synchronize()
        g = run_gpu(qs, ts, 0.8, True)
        assert np.array_equal(nk.numpy(), g["n_keep"]) and np.array_equal(kq.numpy(), g["keep_q"])
        assert np.array_equal(kt.numpy(), g["keep_t"])
        for b in range(batch):
            idx, dist = om.knn2_numpy(qs[b], ts[b])
            rev_idx, _ = om.rev_nearest_numpy(qs[b], ts[b])
            oq, ot = om.filter_numpy(idx, dist, 0.8, rev_idx)
            n = int(nk[b])
            assert n == len(oq) and np.array_equal(kq[b, :n].numpy(), oq) and np.array_equal(kt[b, :n].numpy(), ot)
        qs, ts = synth.descriptors_batch(batch, nq, nt, seed=77 + batch)              # other data, same buffers
    assert hm.h2d_bytes == batch * (nq + nt) * 32 and hm.d2h_bytes == 4 * batch * (2 * nq + 1)


def test_host_matcher_pipelined_calls_give_the_same_lists():
    """run(join=False) three times in a row (the bench's end-to-end loop: consecutive batches pipeline on the two
    internal streams), then join(): the pinned results are those of a plain run()."""
    from visual_slam_b200.matching import HostMatcher
    batch, nq, nt = 11, 300, 257
    qs, ts = synth.descriptors_batch(batch, nq, nt, seed=5)
    hm = HostMatcher(batch, nq, nt, 0.8, True, chunks=4)
    hm.load(qs, ts)
    kq, kt, nk = hm.run()
    torch.cuda.synchronize()
    ref = (kq.numpy().copy(), kt.numpy().copy(), nk.numpy().copy())
    kq.zero_(); kt.zero_(); nk.zero_()
    for _ in range(3):
        kq, kt, nk = hm.run(join=False)
    hm.join()
    torch.cuda.synchronize()
    assert np.array_equal(nk.numpy(), ref[2]) and np.array_equal(kq.numpy(), ref[0]) and np.array_equal(kt.numpy(), ref[1])

