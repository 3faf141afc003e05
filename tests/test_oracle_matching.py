"""CPU: pins the matching oracle against the golden vectors produced by the reference's own
FeatureMatcher (tests/golden/make_matching_golden.py) and against live cv2."""
import os

import numpy as np
import pytest

from oracle import matching as om
from visual_slam_b200 import synth


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "matching_icl_orb.npz"))


@pytest.mark.parametrize("case,ratio", [("icl", 0.8), ("tie", 0.95)])
def test_numpy_restatement_matches_reference_golden(gold, case, ratio):
    q, t = gold[case + "_desc1"], gold[case + "_desc2"]
    idx, dist = om.knn2_numpy(q, t)
    assert np.array_equal(idx, gold[case + "_idx"])
    assert np.array_equal(dist, gold[case + "_dist"])
    kq, kt = om.filter_numpy(idx, dist, ratio)
    assert np.array_equal(kq, gold[case + "_keep_q"])
    assert np.array_equal(kt, gold[case + "_keep_t"])
    rev_idx, _ = om.rev_nearest_numpy(q, t)
    mutual = [(int(a), int(idx[a, 0])) for a in range(len(q)) if rev_idx[idx[a, 0]] == a]
    assert mutual == [tuple(r) for r in gold[case + "_mutual"]]


@pytest.mark.parametrize("nq,nt", [(1, 2), (2, 1), (31, 33), (33, 31), (257, 129)])
def test_numpy_vs_cv2_small_and_ragged(nq, nt):
    cv2 = pytest.importorskip("cv2")  # noqa: F841
    for gen in (synth.low_entropy_pair, lambda a, b, seed: synth.descriptors_pair(a, b, seed)[:2]):
        q, t = gen(nq, nt, seed=nq * 100 + nt)
        i1, d1 = om.knn2_numpy(q, t)
        i2, d2 = om.knn2_cv2(q, t)
        assert np.array_equal(i1, i2) and np.array_equal(d1, d2)
        if nq and nt:
            rev_idx, _ = om.rev_nearest_numpy(q, t)
            mutual = sorted((a, int(i1[a, 0])) for a in range(nq) if rev_idx[i1[a, 0]] == a)
            assert mutual == om.cross_check_cv2(q, t)


def test_duplicate_train_rows_lower_index_wins():
    q, t, _ = synth.descriptors_pair(64, 64, seed=3)
    t[10] = t[5]
    q[0] = t[5]
    idx, dist = om.knn2_numpy(q, t)
    assert idx[0, 0] == 5 and idx[0, 1] == 10 and dist[0, 0] == 0 and dist[0, 1] == 0


def test_c2_generator_ratio_survivors():
    q, t, truth = synth.descriptors_pair(512, 512, seed=0)
    idx, dist = om.knn2_numpy(q, t)
    kq, kt = om.filter_numpy(idx, dist, 0.8)
    planted = truth >= 0
    assert planted.sum() == round(0.7 * 512)
    assert np.array_equal(idx[planted, 0], truth[planted])
    assert set(kq.tolist()) == set(np.nonzero(planted)[0].tolist())
