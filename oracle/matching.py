"""Oracle for descriptor matching (TEST INFRASTRUCTURE -- never imported by the product).

Two independent statements of the same result:
  * `knn2_cv2` / `match_features_cv2`: the reference's own calls -- cv2.BFMatcher(NORM_HAMMING)
    .knnMatch(k=2) (reference src/v2/frame.py:18,23 with the ORB line frame.py:8) followed by
    the ratio loop frame.py:30-42.  OpenCV is a third-party wheel (version un-pinned by the
    reference; 4.13.0 in this image).
  * `knn2_numpy` / `filter_numpy`: a NumPy restatement (popcount table + stable sort) that needs
    no cv2.  Pinned against cv2 in tests/test_oracle_matching.py, ties included.
Cross-check semantics follow src/slam.py:40-47 (argmin both ways, first minimum wins).
"""
import numpy as np

_POP8 = np.array([bin(i).count("1") for i in range(256)], dtype=np.uint16)


def hamming_matrix(q, t):
    """All-pairs Hamming distances, uint16 [nq, nt] (what cv::batchDistance computes)."""
    q = np.ascontiguousarray(q, dtype=np.uint8)
    t = np.ascontiguousarray(t, dtype=np.uint8)
    out = np.zeros((q.shape[0], t.shape[0]), dtype=np.uint16)
    for b in range(q.shape[1]):
        out += _POP8[q[:, b][:, None] ^ t[:, b][None, :]]
    return out


def knn2_numpy(q, t):
    """Top-2 per query in ascending (distance, trainIdx) order; missing entries are -1."""
    nq, nt = q.shape[0], t.shape[0]
    idx = np.full((nq, 2), -1, dtype=np.int32)
    dist = np.full((nq, 2), -1, dtype=np.int32)
    if nq == 0 or nt == 0:
        return idx, dist
    D = hamming_matrix(q, t).astype(np.int64)
    order = np.argsort(D, axis=1, kind="stable")[:, :2]  # stable => lower trainIdx wins ties
    k = order.shape[1]
    idx[:, :k] = order
    dist[:, :k] = np.take_along_axis(D, order, axis=1)
    return idx, dist


def rev_nearest_numpy(q, t):
    """For every train row: (nearest query index, distance), first-min (src/slam.py:41)."""
    D = hamming_matrix(q, t).astype(np.int64)
    ids2 = np.argmin(D, axis=0)
    return ids2.astype(np.int32), D[ids2, np.arange(t.shape[0])].astype(np.int32)


def filter_numpy(idx, dist, ratio=0.8, rev_idx=None):
    """frame.py:33 ratio test in float64 on float32-exact integers; optional mutual-NN check.
    Returns kept (queryIdx, trainIdx) in query order."""
    d0 = dist[:, 0].astype(np.float32).astype(np.float64)
    d1 = dist[:, 1].astype(np.float32).astype(np.float64)
    keep = (idx[:, 1] >= 0) & (d0 < ratio * d1)
    if rev_idx is not None:
        qs = np.arange(idx.shape[0])
        safe_t = np.where(idx[:, 0] >= 0, idx[:, 0], 0)
        keep &= rev_idx[safe_t] == qs
    kq = np.nonzero(keep)[0].astype(np.int32)
    return kq, idx[kq, 0].astype(np.int32)


def knn2_cv2(q, t):
    import cv2
    raw = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(np.ascontiguousarray(q), np.ascontiguousarray(t), k=2)
    nq = q.shape[0]
    idx = np.full((nq, 2), -1, dtype=np.int32)
    dist = np.full((nq, 2), -1, dtype=np.int32)
    for i, row in enumerate(raw):
        for j, m in enumerate(row):
            assert m.queryIdx == i and m.imgIdx == 0
            idx[i, j] = m.trainIdx
            dist[i, j] = int(m.distance)
            assert float(int(m.distance)) == m.distance
    return idx, dist


def cross_check_cv2(q, t):
    """cv2.BFMatcher(NORM_HAMMING, crossCheck=True).knnMatch(k=1): mutual pairs (q, t)."""
    import cv2
    raw = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).knnMatch(
        np.ascontiguousarray(q), np.ascontiguousarray(t), k=1)
    return sorted((m.queryIdx, m.trainIdx) for row in raw for m in row)


def match_features_cv2(kp1, desc1, kp2, desc2, ratio=0.8):
    """Restates FeatureMatcher.match_features (frame.py:20-49) with the Hamming matcher."""
    import cv2
    raw = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(desc1, desc2, k=2)
    matches, pts1, pts2, ft1, ft2 = [], [], [], [], []
    for m, n in raw:
        if m.distance < ratio * n.distance:
            pts2.append(kp2[m.trainIdx]); pts1.append(kp1[m.queryIdx])
            ft1.append(desc1[m.queryIdx]); ft2.append(desc2[m.trainIdx])
            matches.append([m])
    return matches, np.asarray(pts1), np.asarray(ft1), np.asarray(pts2), np.asarray(ft2)
