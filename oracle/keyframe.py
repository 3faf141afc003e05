"""CPU restatement (TEST INFRASTRUCTURE ONLY -- never imported by the product path) of the reference's
key-frame control logic.  Pinned against the reference's own `GetListDiff` through
tests/golden/listdiff_ref.npz (tests/golden/make_listdiff_golden.py)."""
import numpy as np


def get_list_diff(kp1, kp2):
    """src/v2/helper_functions.py:316-326.  The value returned is the list comprehension at :326,
    `x not in kp2` on numpy rows = not (kp2 == x).any(): a row of kp1 is dropped when ANY row of kp2 has
    the same u OR the same v.  (The nested loop at :319-325 computes the row-wise test and discards it.)"""
    kp1 = np.asarray(kp1, dtype=np.float64).reshape(-1, 2)
    kp2 = np.asarray(kp2, dtype=np.float64).reshape(-1, 2)
    out = []
    for i in range(len(kp1)):
        hit = False
        for j in range(len(kp2)):
            if kp1[i, 0] == kp2[j, 0] or kp1[i, 1] == kp2[j, 1]:
                hit = True
                break
        if not hit:
            out.append(i)
    return out


def get_list_diff_rows(kp1, kp2):
    """The row-wise test of the dead loop (:319-325): both coordinates equal."""
    kp1 = np.asarray(kp1, dtype=np.float64).reshape(-1, 2)
    kp2 = np.asarray(kp2, dtype=np.float64).reshape(-1, 2)
    seen = {(float(a), float(b)) for a, b in kp2}
    return [i for i in range(len(kp1)) if (float(kp1[i, 0]), float(kp1[i, 1])) not in seen]


def is_keyframe(i, loop_idx, n_tracked, n_known):
    """src/v2/main.py:221."""
    return (i - loop_idx > 20 or n_tracked < 80) and (n_tracked < 0.9 * n_known)


def cull_due(id_frame):
    """src/v2/main.py:234."""
    return id_frame >= 6 and id_frame % 4 == 0


def discard_outlier_map_points(point_ids, obs_point_rows, n_visible_frames=3):
    """src/v2/map.py:124-131: keep the points seen in at least n_visible_frames frames (order kept).
    point_ids[r] is the id of row r, obs_point_rows the point row of every observation."""
    cnt = np.bincount(np.asarray(obs_point_rows, dtype=np.int64), minlength=len(point_ids))
    return [pid for r, pid in enumerate(point_ids) if cnt[r] >= n_visible_frames]
