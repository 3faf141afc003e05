"""CPU: pins the PnP oracle against live OpenCV (the reference's third-party kernel)."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
from oracle import pnp as op  # noqa: E402
from visual_slam_b200 import synth  # noqa: E402

K = synth.ICL_K
KM = np.array([[K[0], 0, K[2]], [0, K[1], K[3]], [0, 0, 1.0]])


def test_rodrigues_matches_cv2():
    rng = np.random.default_rng(0)
    for _ in range(20):
        r = rng.normal(size=3) * rng.uniform(0.01, 3.0)
        R = op.rodrigues(r)
        assert np.allclose(R, cv2.Rodrigues(r)[0], atol=1e-14)
        assert np.allclose(op.rodrigues_inv(R), cv2.Rodrigues(R)[0].ravel(), atol=1e-10)


def test_projection_and_error_match_cv2_float_semantics():
    X, uv, rvec, tvec, _ = synth.tracking_pair(3, 500)
    Rt = np.concatenate([op.rodrigues(rvec), tvec[:, None]], axis=1)
    want, _ = cv2.projectPoints(X.reshape(-1, 1, 3), rvec, tvec, KM, None)
    got = op.project_f32(Rt, X, K)
    want32 = want.reshape(-1, 2).astype(np.float32)
    assert np.array_equal(got, want32)


def test_p3p_plus1_matches_cv2_p3p():
    ok_count = 0
    for seed in range(40):
        X, uv, rvec, tvec, inl = synth.tracking_pair(seed, 64, outlier_frac=0.0, pix_sigma=0.0)
        s = np.random.default_rng(seed).permutation(64)[:4]
        Rt = op.p3p_plus1(X[s], uv[s], K)
        ok, rv, tv = cv2.solvePnP(X[s].astype(np.float64), uv[s].astype(np.float64), KM, None, flags=cv2.SOLVEPNP_P3P)
        if not ok or Rt is None:
            continue
        # noise-free data: both must recover the true pose
        assert np.allclose(Rt[:, :3], op.rodrigues(rvec), atol=1e-5)
        assert np.allclose(Rt[:, 3], tvec, atol=1e-5)
        assert np.allclose(Rt[:, :3], cv2.Rodrigues(rv)[0], atol=1e-5) and np.allclose(Rt[:, 3], tv.ravel(), atol=1e-5)
        ok_count += 1
    assert ok_count >= 30


def test_ransac_update_num_iters_matches_known_values():
    assert op.ransac_update_num_iters(0.99, 0.3, 4, 100) == round(np.log(0.01) / np.log(1 - 0.7 ** 4))
    assert op.ransac_update_num_iters(0.99, 0.95, 4, 100) == 100
    assert op.ransac_update_num_iters(0.99, 0.0, 4, 100) == 0


def test_sampling_rule_has_no_duplicates():
    w = np.random.default_rng(1).integers(0, 2 ** 32, size=(200, 4), dtype=np.uint64)
    s = op.draw_samples(w, 7)
    assert all(len(set(r)) == 4 for r in s.tolist()) and s.max() < 7 and s.min() >= 0


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_full_path_agrees_with_cv2_solvepnpransac_on_well_separated_data(seed):
    X, uv, rvec, tvec, inl = synth.tracking_pair(seed, 600, outlier_frac=0.3)
    # keep outliers well away from the 8 px gate so cv2's own sampling cannot change the inlier set
    Rt = np.concatenate([op.rodrigues(rvec), tvec[:, None]], axis=1)
    err = np.sqrt(op.reproj_err_f32(Rt, X, uv, K))
    keep = (err < 3.0) | (err > 30.0)
    X, uv = X[keep], uv[keep]
    words = np.random.default_rng(seed + 100).integers(0, 2 ** 32, size=(100, 4), dtype=np.uint64)
    samples = op.draw_samples(words, len(X))
    guess_r, guess_t = np.zeros(3), np.zeros(3)
    got = op.solve_pnp_ransac(X, uv, K, samples, guess_r, guess_t)
    ok, rv, tv, inliers = cv2.solvePnPRansac(objectPoints=X[:, None, :], imagePoints=uv[:, None, :],
                                             cameraMatrix=KM, distCoeffs=np.array([]), rvec=guess_r.reshape(3, 1).copy(),
                                             tvec=guess_t.reshape(3, 1).copy(), useExtrinsicGuess=True)
    assert ok and got["ok"]
    assert np.array_equal(got["inliers"], inliers.ravel())
    assert np.abs(got["rvec"] - rv.ravel()).max() < 1e-6 and np.abs(got["tvec"] - tv.ravel()).max() < 1e-6
    r2, t2 = op.refine_numpy(X, uv, K, got["inliers"], guess_r, guess_t)
    assert np.abs(r2 - got["rvec"]).max() < 1e-6 and np.abs(t2 - got["tvec"]).max() < 1e-6
