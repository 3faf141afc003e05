"""Two-view geometry (map initialisation): batched tensor entries + drop-ins for the reference's
estimateEssential / estimateRelativePose (src/v2/helper_functions.py:47-70, :164-195; callers
src/v2/main.py:103,111).  Arithmetic in csrc/essential.cu through the C ABI; no CPU fallback."""
import ctypes as C

import numpy as np
import torch

from . import _cabi

DEFAULT_ITERS, DEFAULT_CONF = 1000, 0.999   # cv2.findEssentialMat: maxIters default, prob the reference passes
MAX_MODELS = 10


def _k(K):
    return (C.c_double * 4)(*[float(v) for v in K])


def _k4(K):
    """np.matrix / 3x3 array / 4-tuple -> (fx, fy, cx, cy)."""
    Km = np.asarray(K, dtype=np.float64)
    if Km.size == 4:
        return tuple(float(v) for v in Km.ravel())
    Km = Km.reshape(3, 3)
    return (Km[0, 0], Km[1, 1], Km[0, 2], Km[1, 2])


def _need_cuda():
    if not torch.cuda.is_available():
        raise _cabi.VslamError("visual_slam_b200 needs a CUDA device (no CPU fallback)")


def undistort_points(uv, K):
    """cv2.undistortPoints without distortion: uv f32 [..., 2] CUDA tensor -> same shape."""
    _need_cuda()
    uv = uv.contiguous()
    if uv.dtype != torch.float32 or uv.shape[-1] != 2:
        raise ValueError("uv must be float32 [..., 2]")
    out = torch.empty_like(uv)
    with torch.cuda.device(uv.device):
        _cabi.check(_cabi.lib().vslam_undistort_points(_cabi.ptr(uv), uv.numel() // 2, _k(K), _cabi.ptr(out),
                                                       _cabi.stream_ptr()))
    return out


def essential_hypotheses(x1, x2, n_pts, words):
    B, n_max, H = x1.shape[0], x1.shape[1], words.shape[1]
    models = torch.empty((B, H, MAX_MODELS, 9), dtype=torch.float64, device=x1.device)
    n_models = torch.empty((B, H), dtype=torch.int32, device=x1.device)
    with torch.cuda.device(x1.device):
        _cabi.check(_cabi.lib().vslam_essential_hypotheses(
            _cabi.ptr(x1), _cabi.ptr(x2), _cabi.ptr(n_pts), n_max, B, _cabi.ptr(words), H, _cabi.ptr(models),
            _cabi.ptr(n_models), _cabi.stream_ptr()))
    return models, n_models


def essential_score(x1, x2, n_pts, models, n_models, thr):
    B, n_max, H = x1.shape[0], x1.shape[1], models.shape[1]
    counts = torch.empty((B, H, MAX_MODELS), dtype=torch.int32, device=x1.device)
    with torch.cuda.device(x1.device):
        _cabi.check(_cabi.lib().vslam_essential_score(
            _cabi.ptr(x1), _cabi.ptr(x2), _cabi.ptr(n_pts), n_max, B, _cabi.ptr(models), _cabi.ptr(n_models), H,
            float(thr), _cabi.ptr(counts), _cabi.stream_ptr()))
    return counts


def essential_select(x1, x2, n_pts, models, n_models, counts, thr, confidence=DEFAULT_CONF):
    B, n_max, H = x1.shape[0], x1.shape[1], models.shape[1]
    dev = x1.device
    best = torch.empty((B,), dtype=torch.int32, device=dev)
    E = torch.empty((B, 3, 3), dtype=torch.float64, device=dev)
    mask = torch.empty((B, n_max), dtype=torch.uint8, device=dev)
    n_inl = torch.empty((B,), dtype=torch.int32, device=dev)
    score = torch.empty((B,), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib().vslam_essential_select(
            _cabi.ptr(x1), _cabi.ptr(x2), _cabi.ptr(n_pts), n_max, B, _cabi.ptr(models), _cabi.ptr(n_models),
            _cabi.ptr(counts), H, float(thr), float(confidence), _cabi.ptr(best), _cabi.ptr(E), _cabi.ptr(mask),
            _cabi.ptr(n_inl), _cabi.ptr(score), _cabi.stream_ptr()))
    return {"best": best, "E": E, "mask": mask, "n_inliers": n_inl, "score": score}


def essential_ransac_batch(x1, x2, n_pts, words, thr, confidence=DEFAULT_CONF):
    """x1, x2 f32 [B,n_max,2] normalised coordinates, n_pts i32 [B], words int32/uint32 [B,H,5]
    (random bits) -- CUDA tensors.  Returns a dict of device tensors (E, mask, n_inliers, score, best,
    models, n_models, counts)."""
    _need_cuda()
    x1 = x1.contiguous(); x2 = x2.contiguous(); words = words.contiguous()
    if x1.dtype != torch.float32 or x2.dtype != torch.float32 or x1.shape != x2.shape or x1.shape[-1] != 2:
        raise ValueError("x1/x2 must be float32 [B, n_max, 2] of equal shape")
    if words.shape[-1] != 5:
        raise ValueError("words must be [B, H, 5]")
    models, n_models = essential_hypotheses(x1, x2, n_pts, words)
    counts = essential_score(x1, x2, n_pts, models, n_models, thr)
    out = essential_select(x1, x2, n_pts, models, n_models, counts, thr, confidence)
    out.update(models=models, n_models=n_models, counts=counts)
    return out


def essential_budget(n_pts, n_max, n_models, counts, max_iters, confidence=DEFAULT_CONF):
    """Iteration budget left after the first H = n_models.shape[1] hypotheses of a max_iters run: i32 [B] (device)."""
    B, H = n_models.shape[0], n_models.shape[1]
    out = torch.empty((B,), dtype=torch.int32, device=n_models.device)
    with torch.cuda.device(n_models.device):
        _cabi.check(_cabi.lib().vslam_essential_budget(
            _cabi.ptr(n_pts), int(n_max), B, _cabi.ptr(n_models.contiguous()), _cabi.ptr(counts.contiguous()), H,
            int(max_iters), float(confidence), _cabi.ptr(out), _cabi.stream_ptr()))
    return out


def essential_ransac_single(x1, x2, n, words, thr, confidence=DEFAULT_CONF, rounds=(48, 192)):
    """One pair (x1, x2 f32 [1,n,2]; words [1,H,5]), hypotheses generated in rounds: cv2's loop shrinks its budget as
    soon as a well-supported model appears (typically after a few dozen of the 1000 iterations), so the later
    samples are never drawn there; here they are never solved either.  Same result as essential_ransac_batch."""
    dev, H = x1.device, words.shape[1]
    n_pts = torch.tensor([n], dtype=torch.int32, device=dev)
    models = torch.empty((1, H, MAX_MODELS, 9), dtype=torch.float64, device=dev)
    n_models = torch.empty((1, H), dtype=torch.int32, device=dev)
    counts = torch.empty((1, H, MAX_MODELS), dtype=torch.int32, device=dev)
    lo = 0
    for hi in [r for r in rounds if r < H] + [H]:
        m, nm = essential_hypotheses(x1, x2, n_pts, words[:, lo:hi].contiguous())
        models[:, lo:hi], n_models[:, lo:hi] = m, nm
        counts[:, lo:hi] = essential_score(x1, x2, n_pts, m, nm, thr)
        lo = hi
        if hi < H and int(essential_budget(n_pts, x1.shape[1], n_models[:, :hi], counts[:, :hi], H, confidence)[0]) <= hi:
            break
    out = essential_select(x1, x2, n_pts, models[:, :lo].contiguous(), n_models[:, :lo].contiguous(),
                           counts[:, :lo].contiguous(), thr, confidence)
    out.update(models=models[:, :lo], n_models=n_models[:, :lo], counts=counts[:, :lo], hypotheses_solved=lo)
    return out


def recover_pose_batch(E, p1, p2, n_pts, K, distance_thresh=50.0):
    """E f64 [B,3,3], p1/p2 f32 [B,n_max,2] pixels, n_pts i32 [B] -- CUDA tensors.
    Returns dict: R [B,3,3], t [B,3], mask u8 [B,n_max] (0/255), tri f64 [B,4,n_max], good i32 [B,5]."""
    _need_cuda()
    E = E.contiguous(); p1 = p1.contiguous(); p2 = p2.contiguous()
    if E.dtype != torch.float64 or p1.dtype != torch.float32 or p2.dtype != torch.float32:
        raise ValueError("E must be float64, p1/p2 float32")
    B, n_max = p1.shape[0], p1.shape[1]
    dev = p1.device
    R = torch.empty((B, 3, 3), dtype=torch.float64, device=dev)
    t = torch.empty((B, 3), dtype=torch.float64, device=dev)
    mask = torch.empty((B, n_max), dtype=torch.uint8, device=dev)
    tri = torch.empty((B, 4, n_max), dtype=torch.float64, device=dev)
    good = torch.empty((B, 5), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib().vslam_recover_pose(
            _cabi.ptr(E), _cabi.ptr(p1), _cabi.ptr(p2), _cabi.ptr(n_pts), n_max, B, _k(K), float(distance_thresh),
            _cabi.ptr(R), _cabi.ptr(t), _cabi.ptr(mask), _cabi.ptr(tri), _cabi.ptr(good), _cabi.stream_ptr()))
    return {"R": R, "t": t, "mask": mask, "tri": tri, "good": good}


# ---- drop-ins with the reference's signatures ------------------------------------------------------
def estimateEssential(pts1, pts2, K, essTh, *, maxIters=DEFAULT_ITERS, prob=DEFAULT_CONF, seed=0,
                      device="cuda:0"):
    """Drop-in for reference src/v2/helper_functions.py:47-70.  pts (N,2) float32 pixels, K 3x3.
    Returns (E (3,3) f64 with |E|_F = 1, inliers (N,1) uint8 in {0,1}, score float).  Samples come
    from a PCG64 stream seeded with `seed` (deterministic, like cv2's internal fixed-seed RNG).
    Fewer than 5 points, or no model with >= 5 inliers: (None, None, 0.0) like cv2 returning None."""
    p1 = np.ascontiguousarray(np.asarray(pts1, dtype=np.float32).reshape(-1, 2))
    p2 = np.ascontiguousarray(np.asarray(pts2, dtype=np.float32).reshape(-1, 2))
    if p1.shape != p2.shape:
        raise ValueError("pts1 / pts2 disagree in shape")
    n = p1.shape[0]
    if n < 5:
        return None, None, 0.0
    dev = torch.device(device)
    k4 = _k4(K)
    both = undistort_points(torch.from_numpy(np.stack([p1, p2])).to(dev), k4)
    words = np.random.default_rng(seed).integers(0, 2 ** 32, size=(1, int(maxIters), 5), dtype=np.uint64)
    out = essential_ransac_single(both[0:1], both[1:2], n, torch.from_numpy(words.astype(np.uint32).view(np.int32)).to(dev),
                                  float(essTh), float(prob))
    if int(out["best"][0].item()) < 0:
        return None, None, 0.0
    E = out["E"][0].cpu().numpy()
    inliers = out["mask"][0].cpu().numpy().reshape(-1, 1).astype(np.uint8)
    return E, inliers, float(out["score"][0].item())


def estimateRelativePose(tform, inlier_pts1, inlier_pts2, inlier_fts1, inlier_fts2, K, tform_type="Essential",
                         device="cuda:0"):
    """Drop-in for reference src/v2/helper_functions.py:164-195 ("Essential" branch, the only one the
    v2 driver uses).  Returns (R (3,3), t (3,1), validFraction, triangulatedPoints_good (4,M) f64 in the
    first camera's frame, and the four lists of rows kept by recoverPose's mask)."""
    if tform_type != "Essential":
        raise NotImplementedError("only the Essential branch is on the src/v2 hot path (main.py:111)")
    p1 = np.ascontiguousarray(np.asarray(inlier_pts1, dtype=np.float32).reshape(-1, 2))
    p2 = np.ascontiguousarray(np.asarray(inlier_pts2, dtype=np.float32).reshape(-1, 2))
    n = p1.shape[0]
    dev = torch.device(device)
    E = torch.from_numpy(np.asarray(tform, dtype=np.float64).reshape(1, 3, 3).copy()).to(dev)
    out = recover_pose_batch(E, torch.from_numpy(p1)[None].to(dev), torch.from_numpy(p2)[None].to(dev),
                             torch.tensor([n], dtype=torch.int32, device=dev), _k4(K), 50.0)
    R = out["R"][0].cpu().numpy()
    t = out["t"][0].cpu().numpy().reshape(3, 1)
    mask = out["mask"][0].cpu().numpy() == 255
    tri = out["tri"][0].cpu().numpy()
    keep = np.nonzero(mask)[0]
    valid_fraction = len(keep) / n if n else 0.0
    tri_good = tri[:, keep]
    return (R, t, valid_fraction, tri_good, [inlier_pts1[i] for i in keep], [inlier_pts2[i] for i in keep],
            [inlier_fts1[i] for i in keep], [inlier_fts2[i] for i in keep])
