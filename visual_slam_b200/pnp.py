"""RANSAC PnP: batched tensor entry + drop-in for cv2.solvePnPRansac as the reference calls it
(src/v2/main.py:196-197).  Arithmetic in csrc/pnp.cu through the C ABI."""
import ctypes as C

import numpy as np
import torch

from . import _cabi

DEFAULT_ITERS, DEFAULT_THR, DEFAULT_CONF = 100, 8.0, 0.99  # cv2.solvePnPRansac defaults


def _k(K):
    return (C.c_double * 4)(*[float(v) for v in K])


def pnp_hypotheses(X, uv, n_pts, words, K):
    B, n_max, H = X.shape[0], X.shape[1], words.shape[1]
    hyps = torch.empty((B, H, 12), dtype=torch.float64, device=X.device)
    valid = torch.empty((B, H), dtype=torch.uint8, device=X.device)
    with torch.cuda.device(X.device):
        _cabi.check(_cabi.lib().vslam_pnp_hypotheses(
            _cabi.ptr(X), _cabi.ptr(uv), _cabi.ptr(n_pts), n_max, B, _cabi.ptr(words), H, _k(K),
            _cabi.ptr(hyps), _cabi.ptr(valid), _cabi.stream_ptr()))
    return hyps, valid


def pnp_score(X, uv, n_pts, hyps, valid, K, thr=DEFAULT_THR):
    B, n_max, H = X.shape[0], X.shape[1], hyps.shape[1]
    counts = torch.empty((B, H), dtype=torch.int32, device=X.device)
    with torch.cuda.device(X.device):
        _cabi.check(_cabi.lib().vslam_pnp_score(
            _cabi.ptr(X), _cabi.ptr(uv), _cabi.ptr(n_pts), n_max, B, _cabi.ptr(hyps), _cabi.ptr(valid), H,
            _k(K), float(thr), _cabi.ptr(counts), _cabi.stream_ptr()))
    return counts


def pnp_select_refine(X, uv, n_pts, hyps, valid, counts, K, guess, thr=DEFAULT_THR, confidence=DEFAULT_CONF,
                      max_refine_iters=50):
    B, n_max, H = X.shape[0], X.shape[1], hyps.shape[1]
    dev = X.device
    best = torch.empty((B,), dtype=torch.int32, device=dev)
    bits = torch.empty((B, (n_max + 31) // 32), dtype=torch.int32, device=dev)
    n_inl = torch.empty((B,), dtype=torch.int32, device=dev)
    pose = torch.empty((B, 6), dtype=torch.float64, device=dev)
    status = torch.empty((B,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib().vslam_pnp_select_refine(
            _cabi.ptr(X), _cabi.ptr(uv), _cabi.ptr(n_pts), n_max, B, _cabi.ptr(hyps), _cabi.ptr(valid),
            _cabi.ptr(counts), H, _k(K), float(thr), float(confidence), _cabi.ptr(guess), int(max_refine_iters),
            _cabi.ptr(best), _cabi.ptr(bits), _cabi.ptr(n_inl), _cabi.ptr(pose), _cabi.ptr(status),
            _cabi.stream_ptr()))
    return {"best": best, "inlier_bits": bits, "n_inliers": n_inl, "pose": pose, "status": status}


def pnp_ransac_batch(X, uv, n_pts, words, K, guess, thr=DEFAULT_THR, confidence=DEFAULT_CONF):
    """X f32 [B,n_max,3], uv f32 [B,n_max,2], n_pts i32 [B], words int32/uint32 [B,H,4] (random
    bits), guess f64 [B,6] -- all CUDA tensors.  Returns a dict of device tensors."""
    if not torch.cuda.is_available():
        raise _cabi.VslamError("visual_slam_b200 needs a CUDA device (no CPU fallback)")
    X = X.contiguous(); uv = uv.contiguous(); guess = guess.contiguous(); words = words.contiguous()
    if X.dtype != torch.float32 or uv.dtype != torch.float32 or guess.dtype != torch.float64:
        raise ValueError("X/uv must be float32 and guess float64")
    hyps, valid = pnp_hypotheses(X, uv, n_pts, words, K)
    counts = pnp_score(X, uv, n_pts, hyps, valid, K, thr)
    out = pnp_select_refine(X, uv, n_pts, hyps, valid, counts, K, guess, thr, confidence)
    out.update(hyps=hyps, valid=valid, counts=counts)
    return out


def unpack_inliers(bits_row, n):
    """int32 bit words -> sorted inlier indices (host)."""
    w = np.asarray(bits_row).view(np.uint32)
    m = np.unpackbits(w.view(np.uint8), bitorder="little")[:n]
    return np.nonzero(m)[0].astype(np.int32)


def solvePnPRansac(objectPoints, imagePoints, cameraMatrix, distCoeffs=None, rvec=None, tvec=None,
                   useExtrinsicGuess=False, iterationsCount=DEFAULT_ITERS, reprojectionError=DEFAULT_THR,
                   confidence=DEFAULT_CONF, inliers=None, flags=0, seed=0, device="cuda:0"):
    """Drop-in for the call at reference src/v2/main.py:196-197.  Returns (retval, rvec (3,1) f64,
    tvec (3,1) f64, inliers (M,1) int32 or None) like OpenCV 4.13.  Distortion must be empty/zero,
    as in the reference.  Samples come from a PCG64 stream seeded with `seed` (deterministic, like
    cv2's internal fixed-seed RNG)."""
    if distCoeffs is not None and np.asarray(distCoeffs).size and np.any(np.asarray(distCoeffs) != 0):
        raise ValueError("solvePnPRansac: only zero distortion is supported (the reference passes [])")
    X = np.ascontiguousarray(np.asarray(objectPoints, dtype=np.float32).reshape(-1, 3))
    uv = np.ascontiguousarray(np.asarray(imagePoints, dtype=np.float32).reshape(-1, 2))
    if X.shape[0] != uv.shape[0]:
        raise ValueError("objectPoints / imagePoints disagree in length")
    Km = np.asarray(cameraMatrix, dtype=np.float64).reshape(3, 3)
    K = (Km[0, 0], Km[1, 1], Km[0, 2], Km[1, 2])
    n = X.shape[0]
    g = np.zeros(6)
    if useExtrinsicGuess and rvec is not None and tvec is not None:
        g[:3] = np.asarray(rvec, dtype=np.float64).reshape(3)
        g[3:] = np.asarray(tvec, dtype=np.float64).reshape(3)
    if n < 4:
        return False, g[:3].reshape(3, 1).copy(), g[3:].reshape(3, 1).copy(), None
    dev = torch.device(device)
    words = np.random.default_rng(seed).integers(0, 2 ** 32, size=(1, iterationsCount, 4), dtype=np.uint64)
    out = pnp_ransac_batch(torch.from_numpy(X)[None].to(dev), torch.from_numpy(uv)[None].to(dev),
                           torch.tensor([n], dtype=torch.int32, device=dev),
                           torch.from_numpy(words.astype(np.uint32).view(np.int32)).to(dev), K,
                           torch.from_numpy(g)[None].to(dev), reprojectionError, confidence)
    status = int(out["status"][0].item())
    pose = out["pose"][0].cpu().numpy()
    if status == 1:
        return False, g[:3].reshape(3, 1).copy(), g[3:].reshape(3, 1).copy(), None
    inl = unpack_inliers(out["inlier_bits"][0].cpu().numpy(), n)
    return True, pose[:3].reshape(3, 1).copy(), pose[3:].reshape(3, 1).copy(), inl.reshape(-1, 1)
