"""Two-view triangulation on the device: batched tensor entry + drop-in for the
reference's `triangulate` (src/v2/helper_functions.py:281-291; caller src/v2/main.py:284) and the
cheirality/depth filter that follows it (main.py:285-306).  Arithmetic in csrc/triangulate.cu through
the C ABI; no CPU fallback."""
import numpy as np
import torch

from . import _cabi


def triangulate_batch(proj, x1, x2, n_pts, extr=None, z_max=1.0):
    """proj f64 [B,2,3,4], x1/x2 f64 [B,n_max,2], n_pts i32 [B] (CUDA tensors); extr f64 [B,2,3,4]
    world->camera matrices for the depth test or None.
    Returns (Xh [B,n_max,4], X3 [B,n_max,3], depth [B,n_max,2] | None, good u8 [B,n_max] | None)."""
    if not torch.cuda.is_available():
        raise _cabi.VslamError("visual_slam_b200 needs a CUDA device (no CPU fallback)")
    if proj.dtype != torch.float64 or x1.dtype != torch.float64 or x2.dtype != torch.float64:
        raise ValueError("proj, x1, x2 must be float64")
    if proj.shape[1:] != (2, 3, 4) or x1.shape != x2.shape or x1.shape[-1] != 2 or x1.shape[0] != proj.shape[0]:
        raise ValueError("shapes: proj [B,2,3,4], x1/x2 [B,n_max,2]")
    proj, x1, x2, n_pts = proj.contiguous(), x1.contiguous(), x2.contiguous(), n_pts.contiguous()
    B, n_max = x1.shape[0], x1.shape[1]
    dev = x1.device
    Xh = torch.zeros((B, n_max, 4), dtype=torch.float64, device=dev)
    X3 = torch.zeros((B, n_max, 3), dtype=torch.float64, device=dev)
    depth = good = None
    if extr is not None:
        extr = extr.contiguous()
        depth = torch.zeros((B, n_max, 2), dtype=torch.float64, device=dev)
        good = torch.zeros((B, n_max), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib().vslam_triangulate(
            _cabi.ptr(proj), _cabi.ptr(extr) if extr is not None else None, _cabi.ptr(x1), _cabi.ptr(x2),
            _cabi.ptr(n_pts), n_max, B, float(z_max), _cabi.ptr(Xh), _cabi.ptr(X3),
            _cabi.ptr(depth) if depth is not None else None, _cabi.ptr(good) if good is not None else None,
            _cabi.stream_ptr()))
    return Xh, X3, depth, good


def _one(pose1, pose2, pts1, pts2, extr, z_max, device):
    P = np.stack([np.asarray(pose1, dtype=np.float64)[:3], np.asarray(pose2, dtype=np.float64)[:3]])
    if P.shape != (2, 3, 4):
        raise ValueError("pose1/pose2 must be 3x4 projection matrices")
    p1 = np.asarray(pts1, dtype=np.float64)
    p2 = np.asarray(pts2, dtype=np.float64)
    if p1.ndim != 2 or p1.shape[1] < 2 or p1.shape[0] != p2.shape[0]:
        raise ValueError("pts1/pts2 must be (N, >=2) with equal N")
    n = p1.shape[0]
    dev = torch.device(device)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    E = None
    if extr is not None:
        E = t(np.stack([np.asarray(extr[0], dtype=np.float64)[:3], np.asarray(extr[1], dtype=np.float64)[:3]])[None])
    return triangulate_batch(t(P[None]), t(p1[None, :, :2]), t(p2[None, :, :2]),
                             torch.tensor([n], dtype=torch.int32, device=dev), E, z_max)


def triangulate(pose1, pose2, pts1, pts2, device="cuda:0"):
    """Drop-in for helper_functions.py:281 (same argument names and meaning): returns (N, 4) float64
    unit homogeneous points.  The sign of each row is fixed (largest component positive) where the
    reference's is whatever LAPACK returns; main.py:285 divides by w, which removes it."""
    Xh, _, _, _ = _one(pose1, pose2, pts1, pts2, None, 1.0, device)
    return Xh[0].cpu().numpy()


def triangulate_and_filter(Proj1, Proj2, p1, p2, pts1, pts2, z_max=1.0, device="cuda:0"):
    """main.py:284-306 in one call: triangulate, divide by w, keep points whose depth in both cameras
    (world->camera poses p1, p2) lies in (0, z_max).  Returns (X3 (N,3), good_points_idx)."""
    _, X3, _, good = _one(Proj1, Proj2, pts1, pts2, (p1, p2), z_max, device)
    return X3[0].cpu().numpy(), np.nonzero(good[0].cpu().numpy())[0]
