"""Pose plumbing of the tracking loop (reference src/v2/main.py:191-204, src/v2/helper_functions.py:269-278):
batched device entries and drop-ins with the reference's names.  Arithmetic in csrc/poses.cu through the C ABI;
no CPU fallback."""
import numpy as np
import torch

from . import _cabi


def _f64(t, shape_tail, what):
    if t.dtype != torch.float64 or not t.is_cuda or tuple(t.shape[1:]) != shape_tail:
        raise ValueError("%s must be a float64 CUDA tensor of shape (batch,%s)" % (what, ",".join(map(str, shape_tail))))
    return t.contiguous()


def rodrigues_batch(rvec):
    """rvec f64 [B,3] -> R f64 [B,3,3] (cv2.Rodrigues, vector -> matrix)."""
    rvec = _f64(rvec, (3,), "rvec")
    R = torch.empty((rvec.shape[0], 3, 3), dtype=torch.float64, device=rvec.device)
    with torch.cuda.device(rvec.device):
        _cabi.check(_cabi.lib().vslam_rodrigues(_cabi.ptr(rvec), _cabi.ptr(R), rvec.shape[0], _cabi.stream_ptr()))
    return R


def rodrigues_inverse_batch(R):
    """R f64 [B,3,3] -> rvec f64 [B,3] (cv2.Rodrigues, matrix -> vector)."""
    R = _f64(R, (3, 3), "R")
    rvec = torch.empty((R.shape[0], 3), dtype=torch.float64, device=R.device)
    with torch.cuda.device(R.device):
        _cabi.check(_cabi.lib().vslam_rodrigues_inverse(_cabi.ptr(R), _cabi.ptr(rvec), R.shape[0],
                                                        _cabi.stream_ptr()))
    return rvec


def camera_to_world_batch(pose6, want_w2c=False, want_cam7=True):
    """pose6 f64 [B,6] (rvec, tvec; world -> camera, what PnP returns) -> dict with c2w f64 [B,4,4] (camera -> world,
    main.py:201-203), optionally w2c [B,4,4] (transformMatrix) and cam7 [B,7] (qx,qy,qz,qw,tx,ty,tz of c2w: the
    camera state of the BA arrays)."""
    pose6 = _f64(pose6, (6,), "pose6")
    B, dev = pose6.shape[0], pose6.device
    c2w = torch.empty((B, 4, 4), dtype=torch.float64, device=dev)
    w2c = torch.empty((B, 4, 4), dtype=torch.float64, device=dev) if want_w2c else None
    cam7 = torch.empty((B, 7), dtype=torch.float64, device=dev) if want_cam7 else None
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib().vslam_pose_camera_to_world(_cabi.ptr(pose6), _cabi.ptr(w2c), _cabi.ptr(c2w),
                                                           _cabi.ptr(cam7), B, _cabi.stream_ptr()))
    return {"c2w": c2w, "w2c": w2c, "cam7": cam7}


def guess_batch(c2w):
    """c2w f64 [B,4,4] (previous frame's camera -> world pose) -> pose6 f64 [B,6], the extrinsic guess of
    main.py:191-194."""
    c2w = _f64(c2w, (4, 4), "c2w")
    out = torch.empty((c2w.shape[0], 6), dtype=torch.float64, device=c2w.device)
    with torch.cuda.device(c2w.device):
        _cabi.check(_cabi.lib().vslam_pose_guess(_cabi.ptr(c2w), _cabi.ptr(out), c2w.shape[0], _cabi.stream_ptr()))
    return out


def guess_from_cam7_batch(cam7):
    """cam7 f64 [B,7] (qx,qy,qz,qw,tx,ty,tz of the previous frame's camera -> world pose) -> pose6 f64 [B,6]."""
    cam7 = _f64(cam7, (7,), "cam7")
    out = torch.empty((cam7.shape[0], 6), dtype=torch.float64, device=cam7.device)
    with torch.cuda.device(cam7.device):
        _cabi.check(_cabi.lib().vslam_pose_guess_cam7(_cabi.ptr(cam7), _cabi.ptr(out), cam7.shape[0],
                                                      _cabi.stream_ptr()))
    return out


# ---- drop-ins with the reference's names and return objects ------------------------------------------
def _dev(a, shape, device):
    if not torch.cuda.is_available():
        raise _cabi.VslamError("visual_slam_b200 needs a CUDA device (no CPU fallback)")
    return torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(shape))).to(device)


def Rtorvec(R, device="cuda:0"):
    """helper_functions.py:276-278: rotation matrix -> rotation vector, (3, 1) float64 like cv2.Rodrigues."""
    return rodrigues_inverse_batch(_dev(R, (1, 3, 3), device))[0].cpu().numpy().reshape(3, 1)


def transformMatrix(rvec, tvec, device="cuda:0"):
    """helper_functions.py:269-274: np.matrix 4x4 [Rodrigues(rvec) | tvec; 0 0 0 1].  tvec may be (3,), (3, 1) or,
    as main.py:199 produces on OpenCV >= 4.5, (3, 1, 1): like np.matrix it is squeezed to three numbers."""
    pose6 = np.concatenate([np.asarray(rvec, dtype=np.float64).reshape(3), np.asarray(tvec, dtype=np.float64).reshape(3)])
    out = camera_to_world_batch(_dev(pose6, (1, 6), device), want_w2c=True, want_cam7=False)
    return np.matrix(out["w2c"][0].cpu().numpy())


def camera_to_world(rvec, tvec, device="cuda:0"):
    """main.py:199-203 in one call: the camera -> world 4x4 of a PnP result."""
    pose6 = np.concatenate([np.asarray(rvec, dtype=np.float64).reshape(3), np.asarray(tvec, dtype=np.float64).reshape(3)])
    return camera_to_world_batch(_dev(pose6, (1, 6), device), want_cam7=False)["c2w"][0].cpu().numpy()


def pose_guess(W_T_prev, device="cuda:0"):
    """main.py:191-194: (rvec_guess (3, 1), tvec_guess (3,)) from the previous frame's camera -> world pose."""
    g = guess_batch(_dev(np.asarray(W_T_prev, dtype=np.float64), (1, 4, 4), device))[0].cpu().numpy()
    return g[:3].reshape(3, 1).copy(), g[3:].copy()
