"""Feature extraction on the device, same names and results as the reference.

* FeatureExtractor.compute_features   src/v2/frame.py:5-14 in its ORB build variant (frame.py:8):
      pts = cv2.goodFeaturesToTrack(np.mean(img, axis=2).astype(np.uint8), 3000, qualityLevel=0.01, minDistance=7)
      kps = [cv2.KeyPoint(x, y, size=20) for ...];  kp, des = cv2.ORB_create().compute(img, kps)
      return cv2.KeyPoint_convert(kp), des
  -> bit-exact key points (order included) and descriptors against OpenCV 4.13.
* extract_features_batch              the same for a batch of equally sized frames resident on the device.

No CPU fallback: without the CUDA library or a device every entry point raises.
"""
import math

import numpy as np
import torch

from . import _cabi
from .orb_pattern import ORB_PATTERN

_pattern_cache = {}


def _pattern_on(device):
    key = str(device)
    if key not in _pattern_cache:
        _pattern_cache[key] = torch.from_numpy(ORB_PATTERN.copy()).to(device)
    return _pattern_cache[key]


def _angle_terms(angle_deg):
    """computeOrbDescriptors: angle (float32 degrees) * float32(pi/180); a = float(cos), b = float(sin)."""
    ang = np.float32(angle_deg) * np.float32(np.pi / np.float32(180.0))
    return float(np.float32(math.cos(float(ang)))), float(np.float32(math.sin(float(ang))))


def extract_features_batch(imgs, max_corners=3000, quality=0.01, min_distance=7.0, angle=-1.0):
    """imgs: uint8 CUDA tensor (B, H, W, 3), BGR.  Returns a dict of CUDA tensors:
    kp (B, max_corners, 2) f32 + n_kp (B,) i32 + desc (B, max_corners, 32) u8 (rows >= n_kp are undefined),
    corners / n_corners (goodFeaturesToTrack before ORB's border filter), eig (B, H, W) f32, blur (B, H, W) u8."""
    if imgs.dtype != torch.uint8 or not imgs.is_cuda or imgs.dim() != 4 or imgs.shape[3] != 3:
        raise ValueError("imgs must be a uint8 CUDA tensor of shape (B, H, W, 3)")
    if min_distance >= 1 and float(min_distance) != int(min_distance):
        raise ValueError("min_distance must be an integer value (or < 1)")
    if max_corners < 1 or not quality > 0:
        raise ValueError("max_corners must be >= 1 and quality > 0")
    imgs = imgs.contiguous()
    B, H, W, _ = imgs.shape
    dev = imgs.device
    lib = _cabi.lib()
    ws_bytes = lib.vslam_features_workspace_bytes(B, H, W, float(min_distance))
    if ws_bytes < 0:
        raise ValueError("unsupported image size %dx%d" % (W, H))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    out = {
        "eig": torch.empty((B, H, W), dtype=torch.float32, device=dev),
        "blur": torch.empty((B, H, W), dtype=torch.uint8, device=dev),
        "corners": torch.empty((B, max_corners, 2), dtype=torch.float32, device=dev),
        "n_corners": torch.empty(B, dtype=torch.int32, device=dev),
        "kp": torch.empty((B, max_corners, 2), dtype=torch.float32, device=dev),
        "desc": torch.empty((B, max_corners, 32), dtype=torch.uint8, device=dev),
        "n_kp": torch.empty(B, dtype=torch.int32, device=dev),
        "status": torch.empty(B, dtype=torch.int32, device=dev),
    }
    ca, sa = _angle_terms(angle)
    with torch.cuda.device(dev):
        _cabi.check(lib.vslam_extract_features(
            imgs.data_ptr(), B, H, W, int(max_corners), float(quality), float(min_distance), ca, sa,
            _pattern_on(dev).data_ptr(), ws.data_ptr(), ws_bytes, out["eig"].data_ptr(), out["blur"].data_ptr(),
            out["corners"].data_ptr(), out["n_corners"].data_ptr(), out["kp"].data_ptr(), out["desc"].data_ptr(),
            out["n_kp"].data_ptr(), out["status"].data_ptr(), _cabi.stream_ptr()))
    out["_workspace"] = ws  # keeps the scratch alive until the caller has synchronised
    return out


class FeatureExtractor:
    """Drop-in for the reference's FeatureExtractor (src/v2/frame.py:5-14) with the ORB extractor."""

    def __init__(self, device="cuda", max_corners=3000, quality=0.01, min_distance=7):
        self.device = device
        self.max_corners, self.quality, self.min_distance = max_corners, quality, min_distance

    def compute_features(self, img):
        img = np.asarray(img)
        if img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3:
            raise ValueError("img must be a uint8 array of shape (H, W, 3) (cv2.imread layout)")
        o = extract_features_batch(torch.from_numpy(np.ascontiguousarray(img))[None].to(self.device),
                                   self.max_corners, self.quality, self.min_distance)
        if int(o["status"][0]) != 0:
            raise _cabi.VslamError("feature extraction: internal list overflow")
        n = int(o["n_kp"][0])
        return o["kp"][0, :n].cpu().numpy(), o["desc"][0, :n].cpu().numpy()
