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


class FeatureBuffers:
    """Scratch + result buffers of one extract_features_batch call shape, reusable across calls."""

    def __init__(self, batch, H, W, max_corners=3000, min_distance=7.0, device="cuda"):
        lib = _cabi.lib()
        self.shape = (batch, H, W, max_corners, float(min_distance))
        self.ws_bytes = lib.vslam_features_workspace_bytes(batch, H, W, float(min_distance))
        if self.ws_bytes < 0:
            raise ValueError("unsupported frame size %dx%d (16 <= side, at most ~1.6 Mpixel)" % (W, H))
        dev = torch.device(device)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=dev)
        self.out = {
            "eig": torch.empty((batch, H, W), dtype=torch.float32, device=dev),
            "blur": torch.empty((batch, H, W), dtype=torch.uint8, device=dev),
            "corners": torch.empty((batch, max_corners, 2), dtype=torch.float32, device=dev),
            "n_corners": torch.empty(batch, dtype=torch.int32, device=dev),
            "kp": torch.empty((batch, max_corners, 2), dtype=torch.float32, device=dev),
            "desc": torch.empty((batch, max_corners, 32), dtype=torch.uint8, device=dev),
            "n_kp": torch.empty(batch, dtype=torch.int32, device=dev),
            "status": torch.empty(batch, dtype=torch.int32, device=dev),
        }


def extract_features_batch(imgs, max_corners=3000, quality=0.01, min_distance=7.0, angle=-1.0, buffers=None):
    """imgs: uint8 CUDA tensor (B, H, W, 3), BGR.  Returns a dict of CUDA tensors:
    kp (B, max_corners, 2) f32 + n_kp (B,) i32 + desc (B, max_corners, 32) u8 (rows >= n_kp are undefined),
    corners / n_corners (goodFeaturesToTrack before ORB's border filter), eig (B, H, W) f32, blur (B, H, W) u8.
    `buffers` (a FeatureBuffers of the same shape) makes the call allocation-free; its tensors are returned."""
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
    if buffers is None:
        buffers = FeatureBuffers(B, H, W, max_corners, min_distance, dev)
    elif buffers.shape != (B, H, W, max_corners, float(min_distance)):
        raise ValueError("buffers were made for another call shape")
    out = dict(buffers.out)
    ca, sa = _angle_terms(angle)
    with torch.cuda.device(dev):
        _cabi.check(lib.vslam_extract_features(
            imgs.data_ptr(), B, H, W, int(max_corners), float(quality), float(min_distance), ca, sa,
            _pattern_on(dev).data_ptr(), buffers.ws.data_ptr(), buffers.ws_bytes, out["eig"].data_ptr(),
            out["blur"].data_ptr(), out["corners"].data_ptr(), out["n_corners"].data_ptr(), out["kp"].data_ptr(),
            out["desc"].data_ptr(), out["n_kp"].data_ptr(), out["status"].data_ptr(), _cabi.stream_ptr()))
    out["_workspace"] = buffers.ws  # keeps the scratch alive until the caller has synchronised
    return out


class HostFeatureExtractor:
    """End-to-end batched extractor for HOST frames: pinned staging, H2D copy of the BGR frames, the device
    pipeline, D2H of key points + descriptors + counts.  The batch is cut into chunks whose copies and kernels
    overlap on two streams; every buffer is allocated once."""

    def __init__(self, batch, H, W, max_corners=3000, quality=0.01, min_distance=7.0, device="cuda:0", chunks=3):
        self.device = torch.device(device)
        self.batch, self.H, self.W = batch, H, W
        self.max_corners, self.quality, self.min_distance = max_corners, quality, min_distance
        self.chunks = max(1, min(chunks, batch))
        self.step = (batch + self.chunks - 1) // self.chunks
        self.img_pin = torch.empty((batch, H, W, 3), dtype=torch.uint8).pin_memory()
        self.kp_pin = torch.empty((batch, max_corners, 2), dtype=torch.float32).pin_memory()
        self.desc_pin = torch.empty((batch, max_corners, 32), dtype=torch.uint8).pin_memory()
        self.n_kp_pin = torch.empty((batch,), dtype=torch.int32).pin_memory()
        with torch.cuda.device(self.device):
            self.img_dev = torch.empty((batch, H, W, 3), dtype=torch.uint8, device=self.device)
            self.streams = [torch.cuda.Stream(self.device) for _ in range(2)]
            self.bufs = {}
            for c in range(self.chunks):
                n = min(batch, (c + 1) * self.step) - c * self.step
                if n > 0 and (c & 1, n) not in self.bufs:   # one set per stream and chunk size
                    self.bufs[(c & 1, n)] = FeatureBuffers(n, H, W, max_corners, min_distance, self.device)
        self.h2d_bytes = self.img_pin.numel()
        self.d2h_bytes = self.kp_pin.numel() * 4 + self.desc_pin.numel() + self.n_kp_pin.numel() * 4
        self.launches_per_call = 7 * self.chunks

    def load(self, frames):
        """Copies caller frames (numpy uint8 [batch, H, W, 3], BGR) into the pinned staging buffer."""
        self.img_pin.copy_(torch.from_numpy(np.ascontiguousarray(frames)))

    def run(self, join=True):
        """H2D -> extract -> D2H for the staged batch; returns pinned (kp, desc, n_kp).  join=False leaves the two
        internal streams un-joined so that consecutive calls pipeline (the next batch's frame copies run under this
        batch's kernels); call join() before reading the pinned results."""
        cur = torch.cuda.current_stream(self.device)
        start = torch.cuda.Event()
        start.record(cur)
        for c in range(self.chunks):
            lo, hi = c * self.step, min(self.batch, (c + 1) * self.step)
            if lo >= hi:
                break
            st = self.streams[c & 1]
            st.wait_event(start)
            with torch.cuda.stream(st):
                self.img_dev[lo:hi].copy_(self.img_pin[lo:hi], non_blocking=True)
                o = extract_features_batch(self.img_dev[lo:hi], self.max_corners, self.quality, self.min_distance,
                                           buffers=self.bufs[(c & 1, hi - lo)])
                self.kp_pin[lo:hi].copy_(o["kp"], non_blocking=True)
                self.desc_pin[lo:hi].copy_(o["desc"], non_blocking=True)
                self.n_kp_pin[lo:hi].copy_(o["n_kp"], non_blocking=True)
        if join:
            self.join()
        return self.kp_pin, self.desc_pin, self.n_kp_pin

    def join(self):
        """Makes the current stream wait for everything run(join=False) has enqueued."""
        cur = torch.cuda.current_stream(self.device)
        for st in self.streams:
            cur.wait_stream(st)


class FeatureExtractor:
    """Drop-in for the reference's FeatureExtractor (src/v2/frame.py:5-14) with the ORB extractor."""

    def __init__(self, device="cuda", max_corners=3000, quality=0.01, min_distance=7):
        self.device = device
        self.max_corners, self.quality, self.min_distance = max_corners, quality, min_distance
        self._buffers = {}   # frame shape -> FeatureBuffers (the reference calls this once per frame of one camera)

    def compute_features(self, img):
        img = np.asarray(img)
        if img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3:
            raise ValueError("img must be a uint8 array of shape (H, W, 3) (cv2.imread layout)")
        key = img.shape[:2]
        if key not in self._buffers:
            self._buffers[key] = FeatureBuffers(1, key[0], key[1], self.max_corners, self.min_distance, self.device)
        o = extract_features_batch(torch.from_numpy(np.ascontiguousarray(img))[None].to(self.device),
                                   self.max_corners, self.quality, self.min_distance, buffers=self._buffers[key])
        if int(o["status"][0]) != 0:
            raise _cabi.VslamError("feature extraction: internal list overflow")
        n = int(o["n_kp"][0])
        return o["kp"][0, :n].cpu().numpy(), o["desc"][0, :n].cpu().numpy()


def frontend_batch(frames_a, frames_b, ratio=0.8, cross_check=False, max_corners=3000, quality=0.01,
                   min_distance=7.0, buffers=(None, None)):
    """The per-frame front end of the reference's loop (src/v2/main.py:96,185: compute_features on both frames,
    then match_features) for a batch of frame pairs, device-resident throughout: frames (B, H, W, 3) uint8 CUDA
    -> key points + descriptors of both -> ragged kNN-2 + ratio (+ cross-check).  Returns the two feature dicts
    and the match dict (keep_q / keep_t index the key points of frames_a / frames_b; n_keep per pair).
    `buffers`: two FeatureBuffers (one per side) to make the feature stage allocation-free."""
    from .matching import match_batch
    fa = extract_features_batch(frames_a, max_corners, quality, min_distance, buffers=buffers[0])
    fb = extract_features_batch(frames_b, max_corners, quality, min_distance, buffers=buffers[1])
    m = match_batch(fa["desc"], fb["desc"], ratio, cross_check, n_q=fa["n_kp"], n_t=fb["n_kp"])
    return fa, fb, m
