"""Descriptor matching: drop-in for FeatureMatcher (reference src/v2/frame.py:16-49).

`FeatureMatcher.match_features` keeps the reference signature and return objects
(list of [cv2.DMatch], matched keypoints and descriptors in query order); the
arithmetic runs in the sm_100a kernels of csrc/hamming.cu through the C ABI.
`match_batch` is the tensor-native entry used by the batched tracking path.
"""
import numpy as np
import torch

from . import _cabi


def _require_cuda():
    if not torch.cuda.is_available():
        raise _cabi.VslamError("visual_slam_b200 needs a CUDA device (no CPU fallback)")


def _counts(n, B, limit, device, what):
    if n is None:
        return None
    if n.dtype != torch.int32 or not n.is_cuda or n.shape != (B,):
        raise ValueError("%s must be an int32 CUDA tensor of shape (batch,)" % what)
    return n.contiguous()


def knn2_batch(q, t, want_rev=True, n_q=None, n_t=None):
    """q: uint8 cuda [B, nq, 32]; t: uint8 cuda [B, nt, 32].  n_q / n_t (int32 cuda [B], both or neither): real
    rows per pair of a ragged batch, the rest is padding (never matches: idx / dist -1).

    Returns (idx int32 [B,nq,2], dist int32 [B,nq,2], rev_key uint32-as-int32 [B,nt] or None),
    ordered exactly like cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2) (frame.py:23)."""
    _require_cuda()
    if q.dtype != torch.uint8 or t.dtype != torch.uint8:
        raise ValueError("descriptors must be uint8")
    if q.dim() != 3 or t.dim() != 3 or q.shape[2] != 32 or t.shape[2] != 32:
        raise ValueError("descriptors must be [batch, n, 32] (256-bit)")
    if q.shape[0] != t.shape[0]:
        raise ValueError("query/train batch mismatch")
    q = q.contiguous()
    t = t.contiguous()
    B, nq, nt = q.shape[0], q.shape[1], t.shape[1]
    idx = torch.empty((B, nq, 2), dtype=torch.int32, device=q.device)
    dist = torch.empty((B, nq, 2), dtype=torch.int32, device=q.device)
    rev = torch.empty((B, nt), dtype=torch.int32, device=q.device) if want_rev else None
    n_q, n_t = _counts(n_q, B, nq, q.device, "n_q"), _counts(n_t, B, nt, q.device, "n_t")
    if (n_q is None) != (n_t is None):
        raise ValueError("give both n_q and n_t, or neither")
    with torch.cuda.device(q.device):
        if n_q is None:
            _cabi.check(_cabi.lib().vslam_hamming_knn2(
                _cabi.ptr(q), nq, _cabi.ptr(t), nt, _cabi.ptr(idx), _cabi.ptr(dist), _cabi.ptr(rev),
                B, _cabi.stream_ptr()))
        else:
            _cabi.check(_cabi.lib().vslam_hamming_knn2_ragged(
                _cabi.ptr(q), nq, _cabi.ptr(n_q), _cabi.ptr(t), nt, _cabi.ptr(n_t), _cabi.ptr(idx), _cabi.ptr(dist),
                _cabi.ptr(rev), B, _cabi.stream_ptr()))
    return idx, dist, rev


def filter_batch(idx, dist, rev, nt, ratio=0.8, cross_check=False):
    """Ratio test (frame.py:33) + optional mutual-NN check (src/slam.py:40-47), compacted."""
    B, nq = idx.shape[0], idx.shape[1]
    keep_q = torch.empty((B, nq), dtype=torch.int32, device=idx.device)
    keep_t = torch.empty((B, nq), dtype=torch.int32, device=idx.device)
    n_keep = torch.empty((B,), dtype=torch.int32, device=idx.device)
    if cross_check and rev is None:
        raise ValueError("cross_check needs the reverse keys from knn2_batch(want_rev=True)")
    with torch.cuda.device(idx.device):
        _cabi.check(_cabi.lib().vslam_match_filter(
            _cabi.ptr(idx), _cabi.ptr(dist), _cabi.ptr(rev), nq, nt, float(ratio),
            1 if cross_check else 0, _cabi.ptr(keep_q), _cabi.ptr(keep_t), _cabi.ptr(n_keep),
            B, _cabi.stream_ptr()))
    return keep_q, keep_t, n_keep


def match_batch(q, t, ratio=0.8, cross_check=False, n_q=None, n_t=None):
    """Batched kNN-2 + ratio (+ cross-check).  Returns a dict of device tensors.  n_q / n_t: per-pair row counts of a
    ragged batch (e.g. the n_kp of features.extract_features_batch)."""
    idx, dist, rev = knn2_batch(q, t, want_rev=cross_check, n_q=n_q, n_t=n_t)
    keep_q, keep_t, n_keep = filter_batch(idx, dist, rev, t.shape[1], ratio, cross_check)
    return {"idx": idx, "dist": dist, "rev_key": rev, "keep_q": keep_q, "keep_t": keep_t,
            "n_keep": n_keep}


class FeatureMatcher:
    """Same interface as the reference class (src/v2/frame.py:16-49), Hamming/ORB variant
    (frame.py:8).  `cross_check=True` adds the mutual-nearest test of src/slam.py:40-47."""

    def __init__(self, cross_check=False, device="cuda:0"):
        self.cross_check = bool(cross_check)
        self.device = torch.device(device)
        self._pin_in = self._pin_out = None      # pinned staging, grown on demand and kept between calls

    def _staging(self, n_in, n_out):
        if self._pin_in is None or self._pin_in.numel() < n_in:
            self._pin_in = torch.empty(max(n_in, 1 << 16), dtype=torch.uint8).pin_memory()
        if self._pin_out is None or self._pin_out.numel() < n_out:
            self._pin_out = torch.empty(max(n_out, 1 << 12), dtype=torch.int32).pin_memory()
        return self._pin_in, self._pin_out

    def match_features(self, kp1, desc1, kp2, desc2, ratio=0.8):
        import cv2  # only to build cv2.DMatch return objects, as callers expect (main.py:187,216)
        desc1 = np.ascontiguousarray(desc1)
        desc2 = np.ascontiguousarray(desc2)
        if desc1.dtype != np.uint8 or desc2.dtype != np.uint8:
            raise ValueError("match_features: ORB/Hamming path needs uint8 descriptors")
        if desc1.ndim != 2 or desc2.ndim != 2 or desc1.shape[1] != 32 or desc2.shape[1] != 32:
            raise ValueError("match_features: descriptors must be (N, 32) uint8")
        if desc2.shape[0] < 2 and desc1.shape[0] > 0:
            # the reference unpacks `for m, n in rawMatches` (frame.py:30) and raises here too
            raise ValueError("not enough values to unpack (expected 2, got %d)" % desc2.shape[0])
        _require_cuda()
        nq, nt = desc1.shape[0], desc2.shape[0]
        kp1 = np.asarray(kp1)
        kp2 = np.asarray(kp2)
        if nq == 0:
            e = np.asarray([])
            return [], e, e.copy(), e.copy(), e.copy()
        # ONE host->device copy (both descriptor sets through a pinned buffer), the kernels, ONE device->host copy of
        # the packed result (count, kept query / train rows, best distances), one stream synchronisation
        pin_in, pin_out = self._staging((nq + nt) * 32, 1 + 3 * nq)
        pin_in[:nq * 32].view(nq, 32).numpy()[:] = desc1
        pin_in[nq * 32:(nq + nt) * 32].view(nt, 32).numpy()[:] = desc2
        with torch.cuda.device(self.device):
            both = pin_in[:(nq + nt) * 32].to(self.device, non_blocking=True)
            out = match_batch(both[:nq * 32].view(1, nq, 32), both[nq * 32:].view(1, nt, 32), ratio, self.cross_check)
            packed = torch.cat([out["n_keep"], out["keep_q"][0], out["keep_t"][0], out["dist"][0, :, 0]])
            res = pin_out[:1 + 3 * nq]
            res.copy_(packed, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
        r = res.numpy()
        n = int(r[0])
        kq, kt, d0 = r[1:1 + n].copy(), r[1 + nq:1 + nq + n].copy(), r[1 + 2 * nq:]
        matches = [[cv2.DMatch(a, b, 0, d)] for a, b, d in zip(kq.tolist(), kt.tolist(), d0[kq].astype(np.float64).tolist())]
        if n == 0:
            e = np.asarray([])
            return matches, e, e.copy(), e.copy(), e.copy()
        return matches, kp1[kq], desc1[kq], kp2[kt], desc2[kt]


class HostMatcher:
    """End-to-end batched matcher for HOST descriptor buffers: pinned staging, H2D copy, kNN-2 +
    ratio (+ cross-check) on the GPU, D2H of the compacted match lists.  Buffers are reused across
    calls; copies and kernels of consecutive chunks overlap on two streams."""

    def __init__(self, batch, nq, nt, ratio=0.8, cross_check=True, device="cuda:0", chunks=4):
        _require_cuda()
        self.device = torch.device(device)
        self.batch, self.nq, self.nt = batch, nq, nt
        self.ratio, self.cross_check = ratio, cross_check
        self.chunks = max(1, min(chunks, batch))
        self.q_pin = torch.empty((batch, nq, 32), dtype=torch.uint8).pin_memory()
        self.t_pin = torch.empty((batch, nt, 32), dtype=torch.uint8).pin_memory()
        self.keep_q_pin = torch.empty((batch, nq), dtype=torch.int32).pin_memory()
        self.keep_t_pin = torch.empty((batch, nq), dtype=torch.int32).pin_memory()
        self.n_keep_pin = torch.empty((batch,), dtype=torch.int32).pin_memory()
        with torch.cuda.device(self.device):
            self.q_dev = torch.empty((batch, nq, 32), dtype=torch.uint8, device=self.device)
            self.t_dev = torch.empty((batch, nt, 32), dtype=torch.uint8, device=self.device)
            self.streams = [torch.cuda.Stream(self.device) for _ in range(2)]
        self.h2d_bytes = self.q_pin.numel() + self.t_pin.numel()
        self.d2h_bytes = 4 * (self.keep_q_pin.numel() + self.keep_t_pin.numel() + self.n_keep_pin.numel())
        self.launches_per_call = 3 * self.chunks if cross_check else 2 * self.chunks

    def load(self, q, t):
        """Copies caller arrays (numpy uint8 [batch, n, 32]) into the pinned staging buffers."""
        self.q_pin.copy_(torch.from_numpy(np.ascontiguousarray(q)))
        self.t_pin.copy_(torch.from_numpy(np.ascontiguousarray(t)))

    def run(self, join=True):
        """H2D -> match -> D2H for the staged batch; returns pinned (keep_q, keep_t, n_keep), valid once the
        current stream has been synchronised (join=True) or after `join()` + a synchronisation (join=False).
        With join=False consecutive calls pipeline: the chunks of the next batch queue behind the same chunks of
        this one on the two internal streams, so its first host->device copies run under this batch's last
        kernels instead of waiting for its last device->host copy (a serving loop that stages batch i + 1 while
        batch i is matched; the pinned result buffers are overwritten by the next call)."""
        B = self.batch
        step = (B + self.chunks - 1) // self.chunks
        cur = torch.cuda.current_stream(self.device)
        start = torch.cuda.Event()
        start.record(cur)
        for c in range(self.chunks):
            lo, hi = c * step, min(B, (c + 1) * step)
            if lo >= hi:
                break
            st = self.streams[c & 1]
            st.wait_event(start)
            with torch.cuda.stream(st):
                self.q_dev[lo:hi].copy_(self.q_pin[lo:hi], non_blocking=True)
                self.t_dev[lo:hi].copy_(self.t_pin[lo:hi], non_blocking=True)
                out = match_batch(self.q_dev[lo:hi], self.t_dev[lo:hi], self.ratio, self.cross_check)
                self.keep_q_pin[lo:hi].copy_(out["keep_q"], non_blocking=True)
                self.keep_t_pin[lo:hi].copy_(out["keep_t"], non_blocking=True)
                self.n_keep_pin[lo:hi].copy_(out["n_keep"], non_blocking=True)
        if join:
            self.join()
        return self.keep_q_pin, self.keep_t_pin, self.n_keep_pin

    def join(self):
        """Makes the current stream wait for everything `run(join=False)` has enqueued."""
        cur = torch.cuda.current_stream(self.device)
        for st in self.streams:
            cur.wait_stream(st)
