"""Device-resident tracking window: the local map of the reference's tracking loop
(src/v2/main.py:153-214) as device arrays that tracked frames are appended to, and the motion-only bundle
adjustment of src/v2/LocalBA.py:195-229 as one kernel launch over them.  Arithmetic in csrc/ba_window.cuh through
the C ABI (vslam_window_*); no CPU fallback.

Per tracked frame the reference (a) collects the key frame's mapped points by probing every Point object
(map.py:28-44), (b) matches, (c) runs PnP, (d) adds the frame and one observation per ratio match to its Map
objects (main.py:208-210), (e) rebuilds a g2o graph from all of them and optimises (main.py:213-214).  `track`
below does (b)-(e) on the device without a host round trip: match_batch -> gather -> RANSAC PnP -> pose plumbing
-> window append -> motion-only BA; only the LM statistics come back if asked for."""
import ctypes as C

import numpy as np
import torch

from . import _cabi
from .ba_problem import R_to_quat, quat_to_R
from .poses import camera_to_world_batch, guess_from_cam7_batch
from .tracking import gather_correspondences
from .matching import match_batch
from .pnp import pnp_ransac_batch


def _stats(a):
    keys = ("initial_cost", "final_cost", "iterations", "trials", "lambda", "pcg_iterations", "rejected",
            "pcg_breakdowns")
    return dict(zip(keys, [float(v) for v in a]))


class TrackingWindow:
    def __init__(self, max_frames=64, slot=2048, max_points=2048, device="cuda:0"):
        if not torch.cuda.is_available():
            raise _cabi.VslamError("visual_slam_b200 needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.max_frames, self.slot, self.max_points = int(max_frames), int(slot), int(max_points)
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_window_create(C.byref(self._h), self.max_frames, self.slot, self.max_points))
        self.K = None
        self._keep = None

    def close(self):
        if self._h:
            _cabi.lib().vslam_window_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _t(self, a, dtype):
        if isinstance(a, torch.Tensor):
            return a.to(self.device, dtype).contiguous()
        return torch.from_numpy(np.ascontiguousarray(a, dtype={torch.float64: np.float64, torch.float32: np.float32,
                                                               torch.int32: np.int32}[dtype])).to(self.device)

    # ---- construction ---------------------------------------------------------------------------------
    def reset(self, points, kf_uv, kf_pose, K):
        """A new key frame (main.py:153-162, :333-345): points (P,3) the map points it sees, kf_uv (P,2) their pixels in
        it, kf_pose its camera->world pose (4x4 array, or a 7-vector qx,qy,qz,qw,tx,ty,tz), K = (fx, fy, cx, cy)."""
        pts = self._t(points, torch.float64).reshape(-1, 3)
        uv = self._t(kf_uv, torch.float32).reshape(-1, 2)
        if pts.shape[0] != uv.shape[0] or pts.shape[0] < 1:
            raise ValueError("points / kf_uv disagree in length (or are empty)")
        if pts.shape[0] > self.max_points:
            raise ValueError("more points than the window was created for")
        cam7 = self._cam7(kf_pose)
        self.K = tuple(float(k) for k in K)
        k = (C.c_double * 4)(*self.K)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_window_reset(self._h, _cabi.ptr(pts), _cabi.ptr(uv), pts.shape[0],
                                                       _cabi.ptr(cam7), k, _cabi.stream_ptr()))
        self._keep = (pts, uv, cam7)

    def _cam7(self, pose):
        if isinstance(pose, torch.Tensor) and pose.numel() == 7:
            return pose.to(self.device, torch.float64).contiguous().reshape(7)
        pose = np.asarray(pose, dtype=np.float64)
        if pose.size == 7:
            return self._t(pose.reshape(7), torch.float64)
        return self._t(np.concatenate([R_to_quat(pose[:3, :3]), pose[:3, 3]]), torch.float64)

    def append(self, pose, keep_q, keep_t, n_keep, kp_cur):
        """main.py:208-210: one tracked frame.  pose: camera->world (4x4, 7-vector, or the device `cam7` row that
        poses.camera_to_world_batch / tracking.track_batch(want_c2w=True) return); keep_q / keep_t int32 [>= n_keep],
        n_keep int32 [1], kp_cur f32 [nt,2] -- device tensors are used in place (the outputs of match_batch)."""
        cam7 = self._cam7(pose)
        kq, kt = self._t(keep_q, torch.int32).reshape(-1), self._t(keep_t, torch.int32).reshape(-1)
        nk = self._t(n_keep, torch.int32).reshape(-1)[:1]
        kp = self._t(kp_cur, torch.float32).reshape(-1, 2)
        if kq.numel() < 1 or kq.numel() != kt.numel():
            raise ValueError("keep_q / keep_t disagree in length")
        if kq.numel() < self.slot:                      # the kernel reads up to min(n_keep, slot) rows
            pad = torch.full((self.slot - kq.numel(),), -1, dtype=torch.int32, device=self.device)
            kq, kt = torch.cat([kq, pad]), torch.cat([kt, pad])
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_window_append(self._h, _cabi.ptr(cam7), _cabi.ptr(kq), _cabi.ptr(kt),
                                                        _cabi.ptr(nk), _cabi.ptr(kp), _cabi.stream_ptr()))

    # ---- optimisation ---------------------------------------------------------------------------------
    def motion_only_ba(self, max_iterations=10, want_stats=True):
        """LocalBA.py:195-229 on the window.  want_stats=False enqueues the kernel and returns None (no sync)."""
        out = (C.c_double * 8)() if want_stats else None
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_window_motion_ba(self._h, int(max_iterations), out, _cabi.stream_ptr()))
        return _stats(out) if want_stats else None

    def info(self):
        o = (C.c_int32 * 4)()
        _cabi.check(_cabi.lib().vslam_window_info(self._h, o, None, None))
        return {"frames": o[0], "points": o[1], "max_frames": o[2], "slot": o[3]}

    n_frames = property(lambda s: s.info()["frames"])

    def cams(self, first=0, count=None):
        """Device tensor [count, 7] (a copy) of the camera->world poses from frame `first` on; key frame = row 0."""
        n = self.n_frames
        count = n - first if count is None else count
        out = torch.empty((count, 7), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_window_read_cams(self._h, _cabi.ptr(out), int(first), int(count),
                                                           _cabi.stream_ptr()))
        return out

    def pose(self, i):
        c = self.cams()[i].cpu().numpy()
        T = np.eye(4)
        T[:3, :3] = quat_to_R(c[:4])
        T[:3, 3] = c[4:]
        return T

    # ---- the per-frame tracking step, device-resident ------------------------------------------------------
    def set_keyframe_features(self, desc_kf):
        """Descriptors (P,32) uint8 of the key frame's mapped points, in the order of `points` given to reset()."""
        self._desc_kf = (desc_kf if isinstance(desc_kf, torch.Tensor)
                         else torch.from_numpy(np.ascontiguousarray(desc_kf))).to(self.device).contiguous()[None]

    def track(self, desc_cur, kp_cur, words, ratio=0.8, max_iterations=10, want_stats=False):
        """main.py:185-214 for one frame on the device: match against the key frame's mapped points, PnP with the
        previous frame's pose as the guess, camera->world, append, motion-only BA.  desc_cur u8 [nt,32], kp_cur f32
        [nt,2], words i32 [H,4] (device tensors).  Returns the dict of device results (+ LM stats if asked)."""
        pts, _, _ = self._keep
        guess = guess_from_cam7_batch(self.cams(self.n_frames - 1, 1))      # main.py:191-194, previous frame's pose
        m = match_batch(self._desc_kf, desc_cur[None], ratio, False)
        X, uv = gather_correspondences(m["keep_q"], m["keep_t"], m["n_keep"], pts.float()[None], kp_cur[None])
        o = pnp_ransac_batch(X, uv, m["n_keep"], words[None], self.K, guess)
        p = camera_to_world_batch(o["pose"])
        self.append(p["cam7"][0], m["keep_q"][0], m["keep_t"][0], m["n_keep"], kp_cur)
        st = self.motion_only_ba(max_iterations, want_stats)
        o.update(keep_q=m["keep_q"], keep_t=m["keep_t"], n_keep=m["n_keep"], cam7=p["cam7"], c2w=p["c2w"], stats=st)
        return o
