"""Structure-of-arrays map for bundle adjustment.

The reference rebuilds a g2o graph from its object map on EVERY BA call: one dict probe per (point,
frame) pair and one pybind11 call per vertex and edge (src/v2/LocalBA.py:153-172, :198-214; containers
src/v2/map.py:28-44,60-69,120-131 and src/v2/point.py:14-46).  MapSoA keeps the same information as
growing arrays that are appended to as frames, points and observations arrive (amortised O(1)), hands
them to the device solver without any per-element Python work, and keeps the solver object alive: while the
structure is unchanged a repeated optimisation only uploads the state, and after an append the same object is
given the new structure (its device buffers stay while they fit).  This class serves the GLOBAL map and the full
BA that runs per key frame; the per-frame path of the tracking loop (append a frame, motion-only BA) lives on the
device in visual_slam_b200.window.TrackingWindow.

Conventions are the reference's: poses are camera->world 4x4 (Frame.pose, main.py:116-118), frame 0 is
fixed in local BA (LocalBA.py:155-156), key frames are fixed and all points are fixed in motion-only BA
(LocalBA.py:200-206), scale edges join a frame to its parents with the norm of the stored relative
translation as the measurement (LocalBA.py:159-162, :126).
"""
import numpy as np

from .ba import BAEngine
from .ba_problem import BAProblem, R_to_quat, quat_to_R


class _Grow:
    """Append-only numpy array with capacity doubling."""

    def __init__(self, width, dtype):
        self.a = np.zeros((16, width) if width else (16,), dtype=dtype)
        self.n = 0

    def append(self, v):
        if self.n == self.a.shape[0]:
            self.a = np.concatenate([self.a, np.zeros_like(self.a)])
        self.a[self.n] = v
        self.n += 1
        return self.n - 1

    def view(self):
        return self.a[:self.n]


class MapSoA:
    def __init__(self):
        self._cams = _Grow(7, np.float64)       # (qx, qy, qz, qw, tx, ty, tz), camera->world
        self._kf = _Grow(0, bool)
        self._pts = _Grow(3, np.float64)
        self._oc, self._op = _Grow(0, np.int32), _Grow(0, np.int32)
        self._uv = _Grow(2, np.float64)
        self._sa, self._sb, self._sm = _Grow(0, np.int32), _Grow(0, np.int32), _Grow(0, np.float64)
        self._fidx, self._pidx = {}, {}          # external id -> row
        self._orow = {}                          # (point row, frame row) -> observation row
        self._fids, self._pids = [], []
        self._version = 0                        # bumped on every structural change
        self._engine = None                      # (key, BAEngine)
        self.stats = None

    # ---- incremental construction ---------------------------------------------------------
    def AddFrame(self, frame_id, pose, keyframe=False):
        if frame_id in self._fidx:
            raise Exception("Duplicate frame warning")        # map.py:14
        pose = np.asarray(pose, dtype=np.float64)
        self._fidx[frame_id] = self._cams.append(np.concatenate([R_to_quat(pose[:3, :3]), pose[:3, 3]]))
        self._kf.append(bool(keyframe))
        self._fids.append(frame_id)
        self._version += 1

    def AddParent(self, child_id, parent_id, transition):
        """Frame.AddParent: the scale-edge measurement is |translation| of the stored relative pose."""
        self._sa.append(self._fidx[parent_id])
        self._sb.append(self._fidx[child_id])
        self._sm.append(float(np.linalg.norm(np.asarray(transition, dtype=np.float64)[:3, 3])))
        self._version += 1

    def AddPoint3D(self, point_id, point_3d):
        if point_id in self._pidx:
            return
        self._pidx[point_id] = self._pts.append(np.asarray(point_3d, dtype=np.float64).reshape(3))
        self._pids.append(point_id)
        self._version += 1

    def AddObservation(self, point_id, frame_id, uv):
        """Point.AddFrame(frame, uv, descriptor) without the descriptor (BA does not use it)."""
        if point_id not in self._pidx:
            raise Exception("No point yet added")             # map.py:92
        if frame_id not in self._fidx:
            raise Exception("No frame yet added")             # map.py:79
        f, p = self._fidx[frame_id], self._pidx[point_id]
        row = self._orow.get((p, f))
        if row is not None:
            # Point.frames is a dict keyed by frame id (point.py:14-46): a second observation of the same
            # (point, frame) replaces the first -- one BA edge, the last uv; the structure does not change
            self._uv.a[row] = np.asarray(uv, dtype=np.float64).reshape(2)
            return
        self._orow[(p, f)] = self._oc.append(f)
        self._op.append(p)
        self._uv.append(np.asarray(uv, dtype=np.float64).reshape(2))
        self._version += 1

    def UpdatePose(self, new_pose, frame_id):
        if frame_id not in self._fidx:
            raise Exception("No frame yet added")
        new_pose = np.asarray(new_pose, dtype=np.float64)
        self._cams.a[self._fidx[frame_id]] = np.concatenate([R_to_quat(new_pose[:3, :3]), new_pose[:3, 3]])

    def UpdatePoint3D(self, new_point, point_id):
        if point_id not in self._pidx:
            raise Exception("No point yet added")
        self._pts.a[self._pidx[point_id]] = np.asarray(new_point, dtype=np.float64).reshape(3)

    @classmethod
    def from_map(cls, map):
        """One pass over a reference-style Map (frames / points_3d dicts; Point.frames maps a frame id
        to (frame, uv, descriptor))."""
        m = cls()
        for fid, fr in map.frames.items():
            m.AddFrame(fid, fr.GetPose(), keyframe=bool(fr.IsKeyFrame()) if hasattr(fr, "IsKeyFrame") else False)
        for fid, fr in map.frames.items():
            if fid != 0 and hasattr(fr, "GetParentIDs"):
                for parent in fr.GetParentIDs():
                    m.AddParent(fid, parent, fr.GetTransitionWithParentID(parent))
        for pid, pt in map.points_3d.items():
            m.AddPoint3D(pid, pt.Get3dPoint())
            for fid, corr in pt.frames.items():
                if fid in m._fidx and corr is not None:
                    m.AddObservation(pid, fid, corr[1])
        return m

    # ---- views ----------------------------------------------------------------------------
    n_frames = property(lambda s: s._cams.n)
    n_points = property(lambda s: s._pts.n)
    n_observations = property(lambda s: s._oc.n)

    def GetPose(self, frame_id):
        c = self._cams.a[self._fidx[frame_id]]
        T = np.eye(4)
        T[:3, :3] = quat_to_R(c[:4])
        T[:3, 3] = c[4:]
        return T

    def GetAll3DPoints(self):
        return self._pts.view().copy()

    # ---- key-frame bookkeeping ---------------------------------------------------
    def GetNVisibleFrames(self):
        """Per point row: the number of distinct frames observing it (point.py:58-59)."""
        C = max(self._cams.n, 1)
        pairs = np.unique(self._op.view().astype(np.int64) * C + self._oc.view())
        return np.bincount(pairs // C, minlength=self._pts.n)

    def DiscardOutlierMapPoints(self, n_visible_frames=3):
        """map.py:124-131: drop the map points seen in fewer than n_visible_frames frames, together with
        their observations; the order of the remaining points is kept.  Vectorised (no per-point probes)."""
        keep = self.GetNVisibleFrames() >= n_visible_frames
        if keep.all():
            return 0
        new_row = np.cumsum(keep) - 1
        ok = keep[self._op.view()]
        oc, op, uv = self._oc.view()[ok], new_row[self._op.view()[ok]].astype(np.int32), self._uv.view()[ok]
        pts = self._pts.view()[keep]
        self._pids = [pid for pid, k in zip(self._pids, keep) if k]
        self._pidx = {pid: r for r, pid in enumerate(self._pids)}
        for g, v in ((self._oc, oc), (self._op, op), (self._uv, uv), (self._pts, pts)):
            g.a[:len(v)] = v
            g.n = len(v)
        self._orow = {(int(p), int(f)): r for r, (p, f) in enumerate(zip(op, oc))}
        self._version += 1
        return int((~keep).sum())

    def GetImagePointsWithFrameID(self, frame_id):
        """map.py:28-44: (image_points, descriptors, locations_3d, point_ids) of the map points seen by
        frame_id, in map order.  Descriptors are not kept in the arrays: that slot is None."""
        row = self._fidx[frame_id]
        sel = np.nonzero(self._oc.view() == row)[0]
        # a point keeps ONE correspondence per frame (Point.frames is a dict: the last one added wins)
        op = self._op.view()[sel]
        order = np.lexsort((sel, op))
        sel, op = sel[order], op[order]
        last = np.ones(len(op), bool)
        last[:-1] = op[1:] != op[:-1]
        sel, op = sel[last], op[last]
        ids = np.array([self._pids[r] for r in op])
        return self._uv.view()[sel].copy(), None, self._pts.view()[op].copy(), ids

    def problem(self, K, motion_only=False):
        """The BAProblem of the whole map: vectorised, no per-element Python work."""
        C, P = self._cams.n, self._pts.n
        if motion_only:
            cam_fixed = self._kf.view().copy()
            pt_fixed = np.ones(P, bool)
            sa = sb = sm = None
        else:
            cam_fixed = np.zeros(C, bool)
            if 0 in self._fidx:
                cam_fixed[self._fidx[0]] = True
            pt_fixed = np.zeros(P, bool)
            sa, sb, sm = self._sa.view(), self._sb.view(), self._sm.view()
        return BAProblem(self._cams.view(), self._pts.view(), self._oc.view(), self._op.view(), self._uv.view(),
                         K, cam_fixed, pt_fixed, sa, sb, sm)

    # ---- optimisation ---------------------------------------------------------------------
    def _run(self, camera, motion_only, max_iterations, device):
        K = (camera.fx, camera.fy, camera.cx, camera.cy)
        prob = self.problem(K, motion_only)
        key = (self._version, motion_only, device)
        if self._engine is not None and self._engine[0] == key:
            eng = self._engine[1]                       # same structure: only the state is uploaded
            eng.set_state(prob.cams, prob.pts, prob.obs_uv, prob.scale_meas, prob.K)
        elif self._engine is not None and self._engine[0][2] == device:
            eng = self._engine[1]                       # new structure, same engine object: its device buffers are
            eng.set_structure(prob)                     # re-used while they fit (vslam_ba_set_structure)
            self._engine = (key, eng)
        else:
            if self._engine is not None:
                self._engine[1].close()
            eng = BAEngine(prob, device)
            self._engine = (key, eng)
        self.stats = eng.optimize(max_iterations)
        cams, pts = eng.get_state()
        return cams.cpu().numpy(), pts.cpu().numpy()

    def localBundleAdjustement(self, camera, scale=False, max_iterations=10, device="cuda:0", sink=None):
        """LocalBA.py:143-190 on the arrays; `sink` (a reference-style Map) also receives the results
        through UpdatePose / UpdatePoint3D like the reference's write-back (:185-190)."""
        cams, pts = self._run(camera, False, max_iterations, device)
        median_depth = float(np.median(np.linalg.norm(pts, axis=1))) if (scale and len(pts)) else 1.0
        cams[:, 4:] /= median_depth
        pts = pts / median_depth
        self._cams.a[:self._cams.n] = cams
        self._pts.a[:self._pts.n] = pts
        if sink is not None:
            for fid in self._fids:
                sink.UpdatePose(new_pose=self.GetPose(fid), frame_id=fid)
            for pid in self._pids:
                sink.UpdatePoint3D(new_point=self._pts.a[self._pidx[pid]].copy(), point_id=pid)
        return self.stats

    def motionOnlyBundleAdjustement(self, camera, scale=False, max_iterations=10, device="cuda:0", sink=None):
        """LocalBA.py:195-229: key frames fixed, all points fixed, poses written back."""
        cams, pts = self._run(camera, True, max_iterations, device)
        median_depth = float(np.median(np.linalg.norm(pts, axis=1))) if (scale and len(pts)) else 1.0
        cams[:, 4:] /= median_depth
        self._cams.a[:self._cams.n] = cams
        if sink is not None:
            for fid in self._fids:
                sink.UpdatePose(new_pose=self.GetPose(fid), frame_id=fid)
        return self.stats

    def close(self):
        if self._engine is not None:
            self._engine[1].close()
            self._engine = None
