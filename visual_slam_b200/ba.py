"""Bundle adjustment: tensor-level engine + drop-in for the reference's BundleAdjustment class
(src/v2/LocalBA.py:20-229).  All arithmetic runs in csrc/ba*.cu through the C ABI."""
import ctypes as C

import numpy as np
import torch

from . import _cabi
from .ba_problem import BAProblem, R_to_quat, quat_to_R


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class BAEngine:
    """Owns a vslam_ba object for one problem structure; state lives in HBM."""

    def __init__(self, prob, device="cuda:0"):
        if not torch.cuda.is_available():
            raise _cabi.VslamError("visual_slam_b200 needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self._h = C.c_void_p()
        self.set_structure(prob)

    def set_structure(self, prob):
        """(Re)builds the problem structure for `prob` and uploads its state.  An engine that already exists keeps
        its device buffers while the new sizes fit (vslam_ba_set_structure): no cudaMalloc / cudaFree per BA call."""
        self.prob = prob
        oc, op = _i32(prob.obs_cam), _i32(prob.obs_pt)
        cf = np.ascontiguousarray(prob.cam_fixed, dtype=np.uint8)
        pf = np.ascontiguousarray(prob.pt_fixed, dtype=np.uint8)
        sa, sb = _i32(prob.scale_a), _i32(prob.scale_b)
        self._keep = (oc, op, cf, pf, sa, sb)
        p = lambda a: a.ctypes.data_as(C.c_void_p) if a.size else None  # noqa: E731
        args = (prob.n_cams, prob.n_pts, prob.n_obs, prob.n_scale, p(oc), p(op), p(cf), p(pf), p(sa), p(sb))
        with torch.cuda.device(self.device):
            if self._h:
                _cabi.check(_cabi.lib().vslam_ba_set_structure(self._h, *args))
            else:
                _cabi.check(_cabi.lib().vslam_ba_create(C.byref(self._h), *args))
        self.set_state(prob.cams, prob.pts, prob.obs_uv, prob.scale_meas, prob.K)

    def close(self):
        if self._h:
            _cabi.lib().vslam_ba_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dev(self, a):
        if isinstance(a, torch.Tensor):
            return a.to(self.device, torch.float64).contiguous()
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.device)

    def set_state(self, cams, pts, obs_uv, scale_meas, K):
        self._cams, self._pts = self._dev(cams), self._dev(pts)
        self._uv, self._sm = self._dev(obs_uv), self._dev(scale_meas)
        k = (C.c_double * 4)(*[float(v) for v in K])
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_set_state(
                self._h, _cabi.ptr(self._cams), _cabi.ptr(self._pts), _cabi.ptr(self._uv),
                _cabi.ptr(self._sm), k, _cabi.stream_ptr()))

    def get_state(self):
        cams = torch.empty((self.prob.n_cams, 7), dtype=torch.float64, device=self.device)
        pts = torch.empty((self.prob.n_pts, 3), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_get_state(self._h, _cabi.ptr(cams), _cabi.ptr(pts),
                                                      _cabi.stream_ptr()))
        return cams, pts

    def set_solver(self, pcg_tol=1e-10, pcg_max_iter=0):
        _cabi.check(_cabi.lib().vslam_ba_set_solver(self._h, float(pcg_tol), int(pcg_max_iter)))

    def set_linear_solver(self, kind="auto"):
        """"auto" = "pcg" (block-Jacobi PCG, the default and the faster one); "direct": LDL^T in one CTA, n <= 96."""
        _cabi.check(_cabi.lib().vslam_ba_set_linear_solver(self._h, {"auto": 0, "pcg": 1, "direct": 2, "ldlt": 3}[kind]))

    def set_schur_variant(self, variant="units"):
        """Tuning / tests: "units" (sorted units over G, the default) or "chunks" (resident Y per camera chunk; slower)."""
        _cabi.check(_cabi.lib().vslam_ba_set_schur_variant(self._h, {"chunks": 0, "units": 1}[variant]))

    def dims(self):
        out = (C.c_int32 * 8)()
        _cabi.check(_cabi.lib().vslam_ba_dims(self._h, out))
        keys = ("free_cams", "free_pts", "n", "items", "units", "blocks", "pcg_ctas", "pcg_rows")
        return dict(zip(keys, list(out)))

    def linearize(self):
        out = (C.c_double * 2)()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_linearize(self._h, out, _cabi.stream_ptr()))
        return {"cost": out[0], "max_diag": out[1]}

    def time_stage(self, stage, reps=5, flush=None):
        """CUDA-event times (ms) of `reps` launches of one stage: "linearize" (the fused kernel alone),
        "schur" or "trial"; `flush` is an optional device tensor overwritten before each launch (L2)."""
        code = {"linearize": 0, "schur": 1, "trial": 2, "linearize_back_to_back": 3}[stage]
        out = (C.c_float * reps)()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_time_stage(
                self._h, code, int(reps), _cabi.ptr(flush) if flush is not None else None,
                flush.numel() * flush.element_size() if flush is not None else 0, out, _cabi.stream_ptr()))
        return list(out)[:1] if code == 3 else list(out)

    def trace_linearize(self):
        """Per-CTA (kind, start_ns, end_ns) of one fused linearisation launch (profiling)."""
        cap = 3 * (1 << 20)
        out = (C.c_int64 * cap)()
        n = C.c_int32(0)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_trace_linearize(self._h, out, cap, C.byref(n), _cabi.stream_ptr()))
        import numpy as np
        a = np.ctypeslib.as_array(out)
        self.last_trace_marks = a[3 * n.value:3 * n.value + 7].copy()   # stamp before, controller end, stamp after, controller start
        return a[:3 * n.value].reshape(-1, 3).copy()

    def trial(self, lam):
        out = (C.c_double * 4)()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_trial(self._h, float(lam), out, _cabi.stream_ptr()))
        return {"cost": out[0], "scale": out[1], "pcg_status": int(out[2]), "pcg_iters": int(out[3])}

    def accept(self):
        _cabi.check(_cabi.lib().vslam_ba_accept(self._h))

    def read(self, what):
        d = self.dims()
        sizes = {"Hpp": (0, (d["free_cams"], 6, 6)), "bp": (1, (d["n"],)), "Hll": (2, (d["free_pts"], 6)),
                 "bl": (3, (d["free_pts"], 3)), "S": (4, (d["n"], d["n"])), "bhat": (5, (d["n"],)),
                 "dp": (6, (d["n"],)), "dl": (7, (d["free_pts"], 3)),
                 "trial_cams": (8, (self.prob.n_cams, 7)), "trial_pts": (9, (self.prob.n_pts, 3))}
        code, shape = sizes[what]
        out = torch.empty(shape, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_read(self._h, code, _cabi.ptr(out), out.numel(),
                                                 _cabi.stream_ptr()))
        return out

    def optimize(self, max_iterations=10):
        stats = (C.c_double * 8)()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().vslam_ba_optimize(self._h, int(max_iterations), stats,
                                                     _cabi.stream_ptr()))
        keys = ("initial_cost", "final_cost", "iterations", "trials", "lambda", "pcg_iterations",
                "rejected", "pcg_breakdowns")
        return dict(zip(keys, list(stats)))


def optimize_batch(engines, max_iterations=10):
    """Optimises independent problems concurrently (vslam_ba_optimize_batch).  Returns one stats dict per engine."""
    n = len(engines)
    hs = (C.c_void_p * n)(*[e._h for e in engines])
    stats = (C.c_double * (8 * n))()
    if n:
        with torch.cuda.device(engines[0].device):
            _cabi.check(_cabi.lib().vslam_ba_optimize_batch(hs, n, int(max_iterations), stats))
    keys = ("initial_cost", "final_cost", "iterations", "trials", "lambda", "pcg_iterations", "rejected",
            "pcg_breakdowns")
    return [dict(zip(keys, list(stats[8 * i:8 * i + 8]))) for i in range(n)]


def pcg_solve(S, b, tol=1e-10, max_iter=2000):
    """Block-Jacobi PCG on a dense symmetric system (device tensors).  Returns (x, status, iters)."""
    S = S.contiguous(); b = b.contiguous()
    x = torch.zeros_like(b)
    out = (C.c_int32 * 2)()
    with torch.cuda.device(S.device):
        _cabi.check(_cabi.lib().vslam_pcg_solve(_cabi.ptr(S), _cabi.ptr(b), _cabi.ptr(x), S.shape[0],
                                               float(tol), int(max_iter), out, _cabi.stream_ptr()))
    return x, int(out[0]), int(out[1])


_POOL = {}    # device -> BAEngine kept between calls: the reference builds a fresh optimiser for EVERY BA call
              # (main.py:146,213,322); here only the structure arrays change, the device buffers stay


def direct_solve(S, b, kernel="auto"):
    """In-CTA direct solve of a dense symmetric system (device tensors, n <= 96): register-resident Gauss-Jordan
    for n <= 63, shared-memory LDL^T above (kernel="ldlt" forces the latter).  Returns (x, status)."""
    S = S.contiguous(); b = b.contiguous()
    x = torch.zeros_like(b)
    out = C.c_int32(0)
    with torch.cuda.device(S.device):
        _cabi.check(_cabi.lib().vslam_direct_solve(_cabi.ptr(S), _cabi.ptr(b), _cabi.ptr(x),
                                                   -S.shape[0] if kernel == "ldlt" else S.shape[0],
                                                   C.byref(out), _cabi.stream_ptr()))
    return x, int(out.value)


def solve_problem(prob, max_iterations=10, device="cuda:0", pcg_tol=1e-10):
    """Host-buffer entry: optimise a BAProblem in place (H2D, LM on the GPU, D2H).  The engine object of a device
    is re-used from call to call (capacity-based buffers, see BAEngine.set_structure)."""
    key = str(torch.device(device))
    eng = _POOL.get(key)
    if eng is None:
        eng = _POOL[key] = BAEngine(prob, device)
    else:
        eng.set_structure(prob)
    eng.set_solver(pcg_tol)
    stats = eng.optimize(max_iterations)
    cams, pts = eng.get_state()
    prob.cams = cams.cpu().numpy()
    prob.pts = pts.cpu().numpy()
    return stats


def release_pool():
    """Frees the pooled engines (their device memory)."""
    for eng in _POOL.values():
        eng.close()
    _POOL.clear()


class _Pose:
    """Minimal stand-in for the g2o SBACam/Isometry3d estimate returned by get_pose()."""

    def __init__(self, cam7):
        self._c = np.asarray(cam7, dtype=np.float64)

    def matrix(self):
        T = np.eye(4)
        T[:3, :3] = quat_to_R(self._c[0:4])
        T[:3, 3] = self._c[4:7]
        return T

    def position(self):
        return self._c[4:7].copy()

    translation = position


class BundleAdjustment:
    """Drop-in for reference src/v2/LocalBA.py:20-229: same method names and argument meaning.
    `camera` is any object with fx, fy, cx, cy (main.py:24-30)."""

    def __init__(self, camera, device="cuda:0"):
        self.fx, self.fy, self.cx, self.cy = camera.fx, camera.fy, camera.cx, camera.cy
        self.focal_length = (camera.fx, camera.fy)
        self.principal_point = (camera.cx, camera.cy)
        self.baseline = 0
        self.device = device
        self._poses, self._points, self._scale = {}, {}, []
        self._e_pt, self._e_pose, self._e_uv = [], [], []     # edges as three parallel lists (bulk-converted once)
        self._result = None
        self.stats = None

    # -- graph construction (LocalBA.py:56-131) ------------------------------------------------
    def add_pose(self, pose_id, pose, fixed=False):
        pose = np.asarray(pose, dtype=np.float64)
        self._poses[pose_id] = (np.concatenate([R_to_quat(pose[0:3, 0:3]), pose[0:3, 3]]), bool(fixed))

    def add_point(self, point_id, point, fixed=False, marginalized=True):
        if point_id in self._points:
            print("WARNING: tried to add already existing point!")  # LocalBA.py:77
            return
        self._points[point_id] = (point, bool(fixed))

    def add_edge(self, point_id, pose_id, measurement, edge_id=None, information=None, robust_kernel=None):
        if information is not None and not np.array_equal(np.asarray(information), np.identity(2)):
            raise ValueError("only the reference's information matrix I2 is supported")
        self._e_pt.append(point_id)
        self._e_pose.append(pose_id)
        self._e_uv.append(measurement)

    def AddScalingEdge(self, parent_id, child_id, measurement, information=None, robust_kernel=None):
        m = np.asarray(measurement, dtype=np.float64)
        self._scale.append((parent_id, child_id, float(np.linalg.norm(m[0:3, 3]))))  # LocalBA.py:126

    def _problem(self):
        pid = {k: i for i, k in enumerate(self._poses)}
        qid = {k: i for i, k in enumerate(self._points)}
        cams = np.array([v[0] for v in self._poses.values()]).reshape(-1, 7)
        pts = np.asarray([v[0] for v in self._points.values()], dtype=np.float64).reshape(-1, 3)
        cf = np.array([v[1] for v in self._poses.values()], dtype=bool)
        pf = np.array([v[1] for v in self._points.values()], dtype=bool)
        oc = np.fromiter(map(pid.__getitem__, self._e_pose), dtype=np.int32, count=len(self._e_pose))
        op = np.fromiter(map(qid.__getitem__, self._e_pt), dtype=np.int32, count=len(self._e_pt))
        uv = (np.concatenate([np.ravel(u) for u in self._e_uv]) if self._e_uv and not isinstance(self._e_uv[0], np.ndarray)
              else np.concatenate(self._e_uv) if self._e_uv else np.zeros(0)).astype(np.float64).reshape(-1, 2)
        sa = np.array([pid[s[0]] for s in self._scale], dtype=np.int32)
        sb = np.array([pid[s[1]] for s in self._scale], dtype=np.int32)
        sm = np.array([s[2] for s in self._scale], dtype=np.float64)
        prob = BAProblem(cams, pts, oc, op, uv, (self.fx, self.fy, self.cx, self.cy), cf, pf, sa, sb, sm)
        return prob, pid, qid

    def optimize(self, max_iterations=10, verbose=True):
        prob, pid, qid = self._problem()
        self.stats = solve_problem(prob, max_iterations, self.device)
        self._result = (prob, pid, qid)

    def get_pose(self, pose_id):
        prob, pid, _ = self._result if self._result else self._problem()
        return _Pose(prob.cams[pid[pose_id]])

    def get_point(self, point_id):
        prob, _, qid = self._result if self._result else self._problem()
        return prob.pts[qid[point_id]].copy()

    # -- LocalBA.py:143-190 -------------------------------------------------------------------
    def localBundleAdjustement(self, map, last_keyframe_id=None, scale=False, BAwindow=5):
        frame_ids = list(map.frames.keys())
        point_ids = list(map.points_3d.keys())
        if last_keyframe_id is not None:
            point_ids = list(map.GetPointsVisibleToFrames(frame_ids))
        for frame_id in frame_ids:
            frame_obj = map.GetFrame(frame_id)
            if frame_id == 0:
                self.add_pose(pose_id=frame_id, pose=frame_obj.GetPose(), fixed=True)
            else:
                self.add_pose(pose_id=frame_id, pose=frame_obj.GetPose())
                for parent_ID in frame_obj.GetParentIDs():
                    self.AddScalingEdge(parent_id=parent_ID, child_id=frame_id,
                                        measurement=frame_obj.GetTransitionWithParentID(parent_ID))
        self._add_points_and_edges(map, frame_ids, point_ids, fixed=False)
        self.optimize()
        median_depth = 1
        if scale:
            norms = [np.linalg.norm(self.get_point(p)) for p in point_ids]
            median_depth = np.median(np.array(norms))
        for frame_id in frame_ids:
            new_pose = self.get_pose(frame_id).matrix()
            new_pose[0:3, 3] /= median_depth
            map.UpdatePose(new_pose=new_pose, frame_id=frame_id)
        for point_id in point_ids:
            map.UpdatePoint3D(new_point=self.get_point(point_id) / median_depth, point_id=point_id)

    # -- LocalBA.py:195-229 -------------------------------------------------------------------
    def motionOnlyBundleAdjustement(self, map, scale=False, save=False):
        frame_ids = list(map.frames.keys())
        point_ids = list(map.points_3d.keys())
        for frame_id in frame_ids:
            frame_obj = map.GetFrame(frame_id)
            self.add_pose(pose_id=frame_id, pose=frame_obj.GetPose(), fixed=bool(frame_obj.IsKeyFrame()))
        self._add_points_and_edges(map, frame_ids, point_ids, fixed=True)
        self.optimize()
        median_depth = 1
        if scale:
            norms = [np.linalg.norm(self.get_point(p)) for p in point_ids]
            median_depth = np.median(np.array(norms))
        for frame_id in frame_ids:
            new_pose = self.get_pose(frame_id).matrix()
            new_pose[0:3, 3] /= median_depth
            map.UpdatePose(new_pose=new_pose, frame_id=frame_id)

    def _add_points_and_edges(self, map, frame_ids, point_ids, fixed):
        fset = set(frame_ids)
        e_pt, e_pose, e_uv = self._e_pt, self._e_pose, self._e_uv
        for point_id in point_ids:
            point_obj = map.GetPoint(point_id)
            self.add_point(point_id=point_id, point=point_obj.Get3dPoint(), fixed=fixed)
            # the reference probes every frame per point (LocalBA.py:168-172); Point.frames is the
            # dict it probes (point.py:14-46), so walking it directly yields the same edge set (edge ids,
            # point_id * frame_id + 10000000 there, are not used by the solver)
            frames = getattr(point_obj, "frames", None)
            if frames is None:
                for frame_id in frame_ids:
                    corr = point_obj.GetFrame(frame_id)
                    if corr is not None:
                        self.add_edge(point_id=point_id, pose_id=frame_id, measurement=corr[1])
                continue
            for frame_id, corr in frames.items():
                if frame_id in fset and corr is not None:
                    e_pt.append(point_id)
                    e_pose.append(frame_id)
                    e_uv.append(corr[1])
