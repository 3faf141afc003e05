"""A stand-in `g2o` module so that the reference's OWN BA wrapper runs unmodified.  TEST INFRASTRUCTURE ONLY.

The reference drives bundle adjustment through the `g2o` (g2opy) wheel, which is absent from /root/reference
and not installable here.  Its wrapper, /root/reference/src/v2/LocalBA.py, cannot even be
imported without it (`class BundleAdjustment(g2o.SparseOptimizer)`, LocalBA.py:20).  `install()` registers this
module as `sys.modules["g2o"]`: it offers the subset of the g2opy API that LocalBA.py:20-229 and map.py:3 touch
and forwards `SparseOptimizer.optimize` to the CPU restatement in oracle/ba.py.

What that buys: the graph the reference builds from its Map (which frames are fixed, which points, which edges,
scale-edge measurements, vertex ids, the median-depth rescale and the write-back of LocalBA.py:143-229) is the
reference's own code, executed line by line; only the arithmetic behind `optimize()` is the restatement.
PARITY of that arithmetic stays UNPINNED (see oracle/ba.py).

API covered (g2opy names): SparseOptimizer(add_vertex, add_edge, vertex, set_algorithm,
initialize_optimization, optimize, set_verbose, save), BlockSolverSE3, LinearSolverCholmodSE3,
OptimizationAlgorithmLevenberg, CameraParameters, Isometry3d, Quaternion, SE3Quat, SBACam, VertexCam,
VertexSBAPointXYZ, EdgeProjectP2MC, EdgeSBAScale, EdgeSE3, RobustKernelHuber, RobustKernelDCS.
"""
import sys
import types

import numpy as np

from . import ba as _ba


class Quaternion:
    """Eigen::Quaterniond built from a rotation matrix (Eigen's conversion; NOT normalised here)."""

    def __init__(self, *a):
        if len(a) == 1:
            R = np.asarray(a[0], dtype=np.float64)
            tr = R[0, 0] + R[1, 1] + R[2, 2]
            if tr > 0:
                s = np.sqrt(tr + 1.0)
                w = 0.5 * s
                s = 0.5 / s
                x, y, z = (R[2, 1] - R[1, 2]) * s, (R[0, 2] - R[2, 0]) * s, (R[1, 0] - R[0, 1]) * s
            else:
                i = 0
                if R[1, 1] > R[0, 0]:
                    i = 1
                if R[2, 2] > R[i, i]:
                    i = 2
                j, k = (i + 1) % 3, (i + 2) % 3
                s = np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0)
                v = [0.0, 0.0, 0.0]
                v[i] = 0.5 * s
                s = 0.5 / s
                w = (R[k, j] - R[j, k]) * s
                v[j] = (R[j, i] + R[i, j]) * s
                v[k] = (R[k, i] + R[i, k]) * s
                x, y, z = v
            self.q = np.array([x, y, z, w])
        else:
            w, x, y, z = [float(v) for v in a]
            self.q = np.array([x, y, z, w])

    def coeffs(self):
        return self.q.copy()          # (x, y, z, w), Eigen order

    x = lambda s: s.q[0]  # noqa: E731
    y = lambda s: s.q[1]  # noqa: E731
    z = lambda s: s.q[2]  # noqa: E731
    w = lambda s: s.q[3]  # noqa: E731


class Isometry3d:
    def __init__(self, *a):
        if len(a) == 1:
            self.T = np.array(a[0], dtype=np.float64).reshape(4, 4)
        elif len(a) == 2:
            self.T = np.eye(4)
            self.T[:3, :3] = np.asarray(a[0], dtype=np.float64)
            self.T[:3, 3] = np.asarray(a[1], dtype=np.float64).reshape(3)
        else:
            self.T = np.eye(4)

    def orientation(self):
        return Quaternion(self.T[:3, :3])

    rotation = orientation

    def position(self):
        return self.T[:3, 3].copy()

    translation = position

    def matrix(self):
        return self.T.copy()

    def inverse(self):
        R, t = self.T[:3, :3], self.T[:3, 3]
        return Isometry3d(R.T, -R.T @ t)


class SE3Quat:
    """SE3Quat(q, t): the constructor normalises q and makes w >= 0 (g2o se3quat.h)."""

    def __init__(self, q=None, t=None):
        c = np.array([0.0, 0.0, 0.0, 1.0]) if q is None else (q.coeffs() if isinstance(q, Quaternion) else
                                                              Quaternion(q).coeffs())
        if c[3] < 0:
            c = -c
        self.q = c / np.linalg.norm(c)
        self.t = np.zeros(3) if t is None else np.asarray(t, dtype=np.float64).reshape(3).copy()

    def as7(self):
        return np.concatenate([self.q, self.t])

    def matrix(self):
        T = np.eye(4)
        T[:3, :3] = _ba.quat_to_R(self.q[None])[0]
        T[:3, 3] = self.t
        return T

    to_homogeneous_matrix = matrix

    def position(self):
        return self.t.copy()

    translation = position

    def orientation(self):
        o = Quaternion.__new__(Quaternion)
        o.q = self.q.copy()
        return o

    rotation = orientation


class SBACam(SE3Quat):
    def __init__(self, q=None, t=None):
        super().__init__(q, t)
        self.K = None

    def set_cam(self, fx, fy, cx, cy, baseline):
        self.K = (float(fx), float(fy), float(cx), float(cy))
        self.baseline = baseline


class _Vertex:
    def __init__(self):
        self._id, self._est, self._fixed, self._marg = None, None, False, False

    def set_id(self, i):
        self._id = int(i)

    def id(self):
        return self._id

    def set_estimate(self, e):
        self._est = e

    def estimate(self):
        return self._est

    def set_fixed(self, f):
        self._fixed = bool(f)

    def fixed(self):
        return self._fixed

    def set_marginalized(self, m):
        self._marg = bool(m)


class VertexCam(_Vertex):
    pass


class VertexSBAPointXYZ(_Vertex):
    def set_estimate(self, e):
        self._est = np.asarray(e, dtype=np.float64).reshape(3).copy()


class _Edge:
    def __init__(self):
        self._v, self._meas, self._info, self._kernel, self._id = {}, None, None, None, None

    def set_vertex(self, i, v):
        self._v[int(i)] = v

    def set_measurement(self, m):
        self._meas = m

    def set_information(self, info):
        self._info = np.asarray(info, dtype=np.float64)

    def set_id(self, i):
        self._id = i

    def set_robust_kernel(self, k):
        self._kernel = k

    def set_parameter_id(self, a, b):
        pass                                   # no-op for these edge types


class EdgeProjectP2MC(_Edge):
    pass


class EdgeSBAScale(_Edge):
    pass


class EdgeSE3(_Edge):
    pass


class RobustKernelHuber:
    def __init__(self, delta=1.0):
        self.delta = float(delta)


class RobustKernelDCS:
    def __init__(self, delta=1.0):
        self.delta = float(delta)


class LinearSolverCholmodSE3:
    pass


class LinearSolverCSparseSE3:
    pass


class BlockSolverSE3:
    def __init__(self, linear_solver=None):
        self.linear_solver = linear_solver


class OptimizationAlgorithmLevenberg:
    def __init__(self, solver=None):
        self.solver = solver


class CameraParameters:
    def __init__(self, *a):
        self.args = a

    def set_id(self, i):
        self.id = i


class _Problem:
    """The arrays oracle/ba.py works on (same attribute names as the product's BAProblem container, which the
    oracle must not depend on here)."""

    def __init__(self, cams, pts, oc, op, uv, K, cf, pf, sa, sb, sm):
        self.cams, self.pts = cams, pts
        self.obs_cam, self.obs_pt, self.obs_uv = oc, op, uv
        self.K, self.cam_fixed, self.pt_fixed = K, cf, pf
        self.scale_a, self.scale_b, self.scale_meas = sa, sb, sm

    n_cams = property(lambda s: s.cams.shape[0])
    n_pts = property(lambda s: s.pts.shape[0])
    n_obs = property(lambda s: len(s.obs_cam))
    n_scale = property(lambda s: len(s.scale_a))
    n_free_cams = property(lambda s: int((~s.cam_fixed).sum()))
    n_free_pts = property(lambda s: int((~s.pt_fixed).sum()))

    @property
    def cam_col(self):
        col = np.full(self.n_cams, -1, np.int64)
        col[~self.cam_fixed] = np.arange(self.n_free_cams)
        return col

    @property
    def pt_col(self):
        col = np.full(self.n_pts, -1, np.int64)
        col[~self.pt_fixed] = np.arange(self.n_free_pts)
        return col


class SparseOptimizer:
    def __init__(self):
        self._vertices, self._edges = {}, []
        self.algorithm = None
        self.history = None        # filled by optimize(): the restatement's cost / lambda trajectory
        self.problem = None

    def set_algorithm(self, a):
        self.algorithm = a

    def set_verbose(self, v):
        pass

    def add_vertex(self, v):
        if v.id() in self._vertices:
            return False
        self._vertices[v.id()] = v
        return True

    def vertex(self, i):
        return self._vertices.get(int(i))

    def vertices(self):
        return self._vertices

    def add_edge(self, e):
        if any(v is None for v in e._v.values()):
            return False
        self._edges.append(e)
        return True

    def edges(self):
        return self._edges

    def initialize_optimization(self, level=0):
        return True

    def save(self, filename):
        raise NotImplementedError("g2o text export is out of scope (SURVEY 5: checkpoint / resume)")

    def _pack(self):
        cams_v = [v for v in self._vertices.values() if isinstance(v, VertexCam)]
        pts_v = [v for v in self._vertices.values() if isinstance(v, VertexSBAPointXYZ)]
        ci = {v.id(): i for i, v in enumerate(cams_v)}
        pi = {v.id(): i for i, v in enumerate(pts_v)}
        Ks = {v.estimate().K for v in cams_v}
        if len(Ks) > 1:
            raise ValueError("the restatement assumes one set of intrinsics (LocalBA.py:59 uses one)")
        cams = np.array([v.estimate().as7() for v in cams_v]).reshape(-1, 7)
        pts = np.array([v.estimate() for v in pts_v]).reshape(-1, 3)
        oc, op, uv, sa, sb, sm = [], [], [], [], [], []
        for e in self._edges:
            if isinstance(e, EdgeProjectP2MC):
                if not np.array_equal(e._info, np.identity(2)) or not isinstance(e._kernel, RobustKernelHuber) \
                        or abs(e._kernel.delta - _ba.HUBER_DELTA) > 0:
                    raise ValueError("restatement covers information I2 + Huber(sqrt(5.991)) (LocalBA.py:81-82)")
                op.append(pi[e._v[0].id()]); oc.append(ci[e._v[1].id()])
                uv.append(np.asarray(e._meas, dtype=np.float64).reshape(2))
            elif isinstance(e, EdgeSBAScale):
                if not isinstance(e._kernel, RobustKernelDCS) or e._kernel.delta != _ba.DCS_PHI:
                    raise ValueError("restatement covers DCS(1) scale edges (LocalBA.py:115)")
                sa.append(ci[e._v[0].id()]); sb.append(ci[e._v[1].id()]); sm.append(float(e._meas))
            else:
                raise ValueError("edge type outside the src/v2 hot path")
        prob = _Problem(cams, pts, np.array(oc, np.int32), np.array(op, np.int32), np.array(uv).reshape(-1, 2),
                        Ks.pop() if Ks else (1.0, 1.0, 0.0, 0.0),
                        np.array([v.fixed() for v in cams_v], bool), np.array([v.fixed() for v in pts_v], bool),
                        np.array(sa, np.int32), np.array(sb, np.int32), np.array(sm, np.float64))
        return prob, cams_v, pts_v

    def optimize(self, iterations=10, online=False):
        prob, cams_v, pts_v = self._pack()
        self.problem = prob
        self.history = _ba.optimize(prob, iterations)
        for v, c in zip(cams_v, prob.cams):
            est = v.estimate()
            est.q, est.t = c[0:4].copy(), c[4:7].copy()
        for v, p in zip(pts_v, prob.pts):
            v._est = p.copy()
        return self.history["iterations"]


def install():
    """Registers this module as `g2o` (only if the real wheel is absent) and returns it."""
    if "g2o" in sys.modules and not getattr(sys.modules["g2o"], "_IS_VSLAM_SHIM", False):
        return sys.modules["g2o"]
    m = types.ModuleType("g2o")
    m._IS_VSLAM_SHIM = True
    for name, obj in list(globals().items()):
        if isinstance(obj, type) and not name.startswith("_"):
            setattr(m, name, obj)
    sys.modules["g2o"] = m
    return m
