"""Minimal stand-ins for the reference's Map / Frame / Point containers (src/v2/map.py, frame.py,
point.py), as far as LocalBA.py touches them."""
import numpy as np


class _Frame:
    def __init__(self, fid, pose, kf=False):
        self.ID, self.pose, self.kf, self.parents = fid, pose, kf, {}
    def GetPose(self): return self.pose
    def GetParentIDs(self): return self.parents.keys()
    def GetTransitionWithParentID(self, pid): return self.parents[pid]
    def IsKeyFrame(self): return self.kf
    def GetID(self): return self.ID


class _Point:
    def __init__(self, pid, xyz):
        self.ID, self.location_3d, self.frames = pid, xyz, {}
    def GetFrame(self, fid): return self.frames.get(fid)
    def Get3dPoint(self): return self.location_3d
    def GetID(self): return self.ID


class _Map:
    def __init__(self): self.frames, self.points_3d = {}, {}
    def GetFrame(self, fid): return self.frames[fid]
    def GetPoint(self, pid): return self.points_3d[pid]
    def UpdatePose(self, new_pose, frame_id): self.frames[frame_id].pose = new_pose
    def UpdatePoint3D(self, new_point, point_id): self.points_3d[point_id].location_3d = new_point


class _Cam:
    fx, fy, cx, cy = 481.2, 480.0, 319.5, 239.5


def _map_from_problem(p):
    from visual_slam_b200.ba_problem import quat_to_R
    m = _Map()
    for c in range(p.n_cams):
        T = np.eye(4); T[:3, :3] = quat_to_R(p.cams[c, :4]); T[:3, 3] = p.cams[c, 4:]
        m.frames[c] = _Frame(c, T)
    for k in range(p.n_scale):
        a, b = int(p.scale_a[k]), int(p.scale_b[k])
        rel = np.eye(4); rel[:3, 3] = [p.scale_meas[k], 0, 0]
        m.frames[b].parents[a] = rel
    for j in range(p.n_pts):
        m.points_3d[j] = _Point(j, p.pts[j].copy())
    for o in range(p.n_obs):
        m.points_3d[int(p.obs_pt[o])].frames[int(p.obs_cam[o])] = (m.frames[int(p.obs_cam[o])], p.obs_uv[o], None)
    return m
