"""Host containers for the C1 replay tests: the behaviour of the reference's Map / Point / Frame (src/v2/map.py,
point.py, frame.py:51-125) as far as the tracking loop and LocalBA.py use it.  The containers are out of scope
for the product ("host data model"); the tests need them to hold the state between hot-path calls."""
import copy

import numpy as np


class Frame:
    """frame.py:51-125 without image loading: key points / descriptors come from a feature source."""

    def __init__(self, source, index, id):
        self._source, self.index, self.ID = source, index, id
        self.keypoints = self.features = self.pose = None
        self.parents, self.keyframe = {}, False

    def process_frame(self, feature_extractor=None):
        self.keypoints, self.features = self._source(self.index)
        return self.keypoints, self.features, None

    def ClearParent(self): self.parents = {}
    def AddParent(self, parent_frame_id, transition): self.parents[parent_frame_id] = transition
    def GetParentIDs(self): return self.parents.keys()
    def GetTransitionWithParentID(self, parent_id): return self.parents[parent_id]
    def AddPose(self, init_pose): self.pose = init_pose
    def UpdatePose(self, new_pose): self.pose = new_pose
    def GetPose(self): return self.pose
    def SetAsKeyFrame(self): self.keyframe = True
    def IsKeyFrame(self): return self.keyframe
    def GetKeyPoints(self): return self.keypoints
    def GetFeatures(self): return self.features
    def GetID(self): return self.ID
    def AddID(self, new_id): self.ID = new_id


class Point:
    """point.py: a 3-D location and {frame id: (frame, uv, descriptor)} -- one observation per frame."""

    def __init__(self, location, id):
        self.ID, self.location_3d, self.frames = id, location, {}

    def GetID(self): return self.ID
    def GetFrame(self, frame_id): return self.frames.get(frame_id)
    def AddFrame(self, frame, uv, descriptor): self.frames[frame.GetID()] = (frame, uv, descriptor)
    def UpdatePoint(self, new_location): self.location_3d = new_location
    def Get3dPoint(self): return self.location_3d
    def GetNVisibleFrames(self): return len(self.frames)

    def IsVisibleTo(self, frame_id):
        return any(fr.ID == frame_id for fr, _, _ in self.frames.values())      # point.py:37-41 compares frame.ID

    def GetImagePoint(self, frame_id):
        c = self.frames.get(frame_id)
        return None if c is None else (c[1], c[2])


class Map:
    """map.py:6-131."""

    def __init__(self):
        self.frames, self.points_3d = {}, {}

    def AddFrame(self, frame_id, frame):
        if frame_id in self.frames:
            raise Exception("Duplicate frame warning")
        self.frames[frame_id] = frame

    def AddPoint3D(self, point_id, point_3d):
        if point_id in self.points_3d:
            raise Exception("Duplicate point3d warning")
        self.points_3d[point_id] = point_3d

    def GetFrame(self, frame_id): return self.frames[frame_id]
    def GetPoint(self, point_id): return self.points_3d[point_id]

    def UpdatePose(self, new_pose, frame_id):
        if frame_id not in self.frames:
            raise Exception("No frame yet added")
        self.frames[frame_id].UpdatePose(new_pose)

    def UpdatePoint3D(self, new_point, point_id):
        if point_id not in self.points_3d:
            raise Exception("No point yet added")
        self.points_3d[point_id].UpdatePoint(new_point)

    def GetImagePointsWithFrameID(self, frame_id):
        rows = [(pt.GetImagePoint(frame_id), pt) for pt in self.points_3d.values()]
        rows = [(c, pt) for c, pt in rows if c is not None]
        return (np.array([c[0] for c, _ in rows]), np.array([c[1] for c, _ in rows]),
                np.array([pt.Get3dPoint() for _, pt in rows]), np.array([pt.GetID() for _, pt in rows]))

    def GetAll3DPoints(self):
        return np.array([pt.location_3d for pt in self.points_3d.values()]).reshape(-1, 3)

    def GetCopyOfPointObjects(self, frame_id):
        out = {}
        for key, pt in self.points_3d.items():
            if pt.IsVisibleTo(frame_id):
                cp = copy.deepcopy(pt)
                cp.frames = {frame_id: cp.frames[frame_id]}
                out[key] = cp
        return out

    def Store3DPoints(self, points_dict): self.points_3d = {**self.points_3d, **points_dict}

    def AddParentAndPose(self, parent_id, frame_id, frame_obj, rel_pose_trans, pose):
        frame_obj.AddParent(parent_frame_id=parent_id, transition=rel_pose_trans)
        frame_obj.AddPose(init_pose=pose)
        frame_obj.AddID(frame_id)
        self.AddFrame(frame_id=frame_id, frame=frame_obj)

    def AddPointToFrameCorrespondences(self, point_ids, image_points, descriptors, frame_obj):
        for pid, uv, desc in zip(point_ids, image_points, descriptors):
            self.GetPoint(pid).AddFrame(frame_obj, uv, desc)

    def DiscardOutlierMapPoints(self, n_visible_frames=3):
        self.points_3d = {k: p for k, p in self.points_3d.items() if p.GetNVisibleFrames() >= n_visible_frames}
