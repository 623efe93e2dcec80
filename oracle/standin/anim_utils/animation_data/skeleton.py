"""Skeleton / SkeletonNode stand-in: forward kinematics over a joint tree.

Semantics: global = parent_global . local;
root local = [R(frame[3:7]) | frame[:3] + offset]; animated joint i local =
[R(frame[3+4i : 7+4i]) | offset]; fixed joint uses its stored rotation;
end site = [I | offset].  Index rule: reference mg_state_planner.py:288;
node fields: reference utilities/__init__.py:13-16.
"""
import collections

import numpy as np
from transformations import quaternion_matrix, quaternion_from_matrix

SKELETON_NODE_TYPE_ROOT = 0
SKELETON_NODE_TYPE_JOINT = 1
SKELETON_NODE_TYPE_END_SITE = 2
ORIGIN = np.array([0.0, 0.0, 0.0, 1.0])


class SkeletonNode(object):
    def __init__(self, node_name, parent=None, node_type=SKELETON_NODE_TYPE_JOINT):
        self.node_name = node_name
        self.parent = parent
        self.children = []
        self.node_type = node_type
        self.offset = [0.0, 0.0, 0.0]
        self.rotation = np.array([1.0, 0.0, 0.0, 0.0])
        self.fixed = True
        self.quaternion_frame_index = -1
        self.index = -1
        self.cached_global_matrix = None

    def get_local_matrix(self, quaternion_frame):
        if self.node_type == SKELETON_NODE_TYPE_ROOT:
            m = quaternion_matrix(quaternion_frame[3:7])
            m[:3, 3] = [t + o for t, o in zip(quaternion_frame[:3], self.offset)]
            return m
        if self.node_type == SKELETON_NODE_TYPE_END_SITE:
            m = np.identity(4)
            m[:3, 3] = self.offset
            return m
        if not self.fixed:
            i = self.quaternion_frame_index * 4 + 3
            m = quaternion_matrix(quaternion_frame[i:i + 4])
        else:
            m = quaternion_matrix(self.rotation)
        m[:3, 3] = self.offset
        return m

    def get_global_matrix(self, quaternion_frame, use_cache=False):
        if self.cached_global_matrix is not None and use_cache:
            return self.cached_global_matrix
        if self.parent is not None:
            parent_matrix = self.parent.get_global_matrix(quaternion_frame, use_cache)
            self.cached_global_matrix = np.dot(parent_matrix, self.get_local_matrix(quaternion_frame))
        else:
            self.cached_global_matrix = self.get_local_matrix(quaternion_frame)
        return self.cached_global_matrix

    def get_global_position(self, quaternion_frame, use_cache=False):
        return np.dot(self.get_global_matrix(quaternion_frame, use_cache), ORIGIN)[:3]

    def get_global_orientation_quaternion(self, quaternion_frame, use_cache=False):
        return quaternion_from_matrix(self.get_global_matrix(quaternion_frame, use_cache))


class Skeleton(object):
    def __init__(self):
        self.nodes = collections.OrderedDict()
        self.root = None
        self.animated_joints = []
        self.aligning_root_node = None
        self.aligning_root_dir = [0.0, 0.0, 1.0]
        self.head_joint = "Head"
        self.joint_weight_map = collections.OrderedDict()
        self.node_name_frame_map = collections.OrderedDict()
        self.skeleton_model = {"joints": {}}
        self.free_joints_map = dict()

    @classmethod
    def from_desc(cls, desc):
        """Build from morphablegraphs_b200.synthetic.make_skeleton_desc() output."""
        sk = cls()
        type_map = {"root": SKELETON_NODE_TYPE_ROOT, "joint": SKELETON_NODE_TYPE_JOINT,
                    "end_site": SKELETON_NODE_TYPE_END_SITE}
        for idx, (name, nd) in enumerate(desc["nodes"].items()):
            node = SkeletonNode(name, None, type_map[nd["node_type"]])
            node.offset = list(nd["offset"])
            node.rotation = np.array(nd["rotation"], dtype=np.float64)
            node.fixed = bool(nd["fixed"])
            node.quaternion_frame_index = int(nd["quaternion_frame_index"])
            node.index = idx
            sk.nodes[name] = node
        for name, nd in desc["nodes"].items():
            if nd["parent"] is not None:
                sk.nodes[name].parent = sk.nodes[nd["parent"]]
                sk.nodes[nd["parent"]].children.append(sk.nodes[name])
        sk.root = desc["root"]
        sk.animated_joints = list(desc["animated_joints"])
        sk.aligning_root_node = desc["aligning_root_node"]
        sk.aligning_root_dir = list(desc["aligning_root_dir"])
        sk.head_joint = desc["head_joint"]
        sk.joint_weight_map = collections.OrderedDict(desc["joint_weight_map"])
        for i, name in enumerate(sk.animated_joints):
            sk.node_name_frame_map[name] = i
        return sk

    def clear_cached_global_matrices(self):
        for node in self.nodes.values():
            node.cached_global_matrix = None

    def convert_quaternion_frame_to_cartesian_frame(self, quat_frame, node_names=None):
        if node_names is None:
            node_names = list(self.node_name_frame_map.keys())
        return [self.nodes[n].get_global_position(quat_frame) for n in node_names]

    def get_cartesian_coordinates_from_quaternion(self, node_name, quaternion_frame):
        # anim_utils Skeleton method of that name (call site: trajectory_constraint.py:106)
        return self.nodes[node_name].get_global_position(quaternion_frame)

    def get_joint_weights(self):
        return list(self.joint_weight_map.values())

    def get_root_reference_orientation(self):
        return np.array([1.0, 0.0, 0.0, 0.0])
