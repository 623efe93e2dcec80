"""Host-side mirror of the anim_utils Skeleton / SkeletonNode interface, as far as
the hot path reads it (field names: reference `utilities/__init__.py:13-16`,
`construction/cluster_tree_builder.py:82-91`; frame index rule
`mg_state_planner.py:288`).  Per-candidate forward kinematics runs on the GPU
(csrc/mgb_score.cu, mgb_staged.cu); the single-frame `get_global_matrix` kept
here serves the once-per-step host setup the reference also does on the host
(aligning transform of the previous frame, `mg_state_planner.py:92-107`).
"""
import collections

import numpy as np

SKELETON_NODE_TYPE_ROOT = 0
SKELETON_NODE_TYPE_JOINT = 1
SKELETON_NODE_TYPE_END_SITE = 2
_EPS = np.finfo(float).eps * 4.0


def quaternion_matrix3(q):
    """3x3 rotation of a (w,x,y,z) quaternion, scale invariant, identity below
    |q|^2 < 4 eps (transformations.quaternion_matrix semantics)."""
    q = np.array(q, dtype=np.float64)
    n = np.dot(q, q)
    if n < _EPS:
        return np.identity(3)
    q *= np.sqrt(2.0 / n)
    o = np.outer(q, q)
    return np.array([[1.0 - o[2, 2] - o[3, 3], o[1, 2] - o[3, 0], o[1, 3] + o[2, 0]],
                     [o[1, 2] + o[3, 0], 1.0 - o[1, 1] - o[3, 3], o[2, 3] - o[1, 0]],
                     [o[1, 3] - o[2, 0], o[2, 3] + o[1, 0], 1.0 - o[1, 1] - o[2, 2]]])


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

    def get_local_matrix(self, quaternion_frame):
        m = np.identity(4)
        if self.node_type == SKELETON_NODE_TYPE_ROOT:
            m[:3, :3] = quaternion_matrix3(quaternion_frame[3:7])
            m[:3, 3] = np.asarray(quaternion_frame[:3], dtype=np.float64) + np.asarray(self.offset, dtype=np.float64)
            return m
        if self.node_type != SKELETON_NODE_TYPE_END_SITE:
            if not self.fixed:
                i = self.quaternion_frame_index * 4 + 3
                m[:3, :3] = quaternion_matrix3(quaternion_frame[i:i + 4])
            else:
                m[:3, :3] = quaternion_matrix3(self.rotation)
        m[:3, 3] = self.offset
        return m

    def get_global_matrix(self, quaternion_frame, use_cache=False):
        if self.parent is not None:
            return np.dot(self.parent.get_global_matrix(quaternion_frame), self.get_local_matrix(quaternion_frame))
        return self.get_local_matrix(quaternion_frame)

    def get_global_position(self, quaternion_frame, use_cache=False):
        return self.get_global_matrix(quaternion_frame)[:3, 3].copy()


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

    @classmethod
    def from_desc(cls, desc):
        """Build from `synthetic.make_skeleton_desc()` output (or any dict of that shape)."""
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

    def get_joint_weights(self):
        return list(self.joint_weight_map.values())


def heading_degrees(skeleton, frame):
    """Heading of the aligning root node in `frame`, degrees from +z towards +x
    (anim_utils get_global_node_orientation_vector / get_rotation_angle convention
    frozen by oracle/standin; in-tree use `mg_state_planner.py:92-98`)."""
    m = skeleton.nodes[skeleton.aligning_root_node].get_global_matrix(frame)[:3, :3]
    p = np.dot(m, np.asarray(skeleton.aligning_root_dir, dtype=np.float64))
    return float(np.degrees(np.arctan2(p[0], p[2])))


def align_frames(skeleton, new_frames, prev_frames):
    """The chosen motion brought into the frame of the walk so far - anim_utils `align_quaternion_frames` with
    `prev_frames` given (call sites motion_primitive_constraints.py:113, graph_walk.py; semantics SURVEY.md Appendix A,
    in-tree near-copy mg_state_planner.py:92-107): rotate about +y by the heading difference between the last previous
    frame and the first new frame, then translate in x / z so that the first new root lands on the last previous root.
    The same transform the device applies to every candidate inside the score kernels (OP_ALIGN)."""
    new_frames = np.array(new_frames, dtype=np.float64)
    last = np.asarray(prev_frames[-1], dtype=np.float64)
    ang = np.radians(heading_degrees(skeleton, last) - heading_degrees(skeleton, new_frames[0]))
    ang = (ang + np.pi) % (2.0 * np.pi) - np.pi
    c, s = np.cos(ang), np.sin(ang)
    x, z = new_frames[:, 0].copy(), new_frames[:, 2].copy()
    rx, rz = c * x + s * z, -s * x + c * z
    new_frames[:, 0] = rx + (last[0] - rx[0])
    new_frames[:, 2] = rz + (last[2] - rz[0])
    qw, qy = np.cos(0.5 * ang), np.sin(0.5 * ang)      # quaternion_about_axis(ang, +y) (x) root quaternion
    w0, x0, y0, z0 = (new_frames[:, 3 + i].copy() for i in range(4))
    new_frames[:, 3] = qw * w0 - qy * y0
    new_frames[:, 4] = qw * x0 + qy * z0
    new_frames[:, 5] = qw * y0 + qy * w0
    new_frames[:, 6] = qw * z0 - qy * x0
    return new_frames
