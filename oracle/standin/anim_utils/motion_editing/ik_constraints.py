"""Only the names are needed at import time (reference motion_primitive_constraints.py:34)."""


class JointIKConstraint(object):
    def __init__(self, joint_name, position, orientation, keyframe, free_joints, step_idx=-1, frame_range=None, look_at=False):
        self.joint_name = joint_name
        self.position = position
        self.orientation = orientation
        self.keyframe = keyframe
        self.free_joints = free_joints
        self.step_idx = step_idx


class TwoJointIKConstraint(object):
    def __init__(self, *args, **kwargs):
        self.args = args
