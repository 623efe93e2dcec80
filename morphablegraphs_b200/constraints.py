"""Host-side mirror of the reference's constraint containers for the hot path.

Same class names, constructor signatures `(skeleton, constraint_desc, precision,
weight_factor)` and attribute names as reference
`morphablegraphs/constraints/spatial_constraints/keyframe_constraints/*.py` and
`constraints/motion_primitive_constraints.py:56-77`, so that (a) the parity
tests read like reference code and (b) `engine.pack_constraints` works on either
these or the reference's own objects.  They are data holders: evaluation of a
whole candidate batch happens on the GPU (`engine.DeviceConstraints`).
"""
import numpy as np

SPATIAL_CONSTRAINT_TYPE_TRAJECTORY = "trajectory"
SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSITION = "keyframe_position"
SPATIAL_CONSTRAINT_TYPE_KEYFRAME_DIR_2D = "keyframe_2d_direction"
SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSE = "keyframe_pose"
SPATIAL_CONSTRAINT_TYPE_TWO_HAND_POSITION = "keyframe_two_hands"
SPATIAL_CONSTRAINT_TYPE_TRAJECTORY_SET = "trajectory_set"
SPATIAL_CONSTRAINT_TYPE_KEYFRAME_LOOK_AT = "keyframe_look_at"
SPATIAL_CONSTRAINT_TYPE_KEYFRAME_FEET = "keyframe_feet"
SPATIAL_CONSTRAINT_TYPE_CA_CONSTRAINT = "ca_constraint"
SPATIAL_CONSTRAINT_TYPE_KEYFRAME_RELATIVE_POSITION = "keyframe_relative_position"


class SpatialConstraintBase(object):
    """reference spatial_constraint_base.py:29-35"""

    def __init__(self, precision, weight_factor=1.0):
        self.precision = precision
        self.weight_factor = weight_factor
        self.constraint_type = None


class KeyframeConstraintBase(SpatialConstraintBase):
    """reference keyframe_constraint_base.py:28-40"""

    def __init__(self, constraint_desc, precision, weight_factor=1.0):
        assert "canonical_keyframe" in list(constraint_desc.keys())
        super(KeyframeConstraintBase, self).__init__(precision, weight_factor)
        self.semantic_annotation = constraint_desc["semanticAnnotation"]
        self.keyframe_label = constraint_desc["semanticAnnotation"]["keyframeLabel"]
        self.canonical_keyframe = constraint_desc["canonical_keyframe"]


class GlobalTransformConstraint(KeyframeConstraintBase):
    """reference global_transform_constraint.py:38-64; orientation given as
    `qOrientation` (w,x,y,z).  Euler `orientation` input needs anim_utils'
    euler_to_quaternion and is converted by the caller."""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(GlobalTransformConstraint, self).__init__(constraint_desc, precision, weight_factor)
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSITION
        self.skeleton = skeleton
        self.joint_name = constraint_desc["joint"]
        self.position = constraint_desc["position"] if "position" in constraint_desc else None
        if "qOrientation" in constraint_desc:
            self.orientation = constraint_desc["qOrientation"]
        else:
            self.orientation = None
        self.n_canonical_frames = constraint_desc["n_canonical_frames"]


class RelativeTransformConstraint(GlobalTransformConstraint):
    """reference relative_transform_constraint.py:33-44"""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(RelativeTransformConstraint, self).__init__(skeleton, constraint_desc, precision, weight_factor)
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_KEYFRAME_RELATIVE_POSITION
        self.offset = constraint_desc["offset"]


class GlobalTransformCAConstraint(GlobalTransformConstraint):
    """reference global_transform_ca_constraint.py:28-32"""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0, step_idx=-1):
        super(GlobalTransformCAConstraint, self).__init__(skeleton, constraint_desc, precision, weight_factor)
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_CA_CONSTRAINT
        self.step_idx = step_idx


class Direction2DConstraint(KeyframeConstraintBase):
    """reference direction_2d_constraint.py:32-40"""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(Direction2DConstraint, self).__init__(constraint_desc, precision, weight_factor)
        self.skeleton = skeleton
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_KEYFRAME_DIR_2D
        self.direction_constraint = constraint_desc["dir_vector"]
        self.target_dir = np.array([self.direction_constraint[0], self.direction_constraint[2]], dtype=np.float64)
        self.target_dir /= np.linalg.norm(self.target_dir)
        self.target_dir_len = np.linalg.norm(self.target_dir)


class PoseConstraint(KeyframeConstraintBase):
    """reference pose_constraint.py:34-46"""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(PoseConstraint, self).__init__(constraint_desc, precision, weight_factor)
        self.skeleton = skeleton
        self.pose_constraint = constraint_desc["frame_constraint"]
        self.velocity_constraint = constraint_desc.get("velocity_constraint", None)
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSE
        self.node_names = constraint_desc["node_names"]
        self.weights = constraint_desc["weights"]


class TwoHandConstraintSet(KeyframeConstraintBase):
    """reference two_hand_constraint.py:40-54"""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(TwoHandConstraintSet, self).__init__(constraint_desc, precision, weight_factor)
        self.skeleton = skeleton
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_TWO_HAND_POSITION
        self.positions = constraint_desc["positions"]
        self.orientations = constraint_desc["orientations"]
        self.joint_names = constraint_desc["joint"]
        self.target_direction = np.array(self.positions[1], dtype=np.float64) - np.array(self.positions[0], dtype=np.float64)
        self.target_delta = np.linalg.norm(self.target_direction)
        self.target_center = np.array(self.positions[0], dtype=np.float64) + 0.5 * self.target_direction
        self.target_direction /= self.target_delta
        self.n_canonical_frames = constraint_desc["n_canonical_frames"]


class FeetConstraint(KeyframeConstraintBase):
    """reference feet_constraint.py:30-37"""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(FeetConstraint, self).__init__(constraint_desc, precision, weight_factor)
        self.skeleton = skeleton
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_KEYFRAME_FEET
        self.left = constraint_desc["left"]
        self.right = constraint_desc["right"]


class LookAtConstraint(KeyframeConstraintBase):
    """reference look_at_constraint.py:36-44"""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(LookAtConstraint, self).__init__(constraint_desc, precision, weight_factor)
        self.skeleton = skeleton
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_KEYFRAME_LOOK_AT
        self.joint_name = "Head"
        self.target_position = constraint_desc["position"]


SPATIAL_CONSTRAINT_TYPE_LOCAL_TRAJECTORY = "local_trajectory"


def _quaternion_matrix3(q):
    """3x3 part of transformations.quaternion_matrix (w, x, y, z): identity below 4 eps."""
    q = np.array(q, dtype=np.float64)
    n = float(np.dot(q, q))
    if n < np.finfo(float).eps * 4.0:
        return np.identity(3)
    q *= np.sqrt(2.0 / n)
    o = np.outer(q, q)
    return np.array([[1.0 - o[2, 2] - o[3, 3], o[1, 2] - o[3, 0], o[1, 3] + o[2, 0]],
                     [o[1, 2] + o[3, 0], 1.0 - o[1, 1] - o[3, 3], o[2, 3] - o[1, 0]],
                     [o[1, 3] - o[2, 0], o[2, 3] + o[1, 0], 1.0 - o[1, 1] - o[2, 2]]])


def _euler_matrix3_rxyz(ai, aj, ak):
    """3x3 part of transformations.euler_matrix(ai, aj, ak, axes='rxyz'): rotating frame, x then y then z,
    i.e. R = Rx(ai) Ry(aj) Rz(ak)."""
    ci, si = np.cos(ai), np.sin(ai)
    cj, sj = np.cos(aj), np.sin(aj)
    ck, sk = np.cos(ak), np.sin(ak)
    rx = np.array([[1, 0, 0], [0, ci, -si], [0, si, ci]])
    ry = np.array([[cj, 0, sj], [0, 1, 0], [-sj, 0, cj]])
    rz = np.array([[ck, -sk, 0], [sk, ck, 0], [0, 0, 1]])
    return rx.dot(ry).dot(rz)


class JointRotationConstraint(KeyframeConstraintBase):
    """reference joint_rotation_constraint.py:32-53: the joint's own rotation at `frame_index`
    against a target rotation, Frobenius norm of the matrix difference.  `constraint_rotmat` is 4x4
    like the reference's (transformations returns homogeneous matrices); `constraint_type` stays None
    as in the reference, the engine recognises the class by `constraint_rotmat`."""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(JointRotationConstraint, self).__init__(constraint_desc, precision, weight_factor)
        self.skeleton = skeleton
        self.joint_name = constraint_desc["joint_name"]
        self.rotation_type = constraint_desc["rotation_type"]
        self.rotation_constraint = constraint_desc["rotation_constraint"]
        self.frame_idx = constraint_desc["frame_index"]
        self.constraint_rotmat = np.identity(4)
        if self.rotation_type == "euler":
            rad = [float(np.deg2rad(a)) for a in self.rotation_constraint]
            self.constraint_rotmat[:3, :3] = _euler_matrix3_rxyz(rad[0], rad[1], rad[2])
        elif self.rotation_type == "quaternion":
            quat = np.asarray(self.rotation_constraint, dtype=np.float64)
            quat = quat / np.linalg.norm(quat)
            self.constraint_rotmat[:3, :3] = _quaternion_matrix3(quat)
        else:
            raise ValueError("Unknown rotation type!")


class PoseConstraintQuatFrame(KeyframeConstraintBase):
    """reference pose_constraint_quat_frame.py:30-38: `frame_constraint` is a whole quaternion frame
    (root translation + one quaternion per animated joint), compared at frame 0 joint by joint (root
    excluded) with the skeleton's joint weights."""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(PoseConstraintQuatFrame, self).__init__(constraint_desc, precision, weight_factor)
        self.skeleton = skeleton
        self.pose_constraint = constraint_desc["frame_constraint"]
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSE


class LocalTrajectoryConstraint(KeyframeConstraintBase):
    """reference local_trajectory_constraint.py:32-41.  `trajectory` is any object with the reference's
    ParameterizedSpline interface (`query_point_by_absolute_arc_length`, `full_arc_length`,
    `get_last_control_point`), e.g. the reference's own spline or an `ArcLengthTable`."""

    def __init__(self, skeleton, constraint_desc, precision, weight_factor=1.0):
        super(LocalTrajectoryConstraint, self).__init__(constraint_desc, precision, weight_factor)
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_LOCAL_TRAJECTORY
        self.trajectory = constraint_desc["trajectory"]
        self.start_t = constraint_desc["start_t"]
        self.n_canonical_frames = constraint_desc["n_canonical_frames"]
        self.joint_name = constraint_desc["joint_name"]
        self.skeleton = skeleton


class ArcLengthTable(object):
    """A trajectory tabulated by arc length: `points[i]` = the point at arc length `i * step`,
    `(len(points) - 1) * step` = the full arc length, linear in between, `last_control_point`
    beyond the end (what parameterized_spline.py:131-148 returns there).  This is the form in which
    the engine ships ANY trajectory object to the device (`tabulate_trajectory`); constructing one
    directly skips the per-point host queries."""

    N_DEFAULT = 2048

    def __init__(self, points, step, last_control_point=None):
        self.points = np.ascontiguousarray(points, dtype=np.float64)
        self.step = float(step)
        self.full_arc_length = self.step * (len(self.points) - 1)
        self.last_control_point = np.asarray(self.points[-1] if last_control_point is None else last_control_point, dtype=np.float64)

    def get_last_control_point(self):
        return self.last_control_point

    def query_point_by_absolute_arc_length(self, t):
        if t > self.full_arc_length:
            return self.last_control_point
        u = max(float(t), 0.0) / self.step
        i = min(int(u), len(self.points) - 2)
        f = u - i
        return self.points[i] + f * (self.points[i + 1] - self.points[i])


def tabulate_trajectory(trajectory, n=ArcLengthTable.N_DEFAULT):
    """ArcLengthTable of a trajectory object (cached on the object: the table of a planner's path is
    built once, not once per motion-primitive step)."""
    if isinstance(trajectory, ArcLengthTable):
        return trajectory
    cached = getattr(trajectory, "_mgb_arc_table", None)
    full = float(trajectory.full_arc_length)
    if cached is not None and cached[0] == (full, n):
        return cached[1]
    step = full / n
    pts = np.array([np.asarray(trajectory.query_point_by_absolute_arc_length(min(i * step, full)), dtype=np.float64)[:3] for i in range(n + 1)])
    last = trajectory.get_last_control_point() if hasattr(trajectory, "get_last_control_point") else trajectory.spline.get_last_control_point()
    table = ArcLengthTable(pts, step, np.asarray(last, dtype=np.float64)[:3])
    try:
        trajectory._mgb_arc_table = ((full, n), table)
    except AttributeError:
        pass
    return table


class ParameterTable(object):
    """A curve tabulated by its own parameter: `points[i]` = the curve at u = i / (len(points) - 1), linear in
    between.  The form in which the engine ships the spline of a TrajectoryConstraint to the device."""

    N_DEFAULT = 2048

    def __init__(self, points):
        self.points = np.ascontiguousarray(points, dtype=np.float64)

    def query_point_by_parameter(self, u):
        x = min(max(float(u), 0.0), 1.0) * (len(self.points) - 1)
        i = min(int(x), len(self.points) - 2)
        return self.points[i] + (x - i) * (self.points[i + 1] - self.points[i])


def tabulate_by_parameter(trajectory, n=ParameterTable.N_DEFAULT):
    """ParameterTable of an object with the reference's `query_point_by_parameter(u)` (parameterized_spline.py:129-130),
    cached on the object like `tabulate_trajectory`."""
    if isinstance(trajectory, ParameterTable):
        return trajectory
    cached = getattr(trajectory, "_mgb_param_table", None)
    key = (float(getattr(trajectory, "full_arc_length", 0.0)), n)
    if cached is not None and cached[0] == key:
        return cached[1]
    pts = np.array([np.asarray(trajectory.query_point_by_parameter(i / float(n)), dtype=np.float64).reshape(-1)[:3] for i in range(n + 1)])
    table = ParameterTable(pts)
    try:
        trajectory._mgb_param_table = (key, table)
    except AttributeError:
        pass
    return table


class CatmullRomTrajectory(object):
    """A Catmull-Rom curve through `control_points` with an arc-length map over `granularity` equal parameter steps: the
    queries of the reference's ParameterizedSpline over its CatmullRomSpline that the trajectory constraints use
    (splines/catmull_rom_spline.py:120-182: uniform segments, end points doubled, the four neighbouring control points
    blended by the 0.5-scaled Catmull-Rom basis; splines/arc_length_map.py:45-70 and :86-99: accumulated chord lengths at
    the steps, inverted by linear interpolation between the bracketing steps; parameterized_spline.py:131-156)."""

    BASIS = 0.5 * np.array([[-1.0, 3.0, -3.0, 1.0], [2.0, -5.0, 4.0, -1.0], [-1.0, 0.0, 1.0, 0.0], [0.0, 2.0, 0.0, 0.0]])

    def __init__(self, control_points, granularity=1000):
        cp = [np.asarray(p, dtype=np.float64) for p in control_points]
        if len(cp) < 2:
            raise ValueError("Not enough control points in trajectory constraint definition")
        self.number_of_segments = len(cp) - 1
        self.control_points = np.array([cp[0]] + cp + [cp[-1], cp[-1]])
        self.granularity = int(granularity)
        u = np.arange(self.granularity + 1) / float(self.granularity)
        pts = np.array([self.query_point_by_parameter(x) for x in u])
        acc = np.concatenate([[0.0], np.cumsum(np.linalg.norm(np.diff(pts, axis=0), axis=1))])
        self.full_arc_length = float(acc[-1])
        if self.full_arc_length == 0:
            raise ValueError("Not enough control points in trajectory constraint definition")
        self._u = u
        self._rel = acc / self.full_arc_length

    def get_last_control_point(self):
        return np.array(self.control_points[-1])

    def query_point_by_parameter(self, u):
        x = self.number_of_segments * float(u)
        i = min(int(np.floor(x)), self.number_of_segments)
        if i + 1 > self.number_of_segments:
            return np.array(self.control_points[-1])
        t = x - i
        return np.array([t ** 3, t ** 2, t, 1.0]).dot(self.BASIS).dot(self.control_points[i:i + 4])

    def map_relative_arc_length_to_parameter(self, rel):
        if rel <= self._rel[0]:
            return self._u[0]
        if rel >= self._rel[-1]:
            return self._u[-1]
        i = int(np.searchsorted(self._rel, rel, side="right")) - 1     # last step with arc length <= rel
        if self._rel[i] == rel:
            return self._u[i]
        return self._u[i] + (rel - self._rel[i]) / (self._rel[i + 1] - self._rel[i]) * (self._u[i + 1] - self._u[i])

    def query_point_by_absolute_arc_length(self, absolute_arc_length):
        if absolute_arc_length <= self.full_arc_length:
            return self.query_point_by_parameter(self.map_relative_arc_length_to_parameter(absolute_arc_length / self.full_arc_length))
        return self.get_last_control_point()


class TrajectoryConstraint(CatmullRomTrajectory, SpatialConstraintBase):
    """reference trajectory_constraint.py:33-53 (constructor signature and attribute names).  Only the Catmull-Rom spline
    type (0, the reference's default) is restated; the orientation spline is not part of the error and is not kept.
    The engine reads `joint_name`, `min_arc_length`, `full_arc_length`, `query_point_by_parameter` and `weight_factor`,
    so the reference's own object is accepted as well."""

    def __init__(self, joint_name, control_points, orientations, spline_type, min_arc_length, unconstrained_indices, skeleton, precision,
                 weight_factor=1.0, closest_point_search_accuracy=0.001, closest_point_search_max_iterations=5000, granularity=1000):
        if spline_type not in (0, None):
            raise NotImplementedError("only Catmull-Rom trajectories (spline_type 0) are restated")
        CatmullRomTrajectory.__init__(self, control_points, granularity)
        SpatialConstraintBase.__init__(self, precision, weight_factor)
        self.semantic_annotation = None
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_TRAJECTORY
        self.joint_name = joint_name
        self.skeleton = skeleton
        self.min_arc_length = min_arc_length
        self.n_canonical_frames = 0
        self.arc_length = 0.0
        self.unconstrained_indices = unconstrained_indices
        self.range_start = None
        self.range_end = None
        self.is_collision_avoidance_constraint = False

    def set_active_range(self, new_range_start, new_range_end):
        self.range_start = new_range_start
        self.range_end = new_range_end

    def set_number_of_canonical_frames(self, n_canonical_frames):
        self.n_canonical_frames = n_canonical_frames

    def is_active(self, arc_length):
        return self.range_start is not None and self.range_start <= arc_length <= self.range_end

    def get_length_of_residual_vector(self):
        return self.n_canonical_frames


class TrajectorySetConstraint(SpatialConstraintBase):
    """reference trajectory_set_constraint.py:29-40, :55-56: several joint trajectories (objects with the TrajectoryConstraint
    interface: `query_point_by_absolute_arc_length`, `full_arc_length`, `get_last_control_point`, `range_start`,
    `range_end`) walked by the arc length of each joint's own path, starting at `joint_arc_lengths`."""

    def __init__(self, joint_trajectories, joint_names, skeleton, precision, weight_factor):
        super(TrajectorySetConstraint, self).__init__(precision, weight_factor)
        self.skeleton = skeleton
        self.constraint_type = SPATIAL_CONSTRAINT_TYPE_TRAJECTORY_SET
        self.joint_trajectories = joint_trajectories
        self.joint_names = joint_names
        self.joint_arc_lengths = np.zeros(len(self.joint_trajectories))
        self.semantic_annotation = dict()
        self.semantic_annotation["generated"] = True
        self.n_canonical_frames = 0

    def set_number_of_canonical_frames(self, n_canonical_frames):
        self.n_canonical_frames = n_canonical_frames

    def get_length_of_residual_vector(self):
        return self.n_canonical_frames


CONSTRAINT_FACTORY = {
    "TrajectoryConstraint": TrajectoryConstraint,
    "TrajectorySetConstraint": TrajectorySetConstraint,
    "JointRotationConstraint": JointRotationConstraint,
    "PoseConstraintQuatFrame": PoseConstraintQuatFrame,
    "LocalTrajectoryConstraint": LocalTrajectoryConstraint,
    "GlobalTransformConstraint": GlobalTransformConstraint,
    "RelativeTransformConstraint": RelativeTransformConstraint,
    "Direction2DConstraint": Direction2DConstraint,
    "PoseConstraint": PoseConstraint,
    "TwoHandConstraintSet": TwoHandConstraintSet,
    "FeetConstraint": FeetConstraint,
    "LookAtConstraint": LookAtConstraint,
    "GlobalTransformCAConstraint": GlobalTransformCAConstraint,
}


class MotionPrimitiveConstraints(object):
    """reference motion_primitive_constraints.py:47-77 (fields the hot path reads
    and writes: constraints, is_local, skeleton, start_pose, aligning_transform,
    min_error, evaluations, time, use_local_optimization)."""

    def __init__(self):
        self.action_name = None
        self.motion_primitive_name = None
        self.settings = None
        self.constraints = list()
        self.use_local_optimization = False
        self.start_pose = None
        self.skeleton = None
        self.precision = {"pos": 1.0, "rot": 1.0, "smooth": 1.0}
        self.verbose = False
        self.min_error = 0.0
        self.best_parameters = None
        self.evaluations = 0
        self.aligning_transform = None
        self.is_local = False
        self.is_last_step = False
        self.time = 0.0

    def transform_constraints_to_local_cos(self):
        """reference motion_primitive_constraints.py:268-378 (host, once per step):
        targets x inv(aligning_transform); position / relative / CA constraints are
        rebuilt with weight 1.0 (quirk Q5), the others keep their weight."""
        if self.is_local or self.aligning_transform is None:
            return self
        inv = np.linalg.inv(np.asarray(self.aligning_transform, dtype=np.float64))
        out = MotionPrimitiveConstraints()
        out.start_pose = {"orientation": [0, 0, 0], "position": [0, 0, 0]}
        out.skeleton = self.skeleton
        out.is_local = True
        out.use_local_optimization = self.use_local_optimization

        def tp(p, w=1.0):
            return np.dot(inv, [p[0], p[1], p[2], w])[:3]

        def tp_masked(position):
            pos = [position[0], position[1], position[2], 1]
            idx = [i for i in range(3) if pos[i] is None]
            for i in idx:
                pos[i] = 0
            pos = np.dot(inv, pos)[:3].tolist()
            for i in idx:
                pos[i] = None
            return pos

        for c in self.constraints:
            ct = c.constraint_type
            base = {"canonical_keyframe": c.canonical_keyframe, "semanticAnnotation": c.semantic_annotation}
            if ct == SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSITION:
                if c.position is not None:
                    d = dict(base, joint=c.joint_name, position=tp_masked(c.position), n_canonical_frames=c.n_canonical_frames)
                    out.constraints.append(GlobalTransformConstraint(self.skeleton, d, 1.0))
            elif ct == SPATIAL_CONSTRAINT_TYPE_KEYFRAME_RELATIVE_POSITION:
                if c.position is not None:
                    d = dict(base, joint=c.joint_name, position=tp_masked(c.position), n_canonical_frames=c.n_canonical_frames,
                             offset=c.offset)
                    out.constraints.append(RelativeTransformConstraint(self.skeleton, d, 1.0))
            elif ct == SPATIAL_CONSTRAINT_TYPE_CA_CONSTRAINT:
                d = dict(base, joint=c.joint_name, position=tp(c.position).tolist(), n_canonical_frames=c.n_canonical_frames,
                         ca_constraint=True)
                out.constraints.append(GlobalTransformCAConstraint(self.skeleton, d, 1.0))
            elif ct == SPATIAL_CONSTRAINT_TYPE_TWO_HAND_POSITION:
                d = dict(base, joint=c.joint_names, positions=[tp(p) for p in c.positions], orientations=c.orientations,
                         n_canonical_frames=c.n_canonical_frames)
                out.constraints.append(TwoHandConstraintSet(self.skeleton, d, c.precision, c.weight_factor))
            elif ct == SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSE:
                d = dict(base, keyframeLabel="start", frame_constraint=[tp(p) for p in c.pose_constraint],
                         velocity_constraint=c.velocity_constraint, node_names=c.node_names, weights=c.weights)
                out.constraints.append(PoseConstraint(self.skeleton, d, c.precision, c.weight_factor))
            elif ct == SPATIAL_CONSTRAINT_TYPE_KEYFRAME_DIR_2D:
                dv = c.direction_constraint
                d = dict(base, dir_vector=tp(dv, 0.0))
                out.constraints.append(Direction2DConstraint(self.skeleton, d, c.precision, c.weight_factor))
            elif ct == SPATIAL_CONSTRAINT_TYPE_KEYFRAME_LOOK_AT:
                d = dict(base, position=tp(c.target_position))
                out.constraints.append(LookAtConstraint(self.skeleton, d, c.precision, c.weight_factor))
            elif ct == SPATIAL_CONSTRAINT_TYPE_KEYFRAME_FEET:
                d = dict(base, left=tp(c.left), right=tp(c.right))
                out.constraints.append(FeetConstraint(self.skeleton, d, c.precision, c.weight_factor))
        return out
