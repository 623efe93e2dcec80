"""anim_utils.animation_data.utils stand-in (point clouds, small quaternion helpers).
Call sites: reference pose_constraint.py:59-64, global_transform_constraint.py:118-119,
look_at_constraint.py:66, motion_primitive_constraints.py:30."""
from math import sqrt

import numpy as np
from transformations import quaternion_multiply, quaternion_conjugate, quaternion_matrix  # noqa: F401 (re-exported)


def quaternion_rotate_vector(q, vector):
    """(q (x) v (x) q*)[1:], v given as a 4-vector quaternion."""
    tmp = quaternion_multiply(q, vector)
    return quaternion_multiply(tmp, quaternion_conjugate(q))[1:]


def quaternion_from_vector_to_vector(a, b):
    """In-tree copy: reference constraints/foot_step_constraints_builder.py:32-37."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    v = np.cross(a, b)
    w = np.sqrt((np.linalg.norm(a) ** 2) * (np.linalg.norm(b) ** 2)) + np.dot(a, b)
    q = np.array([w, v[0], v[1], v[2]])
    n = np.linalg.norm(q)
    if n > 0.0:
        return q / n
    return np.array([0.0, 0.0, 1.0, 0.0])


def transform_point(point, euler_angles, offset, origin=None, rotation_order=None):
    from .quaternion_frame import euler_to_quaternion
    q = euler_to_quaternion(euler_angles)
    m = quaternion_matrix(q)
    p = np.dot(m[:3, :3], np.asarray(point, dtype=np.float64))
    return p + np.asarray(offset, dtype=np.float64)


def convert_quaternion_frame_to_cartesian_frame(skeleton, quat_frame, node_names=None):
    return skeleton.convert_quaternion_frame_to_cartesian_frame(quat_frame, node_names)


def get_cartesian_coordinates_from_quaternion(skeleton, node_name, quaternion_frame, return_global_matrix=False):
    return skeleton.nodes[node_name].get_global_position(quaternion_frame)


def align_point_clouds_2D(a, b, weights):
    """Weighted 2-D rigid alignment of b onto a (Kovar et al., Motion Graphs)."""
    if len(a) != len(b):
        raise ValueError("two point cloud should have the same number points: " + str(len(a)) + "," + str(len(b)))
    n_points = len(a)
    numerator_left = 0.0
    denominator_left = 0.0
    weighted_sum_a_x = 0.0
    weighted_sum_b_x = 0.0
    weighted_sum_a_z = 0.0
    weighted_sum_b_z = 0.0
    sum_of_weights = 0.0
    for i in range(n_points):
        numerator_left += weights[i] * (a[i][0] * b[i][2] - b[i][0] * a[i][2])
        denominator_left += weights[i] * (a[i][0] * b[i][0] + a[i][2] * b[i][2])
        sum_of_weights += weights[i]
        weighted_sum_a_x += weights[i] * a[i][0]
        weighted_sum_b_x += weights[i] * b[i][0]
        weighted_sum_a_z += weights[i] * a[i][2]
        weighted_sum_b_z += weights[i] * b[i][2]
    numerator_right = 1.0 / sum_of_weights * (weighted_sum_a_x * weighted_sum_b_z - weighted_sum_b_x * weighted_sum_a_z)
    numerator = numerator_left - numerator_right
    denominator_right = 1.0 / sum_of_weights * (weighted_sum_a_x * weighted_sum_b_x + weighted_sum_a_z * weighted_sum_b_z)
    denominator = denominator_left - denominator_right
    theta = np.arctan2(numerator, denominator)
    offset_x = (weighted_sum_a_x - weighted_sum_b_x * np.cos(theta) - weighted_sum_b_z * np.sin(theta)) / sum_of_weights
    offset_z = (weighted_sum_a_z + weighted_sum_b_x * np.sin(theta) - weighted_sum_b_z * np.cos(theta)) / sum_of_weights
    return theta, offset_x, offset_z


def transform_point_cloud(point_cloud, theta, offset_x, offset_z):
    """Rotate about +y by theta, then translate in x and z."""
    c, s = np.cos(theta), np.sin(theta)
    out = []
    for p in point_cloud:
        if p is not None:
            out.append(np.array([c * p[0] + s * p[2] + offset_x, p[1], -s * p[0] + c * p[2] + offset_z]))
    return out


def calculate_point_cloud_distance(a, b):
    assert len(a) == len(b)
    distance = 0.0
    n_points = len(a)
    for i in range(n_points):
        d = [a[i][0] - b[i][0], a[i][1] - b[i][1], a[i][2] - b[i][2]]
        distance += sqrt(d[0] ** 2 + d[1] ** 2 + d[2] ** 2)
    return distance / n_points


def quat_distance(quat_a, quat_b):
    """PARITY UNPINNED (anim_utils is absent).  Restated after the reference's own in-tree twin of this
    comparison, JointRotationConstraint.evaluate_frame (joint_rotation_constraint.py:62-75): normalise
    both quaternions, Frobenius norm of the difference of their rotation matrices."""
    from transformations import quaternion_matrix
    quat_a = np.asarray(quat_a, dtype=np.float64)
    quat_b = np.asarray(quat_b, dtype=np.float64)
    quat_a = quat_a / np.linalg.norm(quat_a)
    quat_b = quat_b / np.linalg.norm(quat_b)
    return float(np.linalg.norm(np.ravel(quaternion_matrix(quat_a) - quaternion_matrix(quat_b))))


def calculate_weighted_frame_distance_quat(quat_frame_a, quat_frame_b, weights):
    """PARITY UNPINNED.  Joint i + 1 weighted by weights[i], root excluded - the loop the reference's own
    residual form of the same comparison runs (pose_constraint_quat_frame.py:60-70)."""
    diff = 0.0
    for i in range(len(weights) - 1):
        q1 = quat_frame_a[(i + 1) * 4 + 3:(i + 2) * 4 + 3]
        q2 = quat_frame_b[(i + 1) * 4 + 3:(i + 2) * 4 + 3]
        diff += quat_distance(q1, q2) * weights[i]
    return diff


def get_arc_length_from_points(points):
    points = np.asarray(points, dtype=np.float64)
    return float(np.sum(np.linalg.norm(points[1:] - points[:-1], axis=1)))
