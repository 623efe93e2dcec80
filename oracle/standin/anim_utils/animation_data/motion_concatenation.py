"""anim_utils.animation_data.motion_concatenation stand-in: 2-D heading alignment.

Forced by in-tree copies: reference mg_state_planner.py:92-107 (aligning
transform), objective_functions.py:35-49 (start-pose variant),
mg_state_queue.py:112-115 (transform of frames), motion_generator/utils.py:69-84
(heading helpers).  The sign convention of get_rotation_angle is chosen so that
the alignment is geometrically correct (after alignment the heading of the new
first frame equals the heading of the previous last frame); tests/ freeze this.
"""
import numpy as np
from transformations import (quaternion_matrix, quaternion_from_matrix, quaternion_multiply,
                             quaternion_about_axis, quaternion_from_euler)


def get_orientation_vector_from_matrix(m, v=(0, 0, 1)):
    p = np.dot(m, np.asarray(v, dtype=np.float64))
    dir_vec = np.array([p[0], p[2]])
    dir_vec /= np.linalg.norm(dir_vec)
    return dir_vec


def get_global_node_orientation_vector(skeleton, node_name, frame, v=(0, 0, 1)):
    m = skeleton.nodes[node_name].get_global_matrix(frame)[:3, :3]
    return get_orientation_vector_from_matrix(m, v)


def point_to_euler_angle(vec):
    """Heading of an (x, z) direction in degrees, measured from +z towards +x
    (the sense in which a positive rotation about +y turns a vector)."""
    return np.rad2deg(np.arctan2(vec[0], vec[1]))


def get_rotation_angle(point1, point2):
    """Degrees to rotate (about +y) heading `point2` onto heading `point1`, in (-180, 180]."""
    a = point_to_euler_angle(point1) - point_to_euler_angle(point2)
    while a > 180.0:
        a -= 360.0
    while a <= -180.0:
        a += 360.0
    return a


def get_node_aligning_2d_transform(skeleton, node_name, prev_frames, new_frames, ref_vector=(0, 0, 1)):
    """From the last of prev_frames to the first of new_frames: rotation about +y
    and x/z translation (y untouched)."""
    m_a = skeleton.nodes[node_name].get_global_matrix(prev_frames[-1])
    m_b = skeleton.nodes[node_name].get_global_matrix(new_frames[0])
    dir_vec_a = get_orientation_vector_from_matrix(m_a[:3, :3], ref_vector)
    dir_vec_b = get_orientation_vector_from_matrix(m_b[:3, :3], ref_vector)
    angle = get_rotation_angle(dir_vec_a, dir_vec_b)
    q = quaternion_about_axis(np.deg2rad(angle), [0, 1, 0])
    m = quaternion_matrix(q)
    first_frame_pos = [new_frames[0][0], new_frames[0][1], new_frames[0][2], 1.0]
    rotated_first_frame_pos = np.dot(m, first_frame_pos)[:3]
    delta = np.asarray(prev_frames[-1][:3]) - rotated_first_frame_pos[:3]
    m[0, 3] = delta[0]
    m[2, 3] = delta[2]
    return m


def transform_quaternion_frames(frames, m, translation_param_range=(0, 3), orientation_param_range=(3, 7)):
    q = quaternion_from_matrix(m)
    for frame in frames:
        ot = list(frame[:3]) + [1.0]
        frame[:3] = np.dot(m, ot)[:3]
        frame[3:7] = quaternion_multiply(q, frame[3:7])
    return frames


def get_transform_from_start_pose(start_pose):
    e = start_pose["orientation"]
    p = start_pose["position"]
    if None not in e:
        rad = np.deg2rad(e)
        m = quaternion_matrix(quaternion_from_euler(*rad))
    else:
        m = np.eye(4)
    if None not in p:
        m[:3, 3] = p
    return m


def align_quaternion_frames_automatically(skeleton, node_name, new_frames, prev_frames):
    m = get_node_aligning_2d_transform(skeleton, node_name, prev_frames, new_frames)
    return transform_quaternion_frames(new_frames, m)


def align_quaternion_frames(skeleton, node_name, new_frames, prev_frames=None, start_pose=None):
    if prev_frames is not None:
        return align_quaternion_frames_automatically(skeleton, node_name, new_frames, prev_frames)
    elif start_pose is not None:
        m = get_transform_from_start_pose(start_pose)
        first_frame_pos = list(new_frames[0][:3]) + [1]
        t_pos = np.dot(m, first_frame_pos)[:3]
        delta = start_pose["position"]
        delta[0] -= t_pos[0]
        delta[2] -= t_pos[2]
        m[:3, 3] = delta
        return transform_quaternion_frames(new_frames, m)
    return new_frames
