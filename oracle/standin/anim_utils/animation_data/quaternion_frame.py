import numpy as np
from transformations import euler_matrix, quaternion_from_matrix, euler_from_matrix, quaternion_matrix

DEFAULT_ROTATION_ORDER = ['Xrotation', 'Yrotation', 'Zrotation']


def _axes(rotation_order):
    return 'r' + ''.join(o[0].lower() for o in rotation_order)


def euler_to_quaternion(euler_angles, rotation_order=DEFAULT_ROTATION_ORDER, filter_values=True):
    """Euler angles in degrees -> (w, x, y, z)."""
    rad = np.deg2rad(euler_angles)
    m = euler_matrix(rad[0], rad[1], rad[2], axes=_axes(rotation_order))
    q = quaternion_from_matrix(m)
    return q[0], q[1], q[2], q[3]


def quaternion_to_euler(q, rotation_order=DEFAULT_ROTATION_ORDER):
    q = np.asarray(q, dtype=np.float64)
    q = q / np.linalg.norm(q)
    return np.rad2deg(euler_from_matrix(quaternion_matrix(q), _axes(rotation_order))).tolist()
