"""Objective functions with the reference's signature `obj(s, data) -> float`,
`data = (mp_node, mp_constraints, prev_frames)` (reference
motion_generator/optimization/objective_functions.py:142-160,
mg_state_planner.py:110-129), plus a `.batch(S, data) -> float[N]` fast path
that scores a whole candidate batch in one launch.

Each call compiles the constraint list once per (constraints object, prev_frames)
and caches the device program on the constraints object.
"""
import numpy as np

from . import engine


def _device_constraints(mp_node, mp_constraints, prev_frames):
    mp = getattr(mp_node, "motion_primitive", mp_node)
    skeleton = mp_constraints.skeleton
    dm = mp.device_model(skeleton) if hasattr(mp, "device_model") else _foreign_model(mp_node, skeleton)
    cache = getattr(mp_constraints, "_mgb_cache", None)
    key = (id(dm), len(mp_constraints.constraints), None if prev_frames is None else id(prev_frames))
    if cache is None or cache[0] != key:
        dc = engine.DeviceConstraints(dm, mp_constraints, prev_frames)
        mp_constraints._mgb_cache = (key, dc)
        cache = mp_constraints._mgb_cache
    return cache[1]


_foreign = {}


def _foreign_model(mp_node, skeleton):
    """DeviceModel for a reference-side (non-mirror) model object, cached by identity."""
    key = (id(mp_node), id(skeleton))
    if key not in _foreign:
        _foreign[key] = engine.DeviceModel(mp_node, skeleton)
    return _foreign[key]


class _Objective(object):
    error_scale = 1.0
    quality_scale = 0.0
    magnitude_scale = 0.0

    def batch(self, S, data, k=0):
        mp_node, mp_constraints, prev_frames = data[:3]
        dc = _device_constraints(mp_node, mp_constraints, prev_frames)
        S = np.ascontiguousarray(np.atleast_2d(S))
        res = dc.score_host(S, k=k, error_scale=self.error_scale, quality_scale=self.quality_scale,
                            magnitude_scale=self.magnitude_scale, return_cost=True)
        mp_constraints.evaluations += len(S)
        return res["cost"].astype(np.float64)

    def batch_device(self, S_dev, data):
        """Costs of a device-resident table [M, D] (float32/float64 CUDA tensor) -> float32 CUDA
        tensor [M]: K8, the node-cost table of a cluster-tree search (space_partitioning.py)."""
        mp_node, mp_constraints, prev_frames = data[:3]
        dc = _device_constraints(mp_node, mp_constraints, prev_frames)
        mp_constraints.evaluations += int(S_dev.shape[0])
        return dc.node_costs(S_dev, error_scale=self.error_scale, quality_scale=self.quality_scale,
                             magnitude_scale=self.magnitude_scale)

    def __call__(self, s, data):
        mp_constraints = data[1]
        v = float(self.batch(np.asarray(s, dtype=np.float64)[None, :], data)[0])
        mp_constraints.min_error = v if self.magnitude_scale == 0.0 else mp_constraints.min_error
        return v


class _ObjSpatialErrorSum(_Objective):
    """reference objective_functions.py:142-160"""
    __name__ = "obj_spatial_error_sum"


class _ObjSpatialErrorSumAndMagnitude(_Objective):
    """reference mg_state_planner.py:110-129: spatial error + 2 |s|"""
    __name__ = "obj_spatial_error_sum_and_magnitude"
    magnitude_scale = 2.0


class ObjSpatialErrorAndNaturalness(_Objective):
    """error_scale * spatial error + quality_scale * (-log p(s)): the scalar the reference's
    `obj_spatial_error_sum_and_naturalness` computes but never returns (objective_functions.py:163-186, Q7)."""
    __name__ = "obj_spatial_error_sum_and_naturalness"

    def __init__(self, error_scale=1.0, quality_scale=1.0):
        self.error_scale = error_scale
        self.quality_scale = quality_scale


obj_spatial_error_sum = _ObjSpatialErrorSum()
obj_spatial_error_sum_and_magnitude = _ObjSpatialErrorSumAndMagnitude()
