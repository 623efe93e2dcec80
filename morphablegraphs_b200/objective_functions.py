"""Objective functions with the reference's signature `obj(s, data) -> float`,
`data = (mp_node, mp_constraints, prev_frames)` (reference
motion_generator/optimization/objective_functions.py:142-160,
mg_state_planner.py:110-129), plus a `.batch(S, data) -> float[N]` fast path
that scores a whole candidate batch in one launch.

Each call compiles the constraint list once per (constraints object, prev_frames)
and caches the device program on the constraints object.
"""
import ctypes as C

import numpy as np

from . import engine, _lib
from ._lib import MgbError


def constraints_key(dm, flat, mp_constraints, prev_frames):
    """Content key of a compiled constraint program: the flattened descriptors (type, joints, time
    slots, targets, weights, masks), the pose points, the evaluation times, `is_local` and the last
    previous frame (what the alignment reads).  In-place mutation of a constraint - the reference's
    planner rewrites `canonical_keyframe` before scoring, graph_walk_planner.py:199-203 - therefore
    yields a new key and a recompile; object identity is never trusted."""
    descs, n, parr, n_points, times = flat
    is_local = bool(getattr(mp_constraints, "is_local", False))
    prev = b""
    if not is_local and prev_frames is not None:
        prev = np.ascontiguousarray(prev_frames[-1], dtype=np.float64).tobytes()
    start_pose = getattr(mp_constraints, "start_pose", None) is not None
    return (id(dm), n, n_points, bytes(descs)[:n * C.sizeof(_lib.ConstraintDesc)],
            bytes(parr)[:n_points * C.sizeof(_lib.PosePoint)], times.tobytes(), is_local, prev, start_pose)


def _device_constraints(mp_node, mp_constraints, prev_frames):
    """DeviceConstraints of (model, constraint list, prev_frames), cached on the constraints object under a
    CONTENT key (see constraints_key): the constraint list is flattened on every call - that is the cheap
    part - and the program is recompiled and re-uploaded only when a byte of it changed."""
    mp = getattr(mp_node, "motion_primitive", mp_node)
    skeleton = mp_constraints.skeleton
    dm = mp.device_model(skeleton) if hasattr(mp, "device_model") else _foreign_model(mp_node, skeleton)
    flat = engine.flatten_constraints(dm, mp_constraints)
    key = constraints_key(dm, flat, mp_constraints, prev_frames)
    cache = getattr(mp_constraints, "_mgb_cache", None)
    if cache is None or cache[0] != key:
        dc = engine.DeviceConstraints(dm, mp_constraints, prev_frames, flat=flat)
        cache = (key, dc)
        try:
            mp_constraints._mgb_cache = cache
        except AttributeError:      # objects with __slots__: no cache, still correct
            pass
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

# the reference's own scalar objectives, recognised by function name when a caller assigns the plain
# function (mg_state_planner.py:250 does `generator.objective = obj_spatial_error_sum_and_magnitude`)
_KNOWN_OBJECTIVES = {
    "obj_spatial_error_sum": (1.0, 0.0, 0.0),                   # objective_functions.py:142-160
    "obj_spatial_error_sum_and_magnitude": (1.0, 0.0, 2.0),     # mg_state_planner.py:110-129
}


def objective_scales(obj):
    """(error_scale, quality_scale, magnitude_scale) of an objective the device path can evaluate:
    one of this module's objective objects, or one of the reference's scalar objective FUNCTIONS
    recognised by `__name__`.  Anything else raises - a term is never dropped silently
    (SURVEY.md 8(b))."""
    if isinstance(obj, _Objective):
        return float(obj.error_scale), float(obj.quality_scale), float(obj.magnitude_scale)
    name = getattr(obj, "__name__", None)
    if callable(obj) and name in _KNOWN_OBJECTIVES:
        return _KNOWN_OBJECTIVES[name]
    raise MgbError("objective %r is not one of the objectives the device path evaluates (%s, or an object of "
                   "morphablegraphs_b200.objective_functions); refusing to score with a different formula"
                   % (obj, ", ".join(sorted(_KNOWN_OBJECTIVES))))
