"""The batched slot the reference already has towards `mgrd`, served by
the engine: `motion_primitive_get_random_samples(mp, n)` and `MGRDSampleFilter` with the reference's
method names (motion_generator/mgrd_sample_filter.py:40-140), selected by
`constrained_sampling_mode == "random_spline"` (motion_primitive_generator.py:149-150,167-177).

`mgrd` itself is a third-party package that is not in the reference tree (`mgrd_sample_filter.py:28-37`
guards its import), so its scoring formula is not available to restate: the scores returned here are
the reference's OWN keyframe-position error (global_transform_constraint.py:102-116, weighted by
`weight_factor` like motion_primitive_constraints.py:120) of the constraints the reference hands to
mgrd - `extract_cartesian_constraints` (:52-60): position constraints only, unset axes replaced by 0.
Parity of this slot against mgrd is therefore unpinned; against the engine's own cost it is exact.
"""
import copy

import numpy as np

from .constraints import MotionPrimitiveConstraints, SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSITION
from .objective_functions import _device_constraints


def motion_primitive_get_random_samples(motion_primitive, n):
    """mgrd's sampler as the reference calls it (motion_primitive_generator.py:168): [n, D] latent vectors."""
    if hasattr(motion_primitive, "sample_low_dimensional_vectors"):
        return np.asarray(motion_primitive.sample_low_dimensional_vectors(n))
    return np.asarray(motion_primitive.sample_low_dimensional_vector(n))


class MGRDSampleFilter(object):
    def __init__(self, pose_constraint_weights=(1, 1)):
        self.pose_constraint_weights = pose_constraint_weights

    @staticmethod
    def extract_cartesian_constraints(mp_constraints):
        """:52-60 - position constraints with `None` axes set to 0, orientation dropped."""
        out = []
        for c in mp_constraints.constraints:
            if c.constraint_type == SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSITION and getattr(c, "position", None) is not None:
                cc = copy.copy(c)
                cc.position = [p if p is not None else 0 for p in c.position]
                cc.orientation = None
                out.append(cc)
        return out

    @staticmethod
    def score_samples(motion_primitive, samples, semantic_pose_constraints, cartesian_constraints, weights=(1, 1),
                      mp_constraints=None):
        """:62-75.  `cartesian_constraints`: list from extract_cartesian_constraints; semantic pose constraints
        are mgrd objects with no counterpart here and must be empty.  Returns float64 scores [N]."""
        if len(semantic_pose_constraints) > 0:
            from .engine import UnsupportedConstraintError
            raise UnsupportedConstraintError("mgrd semantic pose constraints have no device counterpart")
        samples = np.ascontiguousarray(samples, dtype=np.float64)
        if len(cartesian_constraints) == 0:
            return np.zeros(len(samples))
        mpc = MotionPrimitiveConstraints()
        mpc.skeleton = cartesian_constraints[0].skeleton if mp_constraints is None else mp_constraints.skeleton
        mpc.is_local = True
        mpc.constraints = list(cartesian_constraints)
        dc = _device_constraints(motion_primitive, mpc, None)
        return dc.score_host(samples, k=0, return_cost=True)["cost"].astype(np.float64)

    @staticmethod
    def score_samples_cartesian(motion_primitive, samples, mp_constraints):
        """:77-103 (local constraints; the global branch transforms spline coefficients through mgrd objects)."""
        cartesian = MGRDSampleFilter.extract_cartesian_constraints(mp_constraints)
        if len(cartesian) == 0:
            return np.zeros(len(samples))
        if not mp_constraints.is_local and mp_constraints.aligning_transform is not None:
            cartesian = MGRDSampleFilter.extract_cartesian_constraints(mp_constraints.transform_constraints_to_local_cos())
        return MGRDSampleFilter.score_samples(motion_primitive, samples, [], cartesian, mp_constraints=mp_constraints)
