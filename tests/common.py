"""Shared helpers of the test-suite: rebuild a golden case (model, skeleton,
constraints) from seeds + the fixture's stored specs, using the PRODUCT's
host-side mirror classes (no reference needed -> works on the GPU box)."""
import json
import os

import numpy as np

from morphablegraphs_b200 import synthetic as syn
from morphablegraphs_b200 import (Skeleton, MotionStateGraphNode, MotionPrimitiveConstraints, CONSTRAINT_FACTORY)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["s_config1", "s_mixed", "s_mixed_tm", "s_nonlocal", "xs_mixed", "l_mixed", "s_rot", "s_rot_nonlocal", "s_traj", "s_traj_nonlocal"]
RTOL = 1e-4   # BASELINE.json north_star: 1e-4 relative in fp32


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


class Case(object):
    def __init__(self, name):
        g = load_golden(name)
        self.g = g
        self.name = name
        self.shape = str(g["shape"])
        cfg = syn.SHAPES[self.shape]
        self.F = cfg["n_frames"]
        self.model_json = syn.make_model_json(self.shape, translation_maxima=tuple(g["translation_maxima"]))
        self.skeleton_desc = syn.make_skeleton_desc(cfg["n_joints"])
        self.skeleton = Skeleton.from_desc(self.skeleton_desc)
        self.node = MotionStateGraphNode()
        self.node._initialize_from_json(None, self.model_json)
        self.node.set_skeleton(self.skeleton)
        self.specs = json.loads(str(g["specs_json"]))
        self.constraints = MotionPrimitiveConstraints()
        self.constraints.skeleton = self.skeleton
        self.constraints.is_local = bool(g["is_local"])
        self.constraints.constraints = syn.build_constraints(CONSTRAINT_FACTORY, self.skeleton, self.specs)
        self.prev_frames = g["prev_frames"] if "prev_frames" in g.files else None
        self.samples = g["samples"]


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), 1.0)


def describe_constraint(c):
    """Every field of a constraint object the hot path reads, as plain lists / floats: works on the reference's own
    classes and on the product's mirror classes alike (used to compare `transform_constraints_to_local_cos`)."""
    def lst(v):
        if v is None:
            return None
        return [None if x is None else float(x) for x in np.asarray(v, dtype=object).ravel()]

    d = {"type": c.constraint_type, "weight_factor": float(c.weight_factor),
         "canonical_keyframe": None if c.canonical_keyframe is None else float(c.canonical_keyframe)}
    for name in ("joint_name", "joint_names", "node_names"):
        if hasattr(c, name):
            v = getattr(c, name)
            d[name] = v if isinstance(v, str) else list(v)
    for name in ("position", "orientation", "offset", "target_dir", "direction_constraint", "pose_constraint", "weights",
                 "velocity_constraint", "positions", "target_center", "left", "right", "target_position"):
        if hasattr(c, name):
            d[name] = lst(getattr(c, name))
    if hasattr(c, "n_canonical_frames"):
        d["n_canonical_frames"] = int(c.n_canonical_frames)
    return d


def assert_same_description(a, b, tol=1e-12):
    assert a.keys() == b.keys(), (sorted(a.keys()), sorted(b.keys()))
    for k in a:
        va, vb = a[k], b[k]
        if isinstance(va, list) and va and not isinstance(va[0], str):
            assert len(va) == len(vb), k
            for x, y in zip(va, vb):
                assert (x is None) == (y is None), k
                if x is not None:
                    assert abs(x - y) <= tol * max(1.0, abs(y)), (k, x, y)
        elif isinstance(va, float):
            assert abs(va - vb) <= tol * max(1.0, abs(vb)), (k, va, vb)
        else:
            assert va == vb, (k, va, vb)
