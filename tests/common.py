"""Shared helpers of the test-suite: rebuild a golden case (model, skeleton,
constraints) from seeds + the fixture's stored specs, using the PRODUCT's
host-side mirror classes (no reference needed -> works on the GPU box)."""
import json
import os

import numpy as np

from morphablegraphs_b200 import synthetic as syn
from morphablegraphs_b200 import (Skeleton, MotionStateGraphNode, MotionPrimitiveConstraints, CONSTRAINT_FACTORY)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["s_config1", "s_mixed", "s_mixed_tm", "s_nonlocal", "xs_mixed", "l_mixed"]
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
