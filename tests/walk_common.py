"""BASELINE config 3 in miniature (tests/golden/s_walk.npz, oracle/make_golden.py: run_walk_case): a walk
over four S-shaped primitives whose every step was decided by the reference's own generator loop."""
import json

import numpy as np

from morphablegraphs_b200 import synthetic as syn
from morphablegraphs_b200 import Skeleton, MotionStateGraphNode, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
from tests.common import load_golden


class WalkCase(object):
    def __init__(self):
        g = load_golden("s_walk")
        self.g = g
        cfg = syn.SHAPES["S"]
        self.skeleton = Skeleton.from_desc(syn.make_skeleton_desc(cfg["n_joints"]))
        names = json.loads(str(g["node_names_json"]))
        self.nodes = []
        for nm, sd in zip(names, g["model_seeds"]):
            node = MotionStateGraphNode()
            node._initialize_from_json(None, syn.make_model_json("S", seed=int(sd), name=nm))
            node.set_skeleton(self.skeleton)
            self.nodes.append(node)
        self.order = [int(k) for k in g["order"]]

    def step(self, i):
        g = self.g
        prev = g["step%d_prev_frames" % i] if ("step%d_prev_frames" % i) in g.files else None
        mpc = MotionPrimitiveConstraints()
        mpc.skeleton = self.skeleton
        mpc.is_local = prev is None
        mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, self.skeleton, json.loads(str(g["step%d_specs_json" % i])))
        return dict(node=self.nodes[self.order[i]], constraints=mpc, prev_frames=prev, samples=g["step%d_samples" % i],
                    costs=g["step%d_costs" % i], best_idx=int(g["step%d_best_idx" % i]), min_error=float(g["step%d_min_error" % i]),
                    last_frames=g["step%d_last_frames" % i])
