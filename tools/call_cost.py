#!/usr/bin/env python
"""Where the planner-facing call spends its host time at S1M (median of many calls, microseconds).
usage: python tools/call_cost.py"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import engine, MotionPrimitiveGenerator  # noqa: E402
from morphablegraphs_b200 import objective_functions as OF  # noqa: E402

node, skeleton, mpc, prev = bench.build_case("S1M")
dm = node.motion_primitive.device_model(skeleton)
n = 1000000


def med(fn, reps=200):
    out = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        out.append(1e6 * (time.perf_counter() - t0))
    return sorted(out)[len(out) // 2]


flat = engine.flatten_constraints(dm, mpc)
dc = engine.DeviceConstraints(dm, mpc, prev, flat=flat)
counts = dm.component_counts(n, 7)
gen = MotionPrimitiveGenerator(None, {"n_random_samples": n, "use_local_coordinates": prev is None}, skeleton=skeleton, seed=11)
c0 = mpc.constraints[0]
state = [0]


def fresh_call():
    state[0] += 1
    c0.weight_factor = 1.0 + 1e-6 * (state[0] % 2)
    return gen.generate_constrained_sample(node, mpc, prev_frames=prev)


rows = [("flatten_constraints", med(lambda: engine.flatten_constraints(dm, mpc))),
        ("constraints_key", med(lambda: OF.constraints_key(dm, flat, mpc, prev))),
        ("DeviceConstraints() [basis + mgb_constraints_pack + info]", med(lambda: engine.DeviceConstraints(dm, mpc, prev, flat=flat))),
        ("component_counts", med(lambda: dm.component_counts(n, 7))),
        ("sample_score_best (launch + kernel + D2H + sync)", med(lambda: dc.sample_score_best(n, seed=7, counts=counts))),
        ("sample_score_best, n = 100", med(lambda: dc.sample_score_best(100, seed=7, counts=dm.component_counts(100, 7)))),
        ("generate_constrained_sample, cached program", med(lambda: gen.generate_constrained_sample(node, mpc, prev_frames=prev))),
        ("generate_constrained_sample, new program every call", med(fresh_call))]
for name, us in rows:
    print("%-62s %8.1f us" % (name, us))
