#!/usr/bin/env python
"""Per-step host cost of the planner-facing call at L: constraint program compile + table upload (DeviceConstraints),
first launch, the rest of generate_constrained_sample.  usage: python tools/setup_cost.py [S1M|L4M]"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import engine  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "L4M"
node, skeleton, mpc, prev = bench.build_case(wl)
dm = engine.DeviceModel(node, skeleton, 0)
n = 100000
qs = bench.WORKLOADS[wl]["quality_scale"]


def wall(fn, reps=5):
    out = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        out.append(1e3 * (time.perf_counter() - t0))
    return r, sorted(out)[len(out) // 2]


dc, t_dc = wall(lambda: engine.DeviceConstraints(dm, mpc, prev))
counts, t_counts = wall(lambda: dm.component_counts(4000000, seed=5))
_, t_first = wall(lambda: engine.DeviceConstraints(dm, mpc, prev).sample_score(n, seed=7, counts=dm.component_counts(n, seed=7), k=1, quality_scale=qs))
_, t_warm = wall(lambda: dc.sample_score(n, seed=7, counts=dm.component_counts(n, seed=7), k=1, quality_scale=qs))
print("%s: DeviceConstraints() %.2f ms, component_counts(4M) %.2f ms, build + first launch (n=100k) %.2f ms, warm launch %.2f ms"
      % (wl, t_dc, t_counts, t_first, t_warm))
nt = bench.WORKLOADS[wl]["n_per_gpu"]
counts = dm.component_counts(nt, seed=7)
_, t_one = wall(lambda: dm.sample(1, seed=7, index_base=nt // 2, counts=counts, n_total=nt))
_, t_full = wall(lambda: dc.sample_score(nt, seed=7, counts=counts, k=1, quality_scale=qs), reps=3)
from morphablegraphs_b200 import MotionPrimitiveGenerator  # noqa: E402
from morphablegraphs_b200.objective_functions import ObjSpatialErrorAndNaturalness  # noqa: E402
gen = MotionPrimitiveGenerator(None, {"n_random_samples": nt, "use_local_coordinates": prev is None}, skeleton=skeleton, seed=11)
if qs:
    gen.objective = ObjSpatialErrorAndNaturalness(1.0, qs)
_, t_gen_cached = wall(lambda: gen.generate_constrained_sample(node, mpc, prev_frames=prev), reps=3)


def fresh():
    if hasattr(mpc, "_mgb_cache"):
        del mpc._mgb_cache
    return gen.generate_constrained_sample(node, mpc, prev_frames=prev)


_, t_gen_fresh = wall(fresh, reps=3)
print("%s: sample(1 of %d) %.2f ms, sample_score(%d) %.2f ms, generator call cached constraints %.2f ms, fresh constraints %.2f ms"
      % (wl, nt, t_one, nt, t_full, t_gen_cached, t_gen_fresh))
