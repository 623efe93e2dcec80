#!/usr/bin/env python
"""Under torchrun (one rank per GPU): the sharded generator step returns, on every rank, exactly the
sample a single-GPU run of the same job returns (Philox is keyed by the global index, the per-rank
bests are exchanged with the first-minimum rule)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import MotionPrimitiveGenerator, engine, sharding  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    ok, e_dist, idx = True, 0.0, -1
    for wl, n, qs in (("S1M", 1_000_003, 0.0), ("L4M", 200_001, 0.1)):
        from morphablegraphs_b200.objective_functions import ObjSpatialErrorAndNaturalness
        node, skeleton, mpc, prev = bench.build_case(wl)
        gen = MotionPrimitiveGenerator(None, {"n_random_samples": n, "use_local_coordinates": prev is None}, skeleton=skeleton, seed=11)
        if qs:
            gen.objective = ObjSpatialErrorAndNaturalness(1.0, qs)
        # sharded: one launch per rank, the exchange of the per-rank best inside the kernel's epilogue
        s_dist = gen._get_best_fit_sample_using_gmm(node, mpc, prev_frames=prev)
        e_dist = mpc.min_error
        # the same job on this rank alone
        dm = engine.DeviceModel(node, skeleton, local)
        dc = engine.DeviceConstraints(dm, mpc, prev)
        seed = int(np.random.RandomState(11).randint(0, 2 ** 31 - 1))
        counts = dm.component_counts(n, seed)
        e_one, idx, s_one = dc.sample_score_best(n, seed=seed, counts=counts, quality_scale=qs)
        ok = ok and np.array_equal(s_dist, s_one) and e_dist == e_one
        # a generator built WITHOUT a seed: rank 0's stream is broadcast once, so every rank still returns the same sample
        gen_u = MotionPrimitiveGenerator(None, {"n_random_samples": n, "use_local_coordinates": prev is None}, skeleton=skeleton, seed=None)
        s_u = gen_u._get_best_fit_sample_using_gmm(node, mpc, prev_frames=prev)
        got = [None] * world
        dist.all_gather_object(got, (s_u.tolist(), float(mpc.min_error)))
        ok = ok and all(g == got[0] for g in got)
    flags = [None] * world
    dist.all_gather_object(flags, (bool(ok), e_dist, idx))
    if rank == 0:
        print("world", world, "per-rank (ok, min_error, single-GPU idx):", flags)
    dist.destroy_process_group()
    sys.exit(0 if all(f[0] for f in flags) else 1)


if __name__ == "__main__":
    main()
