#!/usr/bin/env python
"""Under torchrun (one rank per GPU): config 5 sharded - (1) the sharded cluster-tree build with injected seeds gives,
on every rank, the tree the reference's own builder produced (tests/golden/s_tree.npz); (2) a 1M-sample sharded build:
all ranks agree, every sample is one kd node, timing; (3) sharded euclidean PCA = single-GPU PCA; (4) sharded K8 search
returns the single-GPU result."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from morphablegraphs_b200 import features as feat, sharding, space_partitioning as sp  # noqa: E402
from morphablegraphs_b200.objective_functions import obj_spatial_error_sum, _device_constraints  # noqa: E402
from tests.test_tree_cpu import TreeCase, _flat_to_dict, _same_cluster_tree  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    tc = TreeCase()
    ok = True
    fails = []
    seeds = lambda depth, gids, gcounts: torch.stack([(gcounts * j) // 4 for j in range(4)], dim=1)
    for max_level in (6, 2):
        tree = sp.ClusterTree(4, max_level, tc.K).construct(tc.data, seeds=seeds)
        ref = json.loads(str(tc.g["cluster_tree_seeded%d_json" % max_level]))
        try:
            _same_cluster_tree(_flat_to_dict(tree.flat), ref, tol=5e-5)
        except AssertionError:
            ok = False
            fails.append("seeded build max_level=%d" % max_level)
    # sharded search on the reference-built tree
    t = sp.ClusterTree.from_dict(json.loads(str(tc.g["cluster_tree_json"])), 4, 6, tc.K)
    t.shard_search = True
    for nc, ref in zip(tc.g["n_candidates"], tc.g["cluster_tree_results"]):
        v, s = t.find_best_example_excluding_search_candidates(obj_spatial_error_sum, (tc.node, tc.mpc, None), int(nc))
        good = abs(v - ref[0]) <= 1e-4 * max(1.0, abs(ref[0])) and np.array_equal(np.asarray(s), ref[1:])
        if not good:
            fails.append("sharded search nc=%d: %r vs %r" % (nc, v, ref[0]))
        ok = ok and good
    # 1M samples
    dm = _device_constraints(tc.node, tc.mpc, None).dmodel
    n = 1_000_000
    S = dm.sample(n, seed=77)
    data = S.cpu().numpy().astype(np.float64)
    sp.ClusterTree(4, 8, tc.K).construct(data[:20000], seed=5)        # warm-up
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    big = sp.ClusterTree(4, 8, tc.K).construct(data, seed=5)
    torch.cuda.synchronize(); dist.barrier()
    dt = time.perf_counter() - t0
    rows = big.flat.kd_row - big.flat.n_nodes
    good = len(rows) == n and np.array_equal(np.sort(rows), np.arange(n))
    if not good:
        fails.append("1M build: kd rows do not cover the samples (%d)" % len(rows))
    ok = ok and good
    sig = torch.tensor([float(big.flat.n_nodes), float(big.flat.points[:big.flat.n_nodes].sum()), float(big.flat.kd_left.sum())],
                       dtype=torch.float64, device="cuda")
    sigs = [torch.zeros_like(sig) for _ in range(world)]
    dist.all_gather(sigs, sig)
    good = all(torch.equal(s_, sigs[0]) for s_ in sigs)
    if not good:
        fails.append("1M build: ranks disagree %r" % [s_.tolist() for s_ in sigs])
    ok = ok and good
    # sharded PCA of the euclidean features of 40k samples vs the same on one GPU
    m = 40000
    clouds = feat.map_samples_to_euclidean_space(dm, S[:m])
    f_all, r_all = feat.pca_by_variance(clouds)
    base, cnt = sharding.shard_range(m, rank, world)
    f_sh, r_sh = feat.pca_by_variance(clouds[base:base + cnt].contiguous(), sharded=True)
    good = f_sh.shape[1] == f_all.shape[1] and np.abs(r_sh - r_all).max() < 1e-5
    if good:
        err = float((f_sh.abs() - f_all[base:base + cnt].abs()).abs().max()) / float(f_all.abs().max())
        good = err < 2e-3
        if not good:
            fails.append("sharded PCA features differ: %g" % err)
    else:
        fails.append("sharded PCA: components %d vs %d, ratio err %g" % (f_sh.shape[1], f_all.shape[1], np.abs(r_sh[:min(len(r_sh), len(r_all))] - r_all[:min(len(r_sh), len(r_all))]).max()))
    ok = ok and good
    flags = [None] * world
    dist.all_gather_object(flags, bool(ok))
    allf = [None] * world
    dist.all_gather_object(allf, fails)
    if rank == 0:
        print(json.dumps({"world": world, "ok": flags, "fails": allf[0], "sharded_build_1M_seconds": dt, "n_nodes": int(big.flat.n_nodes),
                          "levels": big.build_info["levels"]}))
    dist.destroy_process_group()
    sys.exit(0 if all(flags) else 1)


if __name__ == "__main__":
    main()
