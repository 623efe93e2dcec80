"""Host-side logic of the multi-GPU path on CPU: world_size-2 gloo processes exchange their
per-rank best (cost, global index) and apply the reference's first-minimum rule
(reference motion_primitive_generator.py:251-257); shard ranges tile the job."""
import os
import socket

import numpy as np
import pytest

from morphablegraphs_b200 import sharding

torch = pytest.importorskip("torch")


def test_shard_ranges_tile_the_job():
    for n in (0, 1, 7, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            nxt = 0
            for r in range(world):
                base, cnt = sharding.shard_range(n, r, world)
                assert base == nxt and cnt >= 0
                nxt = base + cnt
            assert nxt == n
            sizes = [sharding.shard_range(n, r, world)[1] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_select_best_is_the_reference_rule():
    # serial loop of the reference: strict '<' in index order
    rng = np.random.RandomState(0)
    for _ in range(200):
        n = rng.randint(1, 12)
        costs = rng.randint(0, 4, size=n).astype(float)
        costs[rng.rand(n) < 0.2] = np.nan
        idx = rng.permutation(100)[:n]
        order = np.argsort(idx)
        best_c, best_i = np.inf, None
        for j in order:
            if best_c > costs[j]:
                best_c, best_i = costs[j], idx[j]
        c, i = sharding.select_best(costs, idx)
        if best_i is None:
            assert (c, i) == (np.inf, 0)
        else:
            assert c == best_c and i == best_i


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, cases, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert sharding.is_distributed()
        res = []
        for case in cases:
            cost, idx = case[rank]
            res.append(sharding.exchange_best(cost, idx))
        out[rank] = res
    finally:
        dist.destroy_process_group()


def test_exchange_best_world2_gloo():
    import torch.multiprocessing as mp
    cases = [
        [(3.0, 10), (2.0, 500_010)],                 # plain minimum on rank 1
        [(1.5, 499_999), (1.5, 500_000)],            # tie -> lowest global index
        [(float("nan"), 3), (7.0, 900_000)],         # NaN never wins
        [(float("inf"), 0), (float("inf"), 500_000)],
        [(0.25, 2 ** 40 + 1), (0.5, 5)],             # indices beyond 2**32 survive the float64 carrier
    ]
    expect = [(2.0, 500_010), (1.5, 499_999), (7.0, 900_000), (float("inf"), 0), (0.25, 2 ** 40 + 1)]
    # spawn-context manager: a fork()ed manager process leaves the parent's BLAS thread pool wedged
    mgr = mp.get_context("spawn").Manager()
    out = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, cases, out), nprocs=2, join=True)
    for r in range(2):
        assert [tuple(x) for x in out[r]] == expect


# ---------------------------------------------------------------------------------------------
# sharded cluster-tree build (config 5 on several GPUs): world-size-2 gloo, K7 replaced by the oracle's
# E-step (no GPU here) - checks the all-reduce bookkeeping: every rank ends with the tree a single process builds
# ---------------------------------------------------------------------------------------------
def _tree_worker(rank, world, port, max_level, out):
    import json
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from morphablegraphs_b200 import space_partitioning as sp
        from tests.test_tree_cpu import TreeCase, _oracle_lloyd_level, _flat_to_dict
        sp._lloyd_level = _oracle_lloyd_level
        tc = TreeCase()
        seeds = lambda depth, gids, gcounts: torch.stack([(gcounts * j) // 4 for j in range(4)], dim=1)
        tree = sp.ClusterTree(4, max_level, tc.K).construct(tc.data, seeds=seeds, device="cpu")
        out[rank] = json.dumps(_flat_to_dict(tree.flat))
        # default seeding (distributed k-means++): ranks must still agree with each other
        tree2 = sp.ClusterTree(4, 3, tc.K).construct(tc.data, seed=9, device="cpu")
        out[10 + rank] = json.dumps(_flat_to_dict(tree2.flat))
        out[20 + rank] = int(tree2.build_info["n_kd_nodes"])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("max_level", [6, 2])
def test_sharded_tree_build_world2_gloo(max_level):
    import json
    import torch.multiprocessing as mp
    from tests.test_tree_cpu import TreeCase, _same_cluster_tree
    mgr = mp.get_context("spawn").Manager()
    out = mgr.dict()
    mp.spawn(_tree_worker, args=(2, _free_port(), max_level, out), nprocs=2, join=True)
    tc = TreeCase()
    ref = json.loads(str(tc.g["cluster_tree_seeded%d_json" % max_level]))       # built by the reference's own builder
    for r in range(2):
        _same_cluster_tree(json.loads(out[r]), ref, tol=2e-5)
    assert out[10] == out[11] and out[20] == out[21] == len(tc.data)


# ---------------------------------------------------------------------------------------------
# a generator built WITHOUT a seed must still score the same job on every rank
# ---------------------------------------------------------------------------------------------
def _seed_worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from morphablegraphs_b200 import MotionPrimitiveGenerator
        gen = MotionPrimitiveGenerator(None, {"n_random_samples": 8}, seed=None)     # rank-local entropy
        out[rank] = [gen._step_seed() for _ in range(4)]
        gen2 = MotionPrimitiveGenerator(None, {"n_random_samples": 8}, seed=77)      # explicit: no collective needed
        out[10 + rank] = [gen2._step_seed() for _ in range(2)]
    finally:
        dist.destroy_process_group()


def test_unseeded_generator_shares_step_seeds_world2_gloo():
    import torch.multiprocessing as mp
    mgr = mp.get_context("spawn").Manager()
    out = mgr.dict()
    mp.spawn(_seed_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert list(out[0]) == list(out[1]) and len(set(out[0])) == 4
    assert list(out[10]) == list(out[11])
