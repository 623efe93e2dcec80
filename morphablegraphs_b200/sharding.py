"""Multi-GPU sharding of the sample-and-evaluate step.

Candidates are independent, so rank r of R owns the global indices
[base_r, base_r + n_r) of the step's job; the sampler keys Philox by the global
index and takes the job-wide per-component counts, so the drawn candidates do not
depend on R.  The ONLY exchange on the path is every rank's best (cost, global
index): one all-gather of 16 bytes per rank (NCCL on the GPUs, gloo in the CPU
tests), after which every rank applies the reference's selection rule
(`min_error > error` in a forward loop: first minimum wins, NaN never wins;
reference motion_primitive_generator.py:251-257) to the gathered pairs.
"""
import math

import numpy as np


def shard_range(n_total, rank, world):
    """(index_base, n) of rank's contiguous share; shares differ by at most one candidate."""
    n_total, rank, world = int(n_total), int(rank), int(world)
    q, r = divmod(n_total, world)
    base = rank * q + min(rank, r)
    return base, q + (1 if rank < r else 0)


def select_best(costs, indices):
    """First-minimum rule over (cost, global index) pairs: lowest cost, lowest index on ties,
    NaN never wins; with nothing finite the reference keeps index 0 / inf."""
    best_c, best_i = math.inf, None
    for c, i in zip(costs, indices):
        c, i = float(c), int(i)
        if c != c:
            continue
        if best_i is None or c < best_c or (c == best_c and i < best_i):
            best_c, best_i = c, i
    if best_i is None:
        return math.inf, 0
    return best_c, best_i


def is_distributed():
    try:
        import torch.distributed as dist
    except ImportError:  # pragma: no cover
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def exchange_best(cost, index, device=None, group=None):
    """All-gather this rank's best (cost, global index); every rank returns the same winner.
    float64 carries indices exactly up to 2**53."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    mine = torch.tensor([float(cost), float(index)], dtype=torch.float64, device=device)
    gathered = torch.empty((world, 2), dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(gathered.view(-1), mine, group=group)
    g = gathered.cpu().numpy()
    return select_best(g[:, 0], g[:, 1].astype(np.int64))
