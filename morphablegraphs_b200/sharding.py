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


def broadcast_seed(seed, group=None):
    """Rank 0's seed on every rank (one int64 broadcast; NCCL on the GPUs, gloo in the CPU tests)."""
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    t = torch.tensor([int(seed)], dtype=torch.int64, device=dev)
    dist.broadcast(t, src=0, group=group)
    return int(t.item())


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


def default_spin_limit():
    """Polls (40 ns sleeps) a rank waits for a peer's record before it raises instead of hanging: ~10 s by default - a rank may be
    late by a first-call CUDA initialisation or a constraint compile -, MGB_EXCHANGE_SPIN_LIMIT overrides it."""
    import os
    return int(os.environ.get("MGB_EXCHANGE_SPIN_LIMIT", "250000000"))


class PeerExchange(object):
    """The same exchange without a library collective on the step's critical path: every rank stores its
    16-byte record into every peer's slot array through NVLink peer memory and reduces the records that
    arrive in its own (csrc/mgb_exchange.cu, `mgb_exchange_best`) - one 32-thread launch per step, results
    stay on the device.  Peer pointers come from torch's symmetric-memory allocator (CUDA VMM handles
    exchanged through the process group's store); construction is collective.  Raises if symmetric memory
    is not available - callers then stay on `exchange_best` (NCCL), which is still a GPU path."""

    def __init__(self, device, group=None, slots=None, peer_ptrs=None, rank=None, world=None):
        import ctypes as C
        import torch
        from . import _lib
        self._C, self.torch, self.lib = C, torch, _lib.load()
        self.device = torch.device(device)
        if peer_ptrs is None:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
            nbytes = int(self.lib.mgb_exchange_slot_bytes(self.world))
            self.slots = symm.empty(nbytes // 4, dtype=torch.int32, device=self.device)
            self.slots.zero_()
            g = group if group is not None else dist.group.WORLD
            self.handle = symm.rendezvous(self.slots, g.group_name if hasattr(g, "group_name") else g)
            peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
            torch.cuda.synchronize(self.device)
            dist.barrier(group)                      # every rank's slots are zeroed before anyone publishes
        else:                                        # tests: several "ranks" inside one process
            self.rank, self.world, self.slots, self.handle = int(rank), int(world), slots, None
        self.peer_ptrs = torch.tensor(peer_ptrs, dtype=torch.int64, device=self.device)
        self.out_cost = torch.empty(1, dtype=torch.float32, device=self.device)
        self.out_idx = torch.empty(1, dtype=torch.int64, device=self.device)
        self.err = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.seq = 0

    def exchange(self, best_cost, best_idx, spin_limit=0):
        """best_cost float32[>=1], best_idx int64[>=1] device tensors of this rank -> (out_cost[1], out_idx[1])
        device tensors holding the job-wide winner; asynchronous on the current stream."""
        C, torch = self._C, self.torch
        from . import _lib
        self.seq += 1
        if self.seq >= 2 ** 32:
            self.seq = 2            # 2**32 - 1 is odd: wrapping to an even value keeps the slot parity alternating
        p = lambda t: C.c_void_p(t.data_ptr())
        with torch.cuda.device(self.device):
            _lib.check(self.lib.mgb_exchange_best(p(best_cost), p(best_idx), p(self.peer_ptrs), self.rank, self.world,
                                                  C.c_uint32(self.seq), C.c_int64(int(spin_limit) or default_spin_limit()), p(self.out_cost), p(self.out_idx),
                                                  p(self.err), C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return self.out_cost, self.out_idx

    def next_desc(self, spin_limit=0):
        """mgb_exchange_desc of the NEXT step, for the entries that run the exchange inside the score kernel's epilogue
        (mgb_sample_score_sharded / mgb_sample_score_best_sharded): advances the shared step counter exactly like
        `exchange` does, so the two forms may be mixed as long as every rank makes the same sequence of calls."""
        from . import _lib
        self.seq += 1
        if self.seq >= 2 ** 32:
            self.seq = 2
        d = _lib.ExchangeDesc()
        d.d_peer_slots = self.peer_ptrs.data_ptr()
        d.rank, d.world, d.seq, d.spin_limit = self.rank, self.world, self.seq, int(spin_limit) or default_spin_limit()
        d.d_err = self.err.data_ptr()
        return d

    def check(self):
        """Synchronising: raises if a peer's record never arrived."""
        e = int(self.err.item())
        if e:
            self.err.zero_()        # one timeout must not poison every later step
            from ._lib import MgbError
            raise MgbError("peer exchange timed out waiting for rank %d" % (e - 1))


_PEERS = {}


def peer_exchange(device):
    """Process-wide PeerExchange of `device` (created collectively on first use by every rank at the same
    point of the program), or None when peer memory cannot be set up on ANY rank - then every rank uses
    the NCCL all-gather."""
    import torch
    import torch.distributed as dist
    key = str(device)
    if key not in _PEERS:
        try:
            p = PeerExchange(device)
        except Exception:
            p = None
        flag = torch.tensor([1.0 if p is not None else 0.0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        _PEERS[key] = p if float(flag.item()) == 1.0 else None
    return _PEERS[key]
