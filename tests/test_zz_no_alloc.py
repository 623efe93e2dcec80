"""mgb_topk_noalloc: the allocation-free variant of K5c (scratch in a file-scope device array of the library).

No counterpart in the reference.  Checked through the C-ABI against the
existing entry point mgb_topk on the same cost arrays - integer / bit-exact work, so the tolerance is zero - on two
calls in a row (the scratch array keeps the previous call's contents), and past its stated limits.
This file is named so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest


def _costs(n, seed):
    rs = np.random.RandomState(seed)
    cost = rs.randint(0, max(2, n // 3 + 1), size=n).astype(np.float32)      # many ties
    if n > 10:
        cost[rs.randint(0, n, size=n // 10)] = np.nan
        cost[rs.randint(0, n, size=5)] = np.inf
        cost[rs.randint(0, n, size=3)] = -np.inf
    return cost


@pytest.mark.gpu
def test_topk_noalloc_matches_topk_twice_in_a_row():
    import torch
    from morphablegraphs_b200 import _lib
    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    base = 777
    # every buffer is allocated BEFORE the calls under test; the new entry point gets inputs and outputs only
    cases = []
    for call, (n, k) in enumerate([(1_000_003, 64), (4099, 5), (3, 8), (2_500_000, 1)]):
        cost = _costs(n, 100 + call)
        cases.append(dict(n=n, k=k, cost=cost, d_cost=torch.as_tensor(cost, device="cuda"),
                          ws=torch.zeros(int(lib.mgb_score_workspace_bytes(n, k)), dtype=torch.uint8, device="cuda"),
                          bc=[torch.empty(k, dtype=torch.float32, device="cuda") for _ in range(2)],
                          bi=[torch.empty(k, dtype=torch.int64, device="cuda") for _ in range(2)]))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for c in cases:        # consecutive calls on one stream: the second .. fourth find the previous call's partials in the scratch
        rc = lib.mgb_topk_noalloc(C.c_void_p(c["d_cost"].data_ptr()), c["n"], base, c["k"], C.c_void_p(c["bc"][0].data_ptr()),
                                  C.c_void_p(c["bi"][0].data_ptr()), stream)
        assert rc == 0, lib.mgb_last_error().decode()
    assert torch.cuda.memory_allocated() == before
    for c in cases:
        rc = lib.mgb_topk(C.c_void_p(c["d_cost"].data_ptr()), c["n"], base, c["k"], C.c_void_p(c["bc"][1].data_ptr()),
                          C.c_void_p(c["bi"][1].data_ptr()), C.c_void_p(c["ws"].data_ptr()), c["ws"].numel(), stream)
        assert rc == 0, lib.mgb_last_error().decode()
    torch.cuda.synchronize()
    for c in cases:
        got_c, got_i = c["bc"][0].cpu().numpy(), c["bi"][0].cpu().numpy()
        ref_c, ref_i = c["bc"][1].cpu().numpy(), c["bi"][1].cpu().numpy()
        assert np.array_equal(got_i, ref_i), (c["n"], c["k"])
        assert np.array_equal(got_c, ref_c), (c["n"], c["k"])          # bit-exact: selection copies values
        # and both against numpy: (cost, index) ascending, NaN never selected, missing slots = (inf, index_base)
        cost, k = c["cost"], c["k"]
        ok = np.where(~np.isnan(cost))[0]
        order = ok[np.lexsort((ok, cost[ok]))][:k]
        want_i = np.full(k, base, dtype=np.int64)
        want_c = np.full(k, np.inf, dtype=np.float32)
        want_i[:len(order)] = order + base
        want_c[:len(order)] = cost[order]
        assert np.array_equal(got_i, want_i) and np.array_equal(got_c, want_c)


@pytest.mark.gpu
def test_topk_noalloc_rejects_sizes_beyond_its_limit():
    import torch
    from morphablegraphs_b200 import _lib
    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    d = torch.zeros(8, device="cuda")
    bc = torch.empty(64, dtype=torch.float32, device="cuda")
    bi = torch.empty(64, dtype=torch.int64, device="cuda")
    args = lambda n, k: (C.c_void_p(d.data_ptr()), n, 0, k, C.c_void_p(bc.data_ptr()), C.c_void_p(bi.data_ptr()), stream)
    assert lib.mgb_topk_noalloc(*args(_lib.TOPK_NOALLOC_MAX_N + 1, 4)) != 0
    assert b"at most" in lib.mgb_last_error()
    assert lib.mgb_topk_noalloc(*args(8, 65)) != 0
    assert lib.mgb_topk_noalloc(*args(8, 0)) != 0
    assert lib.mgb_topk_noalloc(*args(-1, 1)) != 0
    assert lib.mgb_topk_noalloc(*args(8, 4)) == 0        # still usable after the refusals
    torch.cuda.synchronize()
    assert list(bi[:4].cpu().numpy()) == [0, 1, 2, 3]
