#!/usr/bin/env python
"""Run ONE kernel of libmgb200.so a few times on cuda:0 (the target of ncu captures).
usage: python tools/run_kernel.py <k1|k2tc|k2simt|k3|k4|k5a|k5c|k6|k1k6|k5b|k8|k9|k10|k11|multi> [--shape S|L] [--n N] [--reps R]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kernel")
    ap.add_argument("--shape", default="S")
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    wl = "S1M" if a.shape == "S" else "L4M"
    node, skeleton, mpc, prev = bench.build_case(wl)
    qs = bench.WORKLOADS[wl]["quality_scale"]
    dm = engine.DeviceModel(node, skeleton, 0)
    dc = engine.DeviceConstraints(dm, mpc, prev)
    n = a.n
    counts = dm.component_counts(n, seed=2024)
    S = dm.sample(n, seed=1, counts=counts)
    coeffs = frames = None
    if a.kernel in ("k2tc", "k2simt", "k3"):
        coeffs = torch.empty((n, dm.Nb, dm.P), dtype=torch.float32, device="cuda")
    if a.kernel == "k3":
        dm.backproject(S, impl=2, out=coeffs)
        frames = torch.empty((n, len(dc.times), dm.P), dtype=torch.float32, device="cuda")
    if a.kernel in ("k4", "k5a"):
        frames = dm.eval_frames(dm.backproject(S), dc.times)
        pos = torch.empty((n, len(dc.times), 5, 3), dtype=torch.float32, device="cuda")
    if a.kernel == "k11":
        import ctypes as C
        from morphablegraphs_b200 import _lib
        X = torch.randn((n, 1500), device="cuda") * 30
        mean = X.mean(dim=0).contiguous()
        up = torch.zeros((1500, 1500), dtype=torch.float64, device="cuda")
    if a.kernel == "multi":
        jobs = [dict(dc=engine.DeviceConstraints(dm, mpc, prev), n=max(1, n // 4), seed=5 + j) for j in range(4)]
    seeds8 = dm.sample(8, seed=77)
    if a.kernel == "k5c":
        import ctypes as C
        from morphablegraphs_b200 import _lib
        lib = _lib.load()
        cost = torch.rand(n, device="cuda")
        tk_ws = torch.zeros(int(lib.mgb_score_workspace_bytes(n, 16)), dtype=torch.uint8, device="cuda")
        tk_c = torch.empty(16, dtype=torch.float32, device="cuda")
        tk_i = torch.empty(16, dtype=torch.int64, device="cuda")
    fns = {
        "k1": lambda: dm.sample(n, seed=7, counts=counts, out=S),
        "k2tc": lambda: dm.backproject(S, impl=2, out=coeffs),
        "k2simt": lambda: dm.backproject(S, impl=1, out=coeffs),
        "k3": lambda: dm.eval_frames(coeffs, dc.times, out=frames),
        "k6": lambda: dc.score(S, k=1, quality_scale=qs),
        "k1k6": lambda: dc.sample_score(n, seed=7, counts=counts, k=1, quality_scale=qs),
        "k5b": lambda: dm.logpdf(S),
        "k4": lambda: dm.fk(frames, ["Hips", "LeftHand", "RightHand", "LeftFoot", "RightFoot"], out=pos),
        "k5a": lambda: dc.score(S, k=1, frames=frames),
        "k8": lambda: dc.node_costs(S),
        "k9": lambda: dc.residuals(S),
        "k10": lambda: dc.lm_refine(seeds8, quality_scale=qs, max_iter=50),
        "k11": lambda: _lib.check(_lib.load().mgb_syrk_upper(C.c_void_p(X.data_ptr()), n, 1500, 1500, C.c_void_p(mean.data_ptr()),
                                                             C.c_void_p(up.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))),
        "multi": lambda: engine.sample_score_multi(jobs),
        "k5c": lambda: _lib.check(lib.mgb_topk(C.c_void_p(cost.data_ptr()), n, 0, 16, C.c_void_p(tk_c.data_ptr()), C.c_void_p(tk_i.data_ptr()),
                                               C.c_void_p(tk_ws.data_ptr()), tk_ws.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream))),
    }
    fn = fns[a.kernel]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(a.reps):
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        print("%s shape=%s n=%d: %.4f ms" % (a.kernel, a.shape, n, e0.elapsed_time(e1)))


if __name__ == "__main__":
    main()
