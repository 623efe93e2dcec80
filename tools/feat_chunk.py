#!/usr/bin/env python
"""Euclidean point clouds of 1M samples (K2 -> K3 -> K4) as a function of the chunk size: with a chunk whose frames fit the L2
the intermediate arrays never reach HBM.  usage: python tools/feat_chunk.py"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import engine, features as feat  # noqa: E402

node, skeleton, mpc, prev = bench.build_case("S1M")
dm = engine.DeviceModel(node, skeleton, 0)
S = dm.sample(1000000, seed=2024)
out = None
for chunk in (32768, 16384, 8192, 4096, 2048, 1024):
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = feat.map_samples_to_euclidean_space(dm, S, chunk=chunk, out=out)
        torch.cuda.synchronize()
        t = time.perf_counter() - t0
    print("chunk %6d: %.1f ms" % (chunk, t * 1e3))

# per-stage device time of one 32768-sample chunk
import numpy as np  # noqa: E402
from morphablegraphs_b200.features import END_EFFECTORS2  # noqa: E402
m = 32768
coeffs = torch.empty((m, dm.Nb, dm.P), dtype=torch.float32, device="cuda")
frames = torch.empty((m, dm.F, dm.P), dtype=torch.float32, device="cuda")
pos = torch.empty((m, dm.F, len(END_EFFECTORS2), 3), dtype=torch.float32, device="cuda")
times = np.arange(dm.F, dtype=np.float64)
stages = [("K2 backproject", lambda: dm.backproject(S[:m], out=coeffs), coeffs.numel() * 4),
          ("K3 eval_frames (50 frames)", lambda: dm.eval_frames(coeffs, times, out=frames), (coeffs.numel() + frames.numel()) * 4),
          ("K4 fk (50 frames x %d joints)" % len(END_EFFECTORS2), lambda: dm.fk(frames, END_EFFECTORS2, out=pos), (frames.numel() + pos.numel()) * 4)]
for name, fn, byts in stages:
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%-34s %.3f ms  %.2f TB/s of its own traffic (%.0f MB)" % (name, ms, byts / ms * 1e-9, byts * 1e-6))
