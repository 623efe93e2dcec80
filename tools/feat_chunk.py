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
