#!/usr/bin/env python
"""Cycle breakdown of the fused tensor-core step kernel (group 0 of CTA 0; diagnostic counters in mgb_tc.cu)."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import engine, _lib  # noqa: E402
lib = _lib.load()
node, skeleton, mpc, prev = bench.build_case("L4M")
dm = engine.DeviceModel(node, skeleton, 0)
dc = engine.DeviceConstraints(dm, mpc, prev)
n = 1000000
counts = dm.component_counts(n, seed=2024)
buf = (C.c_longlong * 16)()
names = ["wait a_empty", "draw+fill z", "wait sampling MMA", "refill [s|1]", "wait quad chunks", "drain quad chunks", "program",
         "wait GMM group", "tiles", "total", "g1 wait chunk", "g1 consume chunk", "g1 total", "mma wait t_empty", "mma wait b_full", "mma wait a_full"]
for qs in (0.1, 0.0):
    for rep in range(2):
        dc.sample_score(n, seed=7, counts=counts, k=1, quality_scale=qs)
        torch.cuda.synchronize()
        lib.mgb_debug_tc_cycles(buf)
    v = list(buf)
    tiles = max(v[8], 1)
    print("quality_scale", qs, "tiles", v[8], "cycles/tile", v[9] // tiles)
    for i in list(range(8)) + [10, 11, 12, 13, 14, 15]:
        print("   %-20s %8d cycles/tile  %5.1f%%" % (names[i], v[i] // tiles, 100.0 * v[i] / max(v[9], 1)))
