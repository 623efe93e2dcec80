#!/usr/bin/env python
"""Host-side pieces of the planner-facing call, microseconds (median of 2000).  usage: python tools/call_micro.py"""
import ctypes as C
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import engine, _lib  # noqa: E402

node, skeleton, mpc, prev = bench.build_case("S1M")
dm = node.motion_primitive.device_model(skeleton)
dc = engine.DeviceConstraints(dm, mpc, prev)
lib = dm.lib


def med(fn, reps=2000):
    out = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        out.append(1e6 * (time.perf_counter() - t0))
    return sorted(out)[len(out) // 2]


counts = np.ascontiguousarray(dm.component_counts(100, 3), dtype=np.int64)
counts_m = np.ascontiguousarray(dm.component_counts(1000000, 3), dtype=np.int64)
cost, idx = C.c_float(), C.c_int64()
row = np.empty(dm.D, dtype=np.float32)
stream = dm._stream()
cp = counts.ctypes.data_as(C.POINTER(C.c_int64))
rp = C.c_void_p(row.ctypes.data)


cpm = counts_m.ctypes.data_as(C.POINTER(C.c_int64))


def raw(n):
    _lib.check(lib.mgb_sample_score_best_sharded(dm.handle, dc.handle, n, 0, cp if n == 100 else cpm, C.c_uint64(3), 1.0, 0.0, 0.0, None, C.byref(cost), C.byref(idx), rp, stream))


def ctx():
    with torch.cuda.device(dm.device):
        pass


print("%-60s %6.1f us" % ("raw ctypes call, n = 100 (launch + kernel + D2H + sync)", med(lambda: raw(100))))
print("%-60s %6.1f us" % ("raw ctypes call, n = 1 000 000", med(lambda: raw(1000000), 300)))
print("%-60s %6.1f us" % ("with torch.cuda.device(dev): pass", med(ctx)))
print("%-60s %6.1f us" % ("dm._stream()", med(dm._stream)))
print("%-60s %6.1f us" % ("np.empty(D) + ctypes pointers + astype(float64)", med(lambda: (np.empty(dm.D, dtype=np.float32).ctypes.data, counts.ctypes.data_as(C.POINTER(C.c_int64)), row.astype(np.float64)))))
print("%-60s %6.1f us" % ("dm.component_counts(1e6, seed)", med(lambda: dm.component_counts(1000000, 5))))
print("%-60s %6.1f us" % ("dc.sample_score_best(100, counts given)", med(lambda: dc.sample_score_best(100, seed=3, counts=counts))))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    raw(100)
e1.record()
torch.cuda.synchronize()
print("%-60s %6.1f us" % ("device time between events per call, n = 100", e0.elapsed_time(e1) * 1e3 / 200))
