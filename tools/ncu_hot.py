#!/usr/bin/env python
"""Top stall-sample SASS lines of an ncu report: python tools/ncu_hot.py <rep> [N]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ci = {n: i for i, n in enumerate(h)}
stall_cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
data = []
for idx, r in enumerate(rows[hi + 1:]):
    if len(r) < len(h) or not r[0].startswith("0x"):
        continue
    smp = int(r[ci["# Samples"]] or 0)
    st = sorted(((int(r[ci[c]] or 0), c) for c in stall_cols), reverse=True)[:2]
    data.append((smp, idx, r[ci["Source"]].strip(), r[ci["Instructions Executed"]], st))
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", len(data))
for d in sorted(data, reverse=True)[:top]:
    print("%6d %5.1f%%  #%-5d %-70s exec=%s %s" % (d[0], 100.0 * d[0] / max(tot, 1), d[1], d[2][:70], d[3],
                                                 " ".join("%s=%d" % (c[6:], v) for v, c in d[4] if v)))
