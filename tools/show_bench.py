#!/usr/bin/env python
"""One-screen summary of a bench.py JSON line: python tools/show_bench.py <file>"""
import json
import sys

d = json.load(open(sys.argv[1]))


def show(name, x):
    e = x.get("e2e", {})
    g = e.get("generator_call", {})
    print("%-11s value %.4g  ms/step %.4g  roof %.3f (%s)  e2e %.4g  generator_call %.4g ms (%s launches)" % (
        name, x["value"], x["ms_per_step"], x["roofline"]["frac"], x["roofline"]["bound"], e.get("value", float("nan")),
        g.get("ms_per_step", float("nan")), g.get("launches_per_step")))
    ob = e.get("option_batch")
    if ob:
        print("            option batch: %d options x %d samples: %.3f ms in one launch, %.3f ms one call per option" % (
            ob["options"], ob["samples_per_option"], ob["ms_one_launch"], ob["ms_one_call_per_option"]))
    if "cpu_baseline" in x:
        print("            cpu", x["cpu_baseline"].get("value"), x["cpu_baseline"].get("kind"))
    if "kernels" in x:
        print("            ", {k: round(v["ms"], 4) for k, v in x["kernels"].items()})


show(d["config"]["workload"], d)
for k, v in d.get("workloads", {}).items():
    if "roofline" in v:
        show(k, v)
    else:
        print("%-11s value %.4g  %s" % (k, v["value"], {a: b for a, b in v.items() if a not in ("value", "description", "unit", "timing")}))
print(d.get("clocks"))
t = d.get("tree")
if t:
    if "error" in t:
        print("tree        ERROR", t["error"])
    else:
        b = t["cluster_tree_build"]
        print("tree        build %.4f s (%d launches, %d nodes)  K7 pass %s  search %s" % (
            b["seconds"], b["launches"], b["n_nodes"],
            {k: "%.1f us (in loop %.1f)" % (v["ms"] * 1e3, v.get("ms_per_pass_in_loop", 0) * 1e3) for k, v in t["K7_lloyd_iteration"].items()},
            {k: round(v["seconds"] * 1e3, 3) for k, v in t["cluster_tree_search"].items() if isinstance(v, dict) and "seconds" in v}))
