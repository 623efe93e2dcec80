#!/bin/bash
# L step kernel visit + config-5 build numbers.  usage (under gpurun): bash tools/gpu_t.sh <tag>
tag=${1:-r02_x}
bash tools/gpu_L.sh $tag 2>&1 | head -24
python tools/tree_bench.py --no-features > gpurun_out/${tag}_tree.json 2> gpurun_out/${tag}_tree.err; tail -2 gpurun_out/${tag}_tree.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_tree.json"))
print(d["K7_lloyd_iteration"])
b = d["cluster_tree_build"]
print("build s", b["seconds"], "first", b["first_build_seconds"], "launches", b["launches"], "nodes", b["n_nodes"])
PY
