#!/bin/bash
# K7 check: tree tests, tree bench, one ncu capture of the level kernel.  usage (under gpurun): bash tools/gpu_k7.sh <tag>
tag=${1:-k7}
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_tree.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python tools/tree_bench.py --no-features > $out/${tag}_tree.json 2> $out/${tag}_tree.err; tail -2 $out/${tag}_tree.err
python - <<PYEOF
import json
d = json.load(open('$out/${tag}_tree.json'))
for k, v in d['K7_lloyd_iteration'].items():
    print(k, round(v['ms'] * 1000, 1), 'us', round(v['hbm_frac'], 3), 'in loop', round(v.get('ms_per_pass_in_loop', 0) * 1000, 1), 'us')
b = d['cluster_tree_build']
print('build', b['seconds'], 'launches', b['launches'], 'nodes', b['n_nodes'])
PYEOF
timeout 600 ncu --set full --clock-control none --import-source on -k regex:kmeans_level -s 3 -c 2 -f -o $out/${tag}_k7 python tools/tree_bench.py --no-features > $out/${tag}_k7_ncu.log 2>&1; echo "ncu rc=$?"
