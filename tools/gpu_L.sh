#!/bin/bash
# Quick visit for the L step kernel: tensor-core parity tests, L4M + stress bench lines (device only), cycle counters of a profile build.
# usage (under gpurun): bash tools/gpu_L.sh <tag> [tests-pattern]
tag=${1:-r02_x}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $out/${tag}_pytest.log
for w in L4M L4M_stress; do
timeout 600 python bench.py --workload $w --only --no-staged --no-cpu-baseline --steps 10 --warmup 3 > $out/${tag}_bench_$w.json 2> $out/${tag}_bench_$w.err; echo "bench $w rc=$?"; tail -3 $out/${tag}_bench_$w.err
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_$w.json'))
print('$w', 'value %.4g'%d['value'], 'ms %.4g'%d['ms_per_step'], 'roof %.3f'%d['roofline']['frac'], 'e2e %.4g'%d['e2e']['value'], 'gen ms %.4g'%d['e2e']['generator_call']['ms_per_step'])
PY
done
MGB_NVCC_EXTRA=-DMGB_TC_PROFILE python -c "from morphablegraphs_b200 import _lib; _lib.build(force=True)" > $out/${tag}_build.log 2>&1
timeout 300 python tools/tc_cycles.py > $out/${tag}_tc_cycles.txt 2>&1; cat $out/${tag}_tc_cycles.txt
