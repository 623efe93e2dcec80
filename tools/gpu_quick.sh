#!/bin/bash
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $out/${tag}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-staged > $out/${tag}_bench_S1M.json 2> $out/${tag}_bench_S1M.err; echo "bench S1M rc=$?"
python -c "
import json; d=json.load(open('$out/${tag}_bench_S1M.json')); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], {k:v['ms'] for k,v in d['kernels'].items()})"
