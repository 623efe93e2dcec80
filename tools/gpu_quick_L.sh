#!/bin/bash
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/${tag}_pytest.log
for w in L4M S1M; do
timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-staged --no-cpu-baseline > $out/${tag}_bench_$w.json 2> $out/${tag}_bench_$w.err; echo "bench $w rc=$?"
python -c "
import json; d=json.load(open('$out/${tag}_bench_$w.json')); print('$w', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], {k:round(v['ms'],4) for k,v in d['kernels'].items()})"
done
