#!/bin/bash
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
timeout 240 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "large or naturalness or host_entry" > $out/${tag}_pytest_L.log 2>&1; rc=$?; echo "pytest L rc=$rc"; tail -5 $out/${tag}_pytest_L.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 300 python bench.py --workload L4M --steps 10 --warmup 3 --no-staged --no-cpu-baseline > $out/${tag}_bench_L4M.json 2> $out/${tag}_bench_L4M.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('$out/${tag}_bench_L4M.json')); print('L4M', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], {k:round(v['ms'],4) for k,v in d['kernels'].items()})"
timeout 600 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest all rc=$?"; tail -3 $out/${tag}_pytest.log
