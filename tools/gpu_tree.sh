#!/bin/bash
# GPU visit: tree tests + L4M bench + full ncu captures of the step kernels
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_tree.py -m gpu -x -q > $out/${tag}_pytest_tree.log 2>&1; echo "pytest tree rc=$?"
tail -30 $out/${tag}_pytest_tree.log
timeout 600 python bench.py --workload L4M --steps 10 --warmup 3 > $out/${tag}_bench_L4M.json 2> $out/${tag}_bench_L4M.err; echo "bench L4M rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mgb_score -c 2 -f -o $out/${tag}_full_k1k6_S \
  python tools/run_kernel.py k1k6 --shape S --n 1000000 --reps 2 > $out/${tag}_ncu_full_S.log 2>&1; echo "ncu full S rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mgb_score -c 2 -f -o $out/${tag}_full_k1k6_L \
  python tools/run_kernel.py k1k6 --shape L --n 1000000 --reps 2 > $out/${tag}_ncu_full_L.log 2>&1; echo "ncu full L rc=$?"
ls -la $out | tail
