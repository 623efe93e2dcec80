#!/bin/bash
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $out/${tag}_pytest.log
timeout 900 python tools/tree_bench.py > $out/${tag}_tree_bench.json 2> $out/${tag}_tree_bench.err; echo "tree bench rc=$?"; tail -3 $out/${tag}_tree_bench.err
timeout 600 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_S1M.json 2> $out/${tag}_bench_S1M.err; echo "bench S1M rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:kmeans_level -c 3 -f -o $out/${tag}_full_k7 \
  python tools/tree_bench.py --m 1000000 --no-features --cpu-m 0 > $out/${tag}_ncu_full_k7.log 2>&1; echo "ncu k7 rc=$?"
ls -la $out | tail -8
