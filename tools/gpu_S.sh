#!/bin/bash
# S step kernel visit: full GPU tests, S1M bench line, then the same bench with the register cap variants of the kernel
tag=${1:-r02_x}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $out/${tag}_pytest.log
timeout 600 python bench.py --workload S1M --only --no-staged --no-cpu-baseline --steps 30 --warmup 5 > $out/${tag}_bench_S1M.json 2> $out/${tag}_bench_S1M.err; python tools/show_bench.py $out/${tag}_bench_S1M.json | head -3
for mb in 3 4; do
MGB_NVCC_EXTRA=-DMGB_SCORE_MIN_BLOCKS=$mb python -c "from morphablegraphs_b200 import _lib; _lib.build(force=True)" > $out/${tag}_build$mb.log 2>&1
timeout 600 python bench.py --workload S1M --only --no-staged --no-cpu-baseline --steps 30 --warmup 5 > $out/${tag}_bench_S1M_mb$mb.json 2> $out/${tag}_bench_S1M_mb$mb.err; echo "min blocks $mb:"; python tools/show_bench.py $out/${tag}_bench_S1M_mb$mb.json | head -1
done
