#!/bin/bash
# One GPU visit: parity tests, bench lines, ncu launch lists and full captures of the step kernels.
# usage (under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $out/${tag}_smi.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $out/${tag}_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_S1M.json 2> $out/${tag}_bench_S1M.err; echo "bench S1M rc=$?"
timeout 600 python bench.py --workload L4M --steps 10 --warmup 3 > $out/${tag}_bench_L4M.json 2> $out/${tag}_bench_L4M.err; echo "bench L4M rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > $out/${tag}_bench_ref_S1M.json 2> $out/${tag}_bench_ref.err; echo "bench ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_S1M.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_S1M.log 2>&1; echo "ncu list S rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_L4M.csv \
  python bench.py --workload L4M --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_L4M.log 2>&1; echo "ncu list L rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mgb_score -c 2 -f -o $out/${tag}_full_k1k6_S \
  python tools/run_kernel.py k1k6 --shape S --n 1000000 --reps 2 > $out/${tag}_ncu_full_S.log 2>&1; echo "ncu full S rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mgb_score -c 2 -f -o $out/${tag}_full_k1k6_L \
  python tools/run_kernel.py k1k6 --shape L --n 1000000 --reps 2 > $out/${tag}_ncu_full_L.log 2>&1; echo "ncu full L rc=$?"
ls -la $out | tail -20
