#!/bin/bash
# One GPU visit of round 2: parity tests, the default bench line, the reference arm, launch list of the same command.
# usage (under gpurun): bash tools/gpu_r2.sh <tag> [ncu]
tag=${1:-r02_x}
out=gpurun_out; mkdir -p $out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $out/${tag}_smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $out/${tag}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"; tail -3 $out/${tag}_bench.err
python tools/show_bench.py $out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "bench ref rc=$?"; cut -c1-300 $out/${tag}_bench_ref.json
if [ "$2" = "ncu" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-tree > $out/${tag}_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mgb_score_tc -c 1 -f -o $out/${tag}_full_L \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --only --no-staged --no-tree > $out/${tag}_ncu_full_L.log 2>&1; echo "ncu full L rc=$?"
fi
ls $out | grep $tag
