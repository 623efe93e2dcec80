#!/bin/bash
# ncu --set full of every kernel the bench line does not already cover (one small run each); raw CSV pages are made here
# on the box so that only text travels back.  usage (under gpurun): bash tools/gpu_ncu_all.sh <tag>
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
for spec in "k1k6 S 1000000 mgb_score_kernel" "k3 S 1000000 mgb_eval" "k4 S 1000000 mgb_fk" "k5a S 1000000 mgb_score_kernel" "k8 S 1000000 mgb_score_kernel" \
            "k9 S 1000000 mgb_residuals" "k10 S 8 mgb_lm" "k11 S 200000 mgb_syrk" "multi S 400 mgb_score_multi" "k2tc S 1000000 mgb_backproject_tc"; do
  set -- $spec
  timeout 300 python tools/run_kernel.py $1 --shape $2 --n $3 --reps 2 > $out/${tag}_run_$1.log 2>&1; echo "$1 plain rc=$? $(grep 'ms$' $out/${tag}_run_$1.log | tail -1)"
  timeout 300 ncu --set full --clock-control none -k regex:"$4" -c 1 --launch-skip 1 -f -o /tmp/${tag}_$1 \
    python tools/run_kernel.py $1 --shape $2 --n $3 --reps 2 > $out/${tag}_ncu_$1.log 2>&1; echo "$1 ncu rc=$?"
  ncu -i /tmp/${tag}_$1.ncu-rep --page raw --csv > $out/${tag}_ncu_$1_raw.csv 2>/dev/null
done
ls -la $out | grep ${tag}_ncu | grep csv
