#!/bin/bash
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
for spec in "k2tc S 1000000" "k2tc L 250000" "k3 S 1000000" "k1 S 1000000" "k5b L 1000000"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:"mgb_" -c 3 -f -o $out/${tag}_full_$1_$2 \
    python tools/run_kernel.py $1 --shape $2 --n $3 --reps 2 > $out/${tag}_ncu_$1_$2.log 2>&1; echo "$1 $2 rc=$?"; grep "ms$" $out/${tag}_ncu_$1_$2.log | tail -1
done
ls -la $out/${tag}_full_* 
