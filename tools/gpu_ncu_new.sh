#!/bin/bash
# ncu --set full of the kernels added late in round 2: the tiled sampler at D = 103 and the one-pass top-k.
# usage (under gpurun): bash tools/gpu_ncu_new.sh <tag>
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
for spec in "k1 L 1000000 mgb_sample_gemm" "k5c S 4000000 mgb_topk_select"; do
  set -- $spec
  timeout 300 python tools/run_kernel.py $1 --shape $2 --n $3 --reps 3 > $out/${tag}_run_$1.log 2>&1; echo "$1 plain rc=$? $(grep 'ms$' $out/${tag}_run_$1.log | tail -1)"
  timeout 300 ncu --set full --clock-control none -k regex:"$4" -c 2 --launch-skip 1 -f -o $out/${tag}_$1 \
    python tools/run_kernel.py $1 --shape $2 --n $3 --reps 2 > $out/${tag}_ncu_$1.log 2>&1; echo "$1 ncu rc=$?"
  ncu -i $out/${tag}_$1.ncu-rep --page raw --csv > $out/${tag}_ncu_$1_raw.csv 2>/dev/null
done
python tools/call_micro.py 2>&1 | tail -10
