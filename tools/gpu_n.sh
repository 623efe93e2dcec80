#!/bin/bash
# N-GPU visit: distributed generator check (exchange fused into the score kernels), the default bench line at 1 and N GPUs.
# usage (under gpurun --gpus N): bash tools/gpu_n.sh <tag> <N>
tag=${1:-rXX}; N=${2:-2}
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29513 tools/dist_check.py > $out/${tag}_dist_check.log 2>&1; echo "dist check rc=$?"; tail -3 $out/${tag}_dist_check.log
timeout 900 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "bench n=$N rc=$?"; tail -3 $out/${tag}_bench_n$N.err
python tools/show_bench.py $out/${tag}_bench_n$N.json
python -c "
import json
d=json.load(open('$out/${tag}_bench_n$N.json'))
print('strong', d.get('strong_scaling'))
for k,v in d.get('workloads',{}).items(): print(k, 'strong', v.get('strong_scaling'))"
