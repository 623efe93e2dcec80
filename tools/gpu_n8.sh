#!/bin/bash
# 8-GPU visit: S1M / L4M bench at 8 ranks (device arm), distributed generator check, sharded tree check
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $L --master-port 29528 bench.py --gpus 8 --steps 20 --warmup 5 --no-staged --no-cpu-baseline > $out/${tag}_bench_S1M_n8.json 2> $out/${tag}_bench_S1M_n8.err; echo "S1M n=8 rc=$?"
timeout 300 $L --master-port 29529 bench.py --gpus 8 --workload L4M --steps 10 --warmup 3 --no-staged --no-cpu-baseline > $out/${tag}_bench_L4M_n8.json 2> $out/${tag}_bench_L4M_n8.err; echo "L4M n=8 rc=$?"
timeout 200 $L --master-port 29540 tools/dist_check.py > $out/${tag}_dist_check.log 2>&1; echo "dist check rc=$?"; tail -2 $out/${tag}_dist_check.log
timeout 300 $L --master-port 29541 tools/dist_tree_check.py > $out/${tag}_dist_tree.log 2>&1; echo "dist tree rc=$?"; tail -3 $out/${tag}_dist_tree.log
for f in $out/${tag}_bench_*_n8.json; do python -c "
import json
d=json.load(open('$f')); print('$f', d['n_gpus'], '%.4g' % d['value'], '%.4f ms' % d['ms_per_step'], 'e2e %.4g' % d['e2e']['value'], 'gen %.4g' % d['e2e']['generator_call']['value'])"; done
