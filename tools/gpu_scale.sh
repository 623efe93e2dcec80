#!/bin/bash
# scaling visit on an N-GPU box: S1M / L4M bench at 1, 2, 4, 8 ranks (device arm only) + distributed generator check
tag=${1:-rXX}
out=gpurun_out; mkdir -p $out
nvidia-smi -L > $out/${tag}_smi.txt
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n))"; fi
  timeout 300 $L bench.py --gpus $n --steps 20 --warmup 5 --no-staged --no-cpu-baseline > $out/${tag}_bench_S1M_n$n.json 2> $out/${tag}_bench_S1M_n$n.err; echo "S1M n=$n rc=$?"
  timeout 300 $L bench.py --gpus $n --workload L4M --steps 10 --warmup 3 --no-staged --no-cpu-baseline > $out/${tag}_bench_L4M_n$n.json 2> $out/${tag}_bench_L4M_n$n.err; echo "L4M n=$n rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 tools/dist_check.py > $out/${tag}_dist_check.log 2>&1; echo "dist check rc=$?"; tail -2 $out/${tag}_dist_check.log
for f in $out/${tag}_bench_*_n*.json; do python -c "
import json
d=json.load(open('$f')); print('$f', d['n_gpus'], '%.4g' % d['value'], '%.4f ms' % d['ms_per_step'], 'e2e %.4g' % d['e2e']['value'], d['config'].get('exchange','')[:40])"; done
