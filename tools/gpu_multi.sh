#!/bin/bash
# N-GPU visit: bench at N (both arms), distributed generator check
tag=${1:-rXX}; N=${2:-2}
out=gpurun_out; mkdir -p $out
nvidia-smi -L > $out/${tag}_smi.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "peer_exchange or mgrd" > $out/${tag}_pytest_peer.log 2>&1; echo "pytest peer rc=$?"; tail -4 $out/${tag}_pytest_peer.log
for n in 1 $N; do
  if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511"; fi
  timeout 600 $L bench.py --gpus $n --steps 20 --warmup 5 > $out/${tag}_bench_S1M_n$n.json 2> $out/${tag}_bench_S1M_n$n.err; echo "S1M n=$n rc=$?"
  timeout 600 $L bench.py --gpus $n --workload L4M --steps 10 --warmup 3 --no-staged > $out/${tag}_bench_L4M_n$n.json 2> $out/${tag}_bench_L4M_n$n.err; echo "L4M n=$n rc=$?"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 3 > $out/${tag}_bench_ref_n$N.json 2> $out/${tag}_bench_ref_n$N.err; echo "ref n=$N rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/dist_check.py > $out/${tag}_dist_check.log 2>&1; echo "dist check rc=$?"; tail -5 $out/${tag}_dist_check.log
for f in $out/${tag}_bench_*_n*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); print(d['n_gpus'], d['value'], d['ms_per_step'], d.get('e2e',{}).get('value'))"; done
