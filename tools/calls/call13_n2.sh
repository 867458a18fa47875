#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
i=0
for opt in "--collective rr" "--collective nccl"; do
  i=$((i+1))
  RAIN_BENCH_DIAG=artrace timeout 200 $TR --master-port $((29830+i)) bench.py --gpus 2 --steps 20 --warmup 3 $opt > gpurun_out/c13_$i.json 2> gpurun_out/c13_$i.err
  python -c "
import json; d=json.load(open('gpurun_out/c13_$i.json')); print('[$opt]', d['value'], d['ms_per_step'], d['allreduce']['ms_per_step'])"
done
for f in gpurun_out/artrace_*; do echo $f; cat $f; done
