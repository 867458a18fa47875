#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
i=0
for opt in "" "--overlap-chunks 2" "--overlap-chunks 2 --overlap-no-room" "--overlap-chunks 4 --overlap-no-room" "--collective nccl"; do
  i=$((i+1))
  timeout 200 $TR --master-port $((29800+i)) bench.py --gpus $N --steps 20 --warmup 3 $opt > gpurun_out/c12_$i.json 2> gpurun_out/c12_$i.err
  python -c "
import json; d=json.load(open('gpurun_out/c12_$i.json')); print('[$opt]', d['value'], d['e2e']['value'], d['ms_per_step'], d['allreduce']['ms_per_step'])"
done
RAIN_VERIFY_OVERLAP=4 timeout 240 $TR --master-port 29811 tools/verify_multigpu.py 30000 160 96 16 > gpurun_out/c12_verify.json 2>/dev/null; grep -o '"ok": [a-z]*' gpurun_out/c12_verify.json
