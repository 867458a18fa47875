#!/bin/bash
# round 2: scaling lines of the final build at N GPUs: hardware check, C3 (weak), C4 (6M Gaussians, 64 views sharded; strong)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29601 tools/verify_multigpu.py 30000 160 96 16 2>/dev/null | grep "^{" > gpurun_out/s_verify_n$N.json; cat gpurun_out/s_verify_n$N.json | cut -c1-200
timeout 300 $TR --master-port 29603 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/s_bench_c3_n$N.json 2> gpurun_out/s_bench_c3_n$N.err; echo "bench C3 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/s_bench_c3_n$N.json')); print('C3', d['value'], d['e2e']['value'], d['ms_per_step'], d['allreduce']['ms_per_step'], d['allreduce']['impl'][:120])"
timeout 600 $TR --master-port 29605 bench.py --gpus $N --workload C4 --steps 6 --warmup 3 > gpurun_out/s_bench_c4_n$N.json 2> gpurun_out/s_bench_c4_n$N.err; echo "bench C4 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/s_bench_c4_n$N.json')); print('C4', d['value'], d['e2e']['value'], d['ms_per_step'], d['allreduce']['ms_per_step'])"
