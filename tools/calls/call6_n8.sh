#!/bin/bash
# round 2, GPU call 6 (8 GPUs): hardware check of the sharded step, all-reduce probe, C3 weak scaling, C4 (6M Gaussians, 64 views) strong scaling
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29601 tools/verify_multigpu.py 30000 160 96 16 > gpurun_out/c6_verify_n$N.json 2> gpurun_out/c6_verify_n$N.err; echo "verify rc=$?"; grep "^{" gpurun_out/c6_verify_n$N.json
timeout 240 $TR --master-port 29602 tools/probe_symm.py 1000000 > gpurun_out/c6_probe_n$N.json 2> gpurun_out/c6_probe_n$N.err; echo "probe rc=$?"; grep "^{" gpurun_out/c6_probe_n$N.json
timeout 300 $TR --master-port 29603 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/c6_bench_c3_n$N.json 2> gpurun_out/c6_bench_c3_n$N.err; echo "bench C3 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/c6_bench_c3_n$N.json')); print('C3', d['value'], d['e2e']['value'], d['ms_per_step'], d.get('allreduce'))"
timeout 300 $TR --master-port 29604 bench.py --gpus $N --steps 20 --warmup 3 --collective nccl > gpurun_out/c6_bench_c3_nccl_n$N.json 2> gpurun_out/c6_bench_c3_nccl_n$N.err; echo "bench C3 nccl rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/c6_bench_c3_nccl_n$N.json')); print('C3 nccl', d['value'], d['e2e']['value'], d['ms_per_step'], d.get('allreduce'))"
timeout 600 $TR --master-port 29605 bench.py --gpus $N --workload C4 --steps 6 --warmup 3 > gpurun_out/c6_bench_c4_n$N.json 2> gpurun_out/c6_bench_c4_n$N.err; echo "bench C4 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/c6_bench_c4_n$N.json')); print('C4', d['value'], d['e2e']['value'], d['ms_per_step'], d.get('allreduce'))"
timeout 300 $TR --master-port 29606 tools/verify_multigpu.py 6000000 1920 1080 $N > gpurun_out/c6_verify_c4_n$N.json 2> gpurun_out/c6_verify_c4_n$N.err; echo "verify C4-size rc=$?"; grep "^{" gpurun_out/c6_verify_c4_n$N.json
