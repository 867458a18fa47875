#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29701 tools/verify_multigpu.py 30000 160 96 16 > gpurun_out/c11_verify_n$N.json 2> gpurun_out/c11_verify_n$N.err; echo "verify rc=$?"; grep "^{" gpurun_out/c11_verify_n$N.json
timeout 240 $TR --master-port 29702 tools/verify_multigpu.py 30001 100 52 5 > gpurun_out/c11_verify_odd_n$N.json 2> gpurun_out/c11_verify_odd_n$N.err; echo "verify odd rc=$?"; grep "^{" gpurun_out/c11_verify_odd_n$N.json
timeout 300 $TR --master-port 29703 tools/verify_multigpu.py 1000000 1920 1080 8 > gpurun_out/c11_verify_c3_n$N.json 2> gpurun_out/c11_verify_c3_n$N.err; echo "verify C3 rc=$?"; grep "^{" gpurun_out/c11_verify_c3_n$N.json
timeout 300 $TR --master-port 29704 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/c11_bench_c3_n$N.json 2> gpurun_out/c11_bench_c3_n$N.err; echo "bench rc=$?"; tail -3 gpurun_out/c11_bench_c3_n$N.err
python -c "
import json; d=json.load(open('gpurun_out/c11_bench_c3_n$N.json')); print('C3', d['value'], d['e2e']['value'], d['ms_per_step'], d.get('allreduce'))"
timeout 300 $TR --master-port 29705 bench.py --gpus $N --steps 20 --warmup 3 --collective nccl > gpurun_out/c11_bench_c3_nccl_n$N.json 2> gpurun_out/c11_bench_c3_nccl_n$N.err
python -c "
import json; d=json.load(open('gpurun_out/c11_bench_c3_nccl_n$N.json')); print('C3 nccl', d['value'], d['e2e']['value'], d['ms_per_step'], d.get('allreduce'))"
