#!/bin/bash
# round 2, GPU call 3 (2 GPUs): NCCL algorithm log, symmetric-memory / multicast probe, hardware check of the sharded step, N=2 bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/c3_topo.txt 2>&1
timeout 200 python tools/gpu_diag_callers.py > gpurun_out/c3_diag_callers.log 2>&1; tail -14 gpurun_out/c3_diag_callers.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING,COLL timeout 200 $TR --master-port 29501 tools/probe_symm.py 1000000 > gpurun_out/c3_probe.json 2> gpurun_out/c3_probe_nccl.log; echo "probe rc=$?"
grep -E "NVLS|Ring|Tree|Algo|algo|nvls" gpurun_out/c3_probe_nccl.log | head -40 > gpurun_out/c3_nccl_algo.txt
cat gpurun_out/c3_probe.json
NCCL_ALGO=NVLS timeout 200 $TR --master-port 29502 tools/probe_symm.py 1000000 > gpurun_out/c3_probe_nvls.json 2> gpurun_out/c3_probe_nvls.err; echo "probe nvls rc=$?"; cat gpurun_out/c3_probe_nvls.json
timeout 300 $TR --master-port 29503 tools/verify_multigpu.py 30000 160 96 8 > gpurun_out/c3_verify.json 2> gpurun_out/c3_verify.err; echo "verify rc=$?"; cat gpurun_out/c3_verify.json
timeout 300 $TR --master-port 29504 tools/verify_multigpu.py 1000000 1920 1080 8 > gpurun_out/c3_verify_c3.json 2> gpurun_out/c3_verify_c3.err; echo "verify C3 rc=$?"; cat gpurun_out/c3_verify_c3.json
timeout 300 python -m pytest tests/test_multigpu_nccl.py -m gpu -q > gpurun_out/c3_pytest_nccl.log 2>&1; tail -3 gpurun_out/c3_pytest_nccl.log
timeout 300 $TR --master-port 29505 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c3_bench_n2.json 2> gpurun_out/c3_bench_n2.err; echo "bench n2 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/c3_bench_n2.json')); print(d['value'], d['e2e']['value'], d.get('allreduce'))"
