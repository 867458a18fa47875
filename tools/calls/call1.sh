#!/bin/bash
# round 2, GPU call 1: parity suite, bench baseline of the cleaned-up tree, oracle-agreement diagnostics, live lanes, ncu
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench rc=$?"
timeout 200 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/c1_bench_ref.json 2> gpurun_out/c1_bench_ref.err; echo "ref rc=$?"
timeout 200 python tools/gpu_diag_oracle.py C1 C2 > gpurun_out/c1_diag_oracle.log 2>&1; echo "diag rc=$?"
timeout 200 python tools/live_lanes.py C3 4 > gpurun_out/c1_live_lanes.log 2>&1; echo "live rc=$?"
timeout 280 python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c1_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/c1_launches.csv python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c1_ncu1.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|onesweep|preprocess|duplicate|scan_chained|tile_ranges|histogram' -s 300 -c 40 -o gpurun_out/c1_full python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c1_ncu2.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out | tail -20
