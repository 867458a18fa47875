#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 100 python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c8_plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'onesweep' -s 12 -c 8 -o gpurun_out/c8_full python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c8_ncu.log 2>&1
echo "ncu rc=$?"
