#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 100 python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c7_plain.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'onesweep|tile_diff|tile_plan|scan_chained|histogram|preprocess_views' -s 30 -c 40 --csv --log-file gpurun_out/c7_launches.csv python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c7_ncu.log 2>&1
echo "ncu rc=$?"
