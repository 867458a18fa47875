#!/bin/bash
# one ncu --set full capture of the sort / binning kernels of a bench run (after the same command ran clean)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0"
timeout 100 $CMD > gpurun_out/ncu_sort_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_sort_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'onesweep|duplicate|scan_chained|histogram' -s 40 -c 14 -o gpurun_out/ncu_sort $CMD > gpurun_out/ncu_sort.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/ncu_sort.ncu-rep
