#!/bin/bash
# one ncu --set full capture of the per-Gaussian kernels (view-batched geometry pass and backward) of a bench run
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0"
timeout 100 $CMD > gpurun_out/ncu_pg_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_pg_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'backward_views|preprocess_views|scan_chained|tile_diff|tile_plan' -s 10 -c 10 -o gpurun_out/ncu_pg $CMD > gpurun_out/ncu_pg.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/ncu_pg.ncu-rep
