#!/bin/bash
# ncu --set full capture of the per-view kernels (tile renderers, pair emission, tile sort, loss) of a bench run
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
T=${1:-v11}
BCMD="python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0"
timeout 100 $BCMD > gpurun_out/plain_$T.log 2>&1 || { echo "plain run failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'render_|duplicate|onesweep|loss' -s 96 -c 18 -o gpurun_out/ncu_view_$T $BCMD > gpurun_out/ncu2_$T.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/ncu_view_$T.ncu-rep
