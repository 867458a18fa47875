#!/bin/bash
# round 2: single-GPU measurement set of the current build (bench C3 / C2 / C4, reference arm, dist2, live lanes, ncu)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
T=${1:-v10}
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_ours_$T.json 2> gpurun_out/bench_ours_$T.err; echo "bench rc=$?"
timeout 200 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_reference_$T.json 2> gpurun_out/bench_ref_$T.err; echo "ref rc=$?"
timeout 200 python bench.py --workload C2 --steps 20 --warmup 3 > gpurun_out/bench_c2_$T.json 2> gpurun_out/bench_c2_$T.err; echo "c2 rc=$?"
timeout 200 python bench.py --impl reference --workload C2 --steps 5 --warmup 2 > gpurun_out/bench_c2_ref_$T.json 2>/dev/null
timeout 400 python bench.py --workload C4 --steps 3 --warmup 3 --cpu-renders 1 --dropin-iters 20 > gpurun_out/bench_c4_n1_$T.json 2> gpurun_out/bench_c4_$T.err; echo "c4 rc=$?"
timeout 300 python bench.py --impl reference --workload C4 --steps 1 --warmup 1 > gpurun_out/bench_c4_ref_$T.json 2>/dev/null
timeout 200 python tools/bench_dist2.py > gpurun_out/bench_c5_$T.json 2> gpurun_out/bench_c5_$T.err; echo "c5 rc=$?"
timeout 200 python tools/live_lanes.py C3 4 > gpurun_out/live_lanes_$T.log 2>&1; echo "live rc=$?"
timeout 100 python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 4 > gpurun_out/plain_$T.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/ncu_launches_$T.csv python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 4 > gpurun_out/ncu1_$T.log 2>&1
echo "ncu launches rc=$?"
NCU="ncu --set full --clock-control none --import-source on"
BCMD="python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0"
# two captures (one report with every kernel of a step and its source pages does not fit the 64 MiB that travel back):
# the per-view kernels (tile renderers, pair emission, tile sort, loss) and the per-batch kernels (geometry pass, depth sort, scan, backward)
timeout 600 $NCU -k regex:'render_|duplicate|onesweep|loss' -s 96 -c 18 -o gpurun_out/ncu_view_$T $BCMD > gpurun_out/ncu2_$T.log 2>&1; echo "ncu per-view rc=$?"
timeout 600 $NCU -k regex:'preprocess_views|tile_diff|tile_plan|histogram|scan_chained|backward_views|zero_segments' -s 22 -c 12 -o gpurun_out/ncu_batch_$T $BCMD > gpurun_out/ncu3_$T.log 2>&1
echo "ncu per-batch rc=$?"; ls -la gpurun_out/*.ncu-rep
for f in ours reference c2 c2_ref c4_n1 c4_ref; do python -c "
import json
try:
    d=json.load(open('gpurun_out/bench_${f}_$T.json')); print('$f', d['value'], d.get('e2e',{}).get('value'), d.get('dropin',{}).get('value'))
except Exception as e: print('$f', 'ERR', e)"; done
