#!/bin/bash
# round 2: single-GPU measurement set of the current build (bench C3 / C2 / C4, reference arm, dist2, live lanes, ncu)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
T=${1:-v9}
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
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|onesweep|preprocess|duplicate|scan_chained|tile_|histogram|backward_views|loss' -s 120 -c 28 -o gpurun_out/ncu_full_$T python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 4 > gpurun_out/ncu2_$T.log 2>&1
echo "ncu full rc=$?"
for f in ours reference c2 c2_ref c4_n1 c4_ref; do python -c "
import json
try:
    d=json.load(open('gpurun_out/bench_${f}_$T.json')); print('$f', d['value'], d.get('e2e',{}).get('value'), d.get('dropin',{}).get('value'))
except Exception as e: print('$f', 'ERR', e)"; done
