#!/bin/bash
# where do the occasional slow timed regions come from?  12 runs with 3 warm-up steps, per-step device times of each
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
OUT=gpurun_out/steptimes.txt
: > $OUT
for i in 1 2 3 4 5 6 7 8 9 10 11 12 13 14 15 16; do
  RAIN_BENCH_DIAG=steptimes timeout 120 python bench.py --steps 20 --warmup 3 --cpu-renders 0 --dropin-iters 0 2> gpurun_out/st.err | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('run=$i value', d['value'], 'e2e', d['e2e']['value'], d['clocks'])" >> $OUT
  grep steptimes_ms gpurun_out/st.err >> $OUT
done
cat $OUT | cut -c1-260
