#!/bin/bash
# does the clock-polling nvidia-smi process disturb the timed steps?  the same bench command under four polling set-ups, 7 runs each
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
OUT=gpurun_out/smi_polling.txt
: > $OUT
for cfg in "50 1" "100 1" "100 0" "0 0"; do
  set -- $cfg
  for i in 1 2 3 4 5 6 7; do
    RAIN_BENCH_SMI_MS=$1 RAIN_BENCH_SMI_FULL=$2 timeout 120 python bench.py --steps 20 --warmup 5 --cpu-renders 0 --dropin-iters 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('period_ms=$1 full_line=$2 run=$i value', d['value'], 'e2e', d['e2e']['value'], 'samples', d['clocks'].get('samples'))" >> $OUT
  done
done
cat $OUT
