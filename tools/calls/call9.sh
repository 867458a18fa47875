#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for v in base noexp base noexp; do
  if [ $v = base ]; then CMD="python bench.py"; else CMD="python tools/bench_alt.py rain_b200/alt/librainrast_$v.so"; fi
  timeout 120 $CMD --steps 25 --warmup 4 --cpu-renders 0 --dropin-iters 60 > gpurun_out/c9_bench_$v.json 2> gpurun_out/c9_bench_$v.err
  python -c "
import json; d=json.load(open('gpurun_out/c9_bench_$v.json')); print('$v', d['value'], d['e2e']['value'], d['dropin']['value'], d['clocks']['sm_mhz'], {k:v['ms'] for k,v in d['stages'].items() if k in ('duplicate','tile_sort','ranges','render_fwd')})"
done
