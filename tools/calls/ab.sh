#!/bin/bash
# A/B of experiment builds (rain_b200/alt/librainrast_<v>.so) against the product library: bash tools/calls/ab.sh v1 v2 ...
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for round in 1 2; do
for v in base "$@"; do
  if [ $v = base ]; then CMD="python bench.py"; else CMD="python tools/bench_alt.py rain_b200/alt/librainrast_$v.so"; fi
  timeout 200 $CMD --steps 20 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/ab_${v}_$round.json 2> gpurun_out/ab_${v}_$round.err
  python -c "
import json; d=json.load(open('gpurun_out/ab_${v}_$round.json')); print('$v', $round, d['value'], d['e2e']['value'], {k:v['ms'] for k,v in d['stages'].items() if k in ('preprocess_fwd','depth_sort','tile_sort','scan','render_fwd','render_bwd','backward_views')})"
done; done
