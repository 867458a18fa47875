#!/bin/bash
# C2 and C4 (one GPU) bench lines of the current build
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
T=${1:-v11}
timeout 200 python bench.py --workload C2 --steps 20 --warmup 5 --cpu-renders 0 > gpurun_out/bench_c2_$T.json 2>/dev/null; echo "c2 rc=$?"
timeout 400 python bench.py --workload C4 --steps 3 --warmup 3 --cpu-renders 0 --dropin-iters 20 > gpurun_out/bench_c4_n1_$T.json 2>/dev/null; echo "c4 rc=$?"
for f in c2 c4_n1; do python -c "
import json
d=json.load(open('gpurun_out/bench_${f}_$T.json')); print('$f', d['value'], d['e2e']['value'], d.get('dropin',{}).get('value'), {k:v['ms'] for k,v in d['stages'].items()})"; done
