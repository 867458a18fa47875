#!/bin/bash
# sort-focused check: the sort / binning / dist2 parity tests, then the bench line
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "sort or dist2 or knn or no_alloc or configs_match_reference or golden or batched" > gpurun_out/sort_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/sort_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/sort_bench.json 2> gpurun_out/sort_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/sort_bench.err
python -c "
import json; d=json.load(open('gpurun_out/sort_bench.json')); print(d['value'], d['e2e']['value'], d.get('dropin',{}).get('value')); print({k:v['ms'] for k,v in d['stages'].items()})"
timeout 200 python tools/bench_dist2.py > gpurun_out/sort_c5.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/sort_c5.json')); print('C5', d['value'])"
