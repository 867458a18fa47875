#!/bin/bash
# round 2, GPU call 2: full parity suite (no -x) with the view-batched backward, bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c2_pytest.log
tail -15 gpurun_out/c2_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/c2_bench.json')); print(d['value'], d['e2e']['value'], d.get('dropin',{}).get('value')); print({k:v['ms'] for k,v in d['stages'].items()})"
