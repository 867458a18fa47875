#!/bin/bash
# full GPU parity suite, then the default bench line: gpurun --timeout 1500 -- "bash tools/calls/tests_and_bench.sh"
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c5_pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/c5_pytest.log | tail -20
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/c5_bench.json 2> gpurun_out/c5_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/c5_bench.err
python -c "
import json; d=json.load(open('gpurun_out/c5_bench.json')); print(d['value'], d['e2e']['value'], d.get('dropin',{}).get('value')); print({k:v['ms'] for k,v in d['stages'].items()})"
