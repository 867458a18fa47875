#!/bin/bash
# round 2, GPU call 4: callers diag, failing tests again, forward variants, ncu of the new forward + batched backward
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 200 python tools/gpu_diag_callers.py > gpurun_out/c4_diag_callers.log 2>&1; echo "diag rc=$?"; tail -8 gpurun_out/c4_diag_callers.log
timeout 600 python -m pytest tests -m gpu -q -k "accumulate_mode or leaf_mode or reference_callers or oracle" > gpurun_out/c4_pytest.log 2>&1; tail -8 gpurun_out/c4_pytest.log
for v in f4 f6 s6; do
  timeout 120 python tools/bench_alt.py rain_b200/alt/librainrast_$v.so --steps 8 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c4_bench_$v.json 2> gpurun_out/c4_bench_$v.err
  python -c "
import json; d=json.load(open('gpurun_out/c4_bench_$v.json')); print('$v', d['value'], d['stages']['render_fwd']['ms'], d['stages']['render_bwd']['ms'])"
done
timeout 120 python bench.py --steps 8 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c4_bench_base.json 2> gpurun_out/c4_bench_base.err
python -c "
import json; d=json.load(open('gpurun_out/c4_bench_base.json')); print('base', d['value'], d['stages']['render_fwd']['ms'], d['stages']['render_bwd']['ms'])"
timeout 200 python tools/live_lanes.py C3 4 > gpurun_out/c4_live_lanes.log 2>&1; echo "live rc=$?"
timeout 100 python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c4_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'render_fwd|backward_views' -s 40 -c 5 -o gpurun_out/c4_full python bench.py --steps 2 --warmup 3 --cpu-renders 0 --dropin-iters 0 > gpurun_out/c4_ncu.log 2>&1
echo "ncu rc=$?"
