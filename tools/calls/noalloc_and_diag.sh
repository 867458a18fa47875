#!/bin/bash
# the no-allocation sort test alone (fresh process), then the backward run-to-run spread diagnostic
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -5 gpurun_out/noalloc_alone.log
timeout 400 python tools/gpu_diag_bwd.py 4 > gpurun_out/diag_bwd.log 2>&1; echo "diag rc=$?"; tail -3 gpurun_out/diag_bwd.log
