"""Hardware check of the multi-GPU result (-m gpu; skipped on a box with fewer than 2 GPUs): tools/verify_multigpu.py under
torch.distributed.run with 2 ranks over NCCL.  V views sharded over the ranks + one all-reduce must equal a single-rank
accumulation of all V views (1e-5), replicas must be bit-identical, and the cross-rank densification statistics must equal
the single-rank ones — the same assertion tests/test_multiview_gloo.py makes with a test double on the CPU."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("P,W,H,V", [(30000, 160, 96, 8), (30001, 100, 52, 5)])
def test_two_ranks_nccl_allreduce_equals_single_rank_accumulation(device, P, W, H, V):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tools", "verify_multigpu.py"), str(P), str(W), str(H), str(V)],
                       capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rep = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert rep["ok"] and rep["replicas_bit_identical"] and rep["world"] == 2
