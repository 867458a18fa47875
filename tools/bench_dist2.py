"""BASELINE.json configs[4] (C5): simple-knn distCUDA2 point initialisation on 5M random points ~ U([-1,1]^3), seed 0.
Prints one JSON line: this library (CUDA events, device resident, warm), the reference CUDA rebuilt for sm_100a
(oracle/_ref) on the same GPU, a bounded host sample of the oracle port, and the exactness check of SURVEY.md 8(d)
(queries brute-forced on the host).   python tools/bench_dist2.py [P] [reps]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from rain_b200 import _C
from oracle.oracle_py import Oracle, RefCuda

P = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = torch.device("cuda:0")
pts_h = (torch.rand(P, 3, generator=torch.Generator().manual_seed(0)) * 2 - 1).float()
pts = pts_h.to(dev)


def timed(fn):
    for _ in range(3):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


ms_mine, d_mine = timed(lambda: _C.distCUDA2(pts))
line = {"metric": "distCUDA2 time (5M random points)", "unit": "ms", "higher_is_better": False, "points": P, "reps": reps,
        "value": round(ms_mine, 3), "points_per_s": round(P / ms_mine * 1e3),
        # 12 B in + 4 B out per point + Morton sort (4 B histogram read + 4 passes x 16 B) + 16 B gather
        "alg_bytes": P * (12 + 4 + 4 + 64 + 16)}
line["gbs"] = round(line["alg_bytes"] / ms_mine / 1e6, 1)
try:
    ref = RefCuda()
    ms_ref, d_ref = timed(lambda: ref.dist2(pts))
    line["reference_cuda_ms"] = round(ms_ref, 3)
    line["mismatches_vs_reference"] = int((d_mine != d_ref).sum().item())
except Exception as e:  # the rebuilt reference is optional here
    line["reference_cuda_ms"] = None; line["reference_error"] = str(e)[:100]
orc = Oracle()
idx = np.random.default_rng(0).choice(P, 10_000, replace=False).astype(np.int32)
t = time.time(); dq = orc.dist2_queries(pts_h.numpy(), idx); dt = time.time() - t
line["mismatches_vs_host_brute_force_10k_queries"] = int((d_mine.cpu().numpy()[idx] != dq).sum())
line["cpu_baseline"] = {"value": round(dt / len(idx) * P * 1e3, 1), "unit": "ms (extrapolated from the sample)", "cores": orc.num_threads(), "kind": "port",
                        "sample": f"{len(idx)} brute-force queries over {P} points in {dt:.1f} s, oracle/liboracle.so"}
print(json.dumps(line))
