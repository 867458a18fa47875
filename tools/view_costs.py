"""Per-camera cost of one fwd+bwd render of a workload's scene on one GPU, and what it means for view sharding: the slowest
rank's share of a 64-view step under block assignment (rank r renders cameras r*k .. r*k+k-1) and under strided assignment
(rank r renders cameras r, r+N, r+2N, ...).     python tools/view_costs.py [C4]    -> one JSON line"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from rain_b200 import multiview as mv
from rain_b200.loss import l2_loss_grad
from rain_b200.synthetic import CONFIGS, make_camera, make_grad, make_scene

wl = sys.argv[1] if len(sys.argv) > 1 else "C4"
cfg = CONFIGS[wl]; P, W, H, D = cfg["P"], cfg["W"], cfg["H"], cfg["D"]
dev = torch.device("cuda:0")
sc = make_scene(P, D, seed=0).to(dev)
NC = 64
cams = [make_camera(v, NC, W, H).to(dev) for v in range(NC)]
tgt = make_grad(W, H, seed=100).mul_(0.1).add_(0.5).clamp_(0, 1).mul_(255).round_().to(torch.uint8).to(dev)
params = dict(means3D=sc.means3D, opacities=sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
bucket = mv.GradientBucket(P, 16, dev)
bg = torch.zeros(3, device=dev)
grad = lambda v, color, depth: l2_loss_grad(color, tgt)
cost, nr = [], []
for v in range(NC):
    best = 1e9
    for rep in range(3):
        bucket.filled = False
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        res = mv.render_views_fwd_bwd(params, cams, [v], bucket, grad, bg=bg, sh_degree=D, keep_outputs=False, pipeline="one_stream")
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    cost.append(best); nr.append(int(res[-1].num_rendered))
out = {"workload": wl, "ms_per_view": [round(c, 3) for c in cost], "num_rendered": nr, "sum_ms": round(sum(cost), 2)}
for N in (2, 4, 8):
    k = NC // N
    block = max(sum(cost[r * k:(r + 1) * k]) for r in range(N))
    strided = max(sum(cost[r::N]) for r in range(N))
    out[f"n{N}"] = {"ideal_ms": round(sum(cost) / N, 2), "block_max_ms": round(block, 2), "strided_max_ms": round(strided, 2)}
print(json.dumps(out))
