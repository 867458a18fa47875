"""Parameter step at 1M Gaussians (59 floats each): rain_b200.optim.GaussianParamStep (csrc/param_step.cu) against what the
reference does per iteration for the same result — autograd backward through the activations, torch.optim.Adam.step()
over six groups, and the activations of the next render (scene/gaussian_model.py:85-105,139-157).  One JSON line."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from rain_b200.optim import GaussianParamStep, GROUPS

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
P, M = 1_000_000, 16
dev = torch.device("cuda:0")
gen = torch.Generator().manual_seed(0)
host = dict(xyz=torch.randn(P, 3, generator=gen), f_dc=torch.randn(P, 1, 3, generator=gen), f_rest=torch.randn(P, M - 1, 3, generator=gen),
            opacity=torch.randn(P, 1, generator=gen), scaling=torch.randn(P, 3, generator=gen) - 3, rotation=torch.randn(P, 4, generator=gen))
lrs = dict(xyz=8e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)
bucket = torch.randn(P * 59, device=dev)
segs, off = {}, 0
for k, shp in (("means3D", (P, 3)), ("shs", (P, M, 3)), ("opacities", (P, 1)), ("scales", (P, 3)), ("rotations", (P, 4))):
    n = 1
    for d in shp: n *= d
    segs[k] = bucket[off:off + n].view(shp); off += n

mine = GaussianParamStep({k: v.to(dev) for k, v in host.items()}, lrs)
params = {k: torch.nn.Parameter(v.to(dev)) for k, v in host.items()}
opt = torch.optim.Adam([{"params": [params[k]], "lr": lrs[k]} for k in GROUPS], lr=0.0, eps=1e-15)


def activate():
    return dict(means3D=params["xyz"], shs=torch.cat((params["f_dc"], params["f_rest"]), dim=1), opacities=torch.sigmoid(params["opacity"]),
                scales=torch.exp(params["scaling"]), rotations=torch.nn.functional.normalize(params["rotation"]))


state = {"act": activate()}


def ref_step():
    act = state["act"]
    torch.autograd.backward([act[k] for k in segs], [segs[k] for k in segs])     # activation backward with the bucket as upstream gradient
    opt.step(); opt.zero_grad(set_to_none=True)
    state["act"] = activate()


def timed(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ms_m = timed(lambda: mine.step(bucket))
ms_r = timed(ref_step)
b = 8 * 59 * 4 * P
print(json.dumps({"metric": "parameter step time (1M Gaussians: activation backward + Adam + re-activation)", "unit": "ms", "higher_is_better": False,
                  "value": round(ms_m, 4), "torch_reference_ms": round(ms_r, 4), "alg_bytes": b, "gbs": round(b / ms_m / 1e6, 1)}))
