"""Generates tests/golden/param_step.npz with the REAL torch.optim.Adam (the configuration of the reference's
training_setup, scene/gaussian_model.py:139-157: six groups, lr per group, eps=1e-15) and autograd through the reference's
activations (scene/gaussian_model.py:85-105).  python tests/golden/make_golden_param_step.py"""
import os
import numpy as np, torch

P, M, STEPS = 37, 16, 3
g = torch.Generator().manual_seed(0)
leaves = dict(xyz=torch.randn(P, 3, generator=g), f_dc=torch.randn(P, 1, 3, generator=g), f_rest=0.2 * torch.randn(P, M - 1, 3, generator=g),
              opacity=2 * torch.randn(P, 1, generator=g), scaling=torch.randn(P, 3, generator=g) - 3, rotation=torch.randn(P, 4, generator=g))
lrs = dict(xyz=0.00016 * 5.0, f_dc=0.0025, f_rest=0.0025 / 20.0, opacity=0.05, scaling=0.005, rotation=0.001)     # arguments/__init__.py:64-71
params = {k: torch.nn.Parameter(v.clone()) for k, v in leaves.items()}
opt = torch.optim.Adam([{"params": [params[k]], "lr": lrs[k], "name": k} for k in params], lr=0.0, eps=1e-15)
out = {f"init_{k}": v.numpy() for k, v in leaves.items()}
out.update({f"lr_{k}": np.float64(v) for k, v in lrs.items()})
for s in range(STEPS):
    act = dict(means3D=params["xyz"], shs=torch.cat((params["f_dc"], params["f_rest"]), dim=1), opacities=torch.sigmoid(params["opacity"]),
               scales=torch.exp(params["scaling"]), rotations=torch.nn.functional.normalize(params["rotation"]))
    G = {k: torch.randn(v.shape, generator=g) * (0.0 if (s == 1 and k == "opacities") else 1.0) for k, v in act.items()}   # one all-zero gradient segment
    opt.zero_grad()
    sum((act[k] * G[k]).sum() for k in act).backward()
    opt.step()
    for k in G:
        out[f"step{s}_g_{k}"] = G[k].numpy()
    for k in params:
        out[f"step{s}_{k}"] = params[k].detach().numpy().copy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "param_step.npz"), **out)
print("ok", len(out))
