"""Fused parameter step: the numpy oracle against golden vectors produced with the real
torch.optim.Adam + autograd through the reference's activations (CPU); the sm_100a kernel against both (GPU)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.param_step_oracle import GROUPS, activate, adam_step  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "param_step.npz"))
ACT = ("means3D", "shs", "opacities", "scales", "rotations")
STEPS = 3


def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def flat_bucket(g):
    """gradient dict -> flat bucket layout means3D | sh | opacity | scales | rotations (rain_b200/multiview.py)"""
    return np.concatenate([g[k].reshape(-1) for k in ACT]).astype(np.float32)


def test_oracle_matches_torch_adam_golden_vectors():
    leaves = {k: G["init_" + k].astype(np.float64) for k in GROUPS}
    m = {k: np.zeros_like(x) for k, x in leaves.items()}; v = {k: np.zeros_like(x) for k, x in leaves.items()}
    lrs = {k: float(G["lr_" + k]) for k in GROUPS}
    prev = {k: G["init_" + k].astype(np.float64) for k in GROUPS}
    for s in range(STEPS):
        adam_step(leaves, m, v, {k: G[f"step{s}_g_{k}"].astype(np.float64) for k in ACT}, lrs, s + 1)
        for k in GROUPS:
            gold = G[f"step{s}_{k}"].astype(np.float64)
            assert rel(leaves[k], gold) <= 1e-6, (s, k)
            assert rel(leaves[k] - prev[k], gold - prev[k]) <= 2e-3, (s, k)      # the update itself (golden parameters are fp32)
            prev[k] = gold


def test_oracle_zero_gradient_is_a_fixed_point_and_activations_are_consistent():
    leaves = {k: G["init_" + k].astype(np.float64) for k in GROUPS}
    before = {k: x.copy() for k, x in leaves.items()}
    m = {k: np.zeros_like(x) for k, x in leaves.items()}; v = {k: np.zeros_like(x) for k, x in leaves.items()}
    act = adam_step(leaves, m, v, {k: np.zeros_like(G[f"step0_g_{k}"], dtype=np.float64) for k in ACT}, {k: 1.0 for k in GROUPS}, 1)
    assert all(np.array_equal(leaves[k], before[k]) for k in GROUPS)
    assert np.allclose(np.linalg.norm(act["rotations"], axis=1), 1.0) and act["shs"].shape[1] == 16 and (act["scales"] > 0).all()


@pytest.mark.gpu
def test_kernel_matches_torch_adam_golden_vectors_and_oracle(device):
    from rain_b200.optim import GaussianParamStep
    leaves = {k: torch.from_numpy(G["init_" + k].copy()).to(device) for k in GROUPS}
    lrs = {k: float(G["lr_" + k]) for k in GROUPS}
    opt = GaussianParamStep(leaves, lrs)
    o_leaves = {k: G["init_" + k].astype(np.float64) for k in GROUPS}
    o_m = {k: np.zeros_like(x) for k, x in o_leaves.items()}; o_v = {k: np.zeros_like(x) for k, x in o_leaves.items()}
    prev = {k: G["init_" + k].astype(np.float64) for k in GROUPS}
    for s in range(STEPS):
        g = {k: G[f"step{s}_g_{k}"] for k in ACT}
        act = opt.step(torch.from_numpy(flat_bucket(g)).to(device))
        o_act = adam_step(o_leaves, o_m, o_v, {k: x.astype(np.float64) for k, x in g.items()}, lrs, s + 1)
        for k in GROUPS:
            mine = leaves[k].cpu().numpy().astype(np.float64); gold = G[f"step{s}_{k}"].astype(np.float64)
            assert rel(mine, gold) <= 1e-6 and rel(mine, o_leaves[k]) <= 1e-6, (s, k)
            assert rel(mine - prev[k], gold - prev[k]) <= 2e-3, (s, k)
            prev[k] = gold
        for k in ACT:
            assert rel(act[k].cpu().numpy(), o_act[k]) <= 1e-6, (s, k)


@pytest.mark.gpu
def test_full_size_properties(device):
    """1M Gaussians: a zero gradient leaves every leaf bit-identical; two replicas stepping on the same bucket stay
    bit-identical (what view-sharded data parallelism relies on after the all-reduce); bad inputs fail loudly."""
    from rain_b200.optim import GaussianParamStep
    P, M = 1_000_003, 16
    gen = torch.Generator().manual_seed(1)
    mk = lambda: dict(xyz=torch.randn(P, 3, generator=gen), f_dc=torch.randn(P, 1, 3, generator=gen), f_rest=torch.randn(P, M - 1, 3, generator=gen),
                      opacity=torch.randn(P, 1, generator=gen), scaling=torch.randn(P, 3, generator=gen), rotation=torch.randn(P, 4, generator=gen))
    host = mk()
    a = {k: v.to(device) for k, v in host.items()}; b = {k: v.to(device) for k, v in host.items()}
    lrs = dict(xyz=1e-3, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)
    oa, ob = GaussianParamStep(a, lrs), GaussianParamStep(b, lrs)
    zero = torch.zeros(P * 59, device=device)
    oa.step(zero)
    assert all(torch.equal(a[k].cpu(), host[k]) for k in GROUPS)
    oa = GaussianParamStep(a, lrs)
    for s in range(2):
        gr = torch.randn(P * 59, device=device, generator=torch.Generator(device=device).manual_seed(s))
        ra, rb = oa.step(gr), ob.step(gr)
    assert all(torch.equal(a[k], b[k]) for k in GROUPS) and all(torch.equal(ra[k], rb[k]) for k in ACT)
    assert torch.allclose(ra["rotations"].norm(dim=1), torch.ones(P, device=device), atol=1e-5)
    with pytest.raises(RuntimeError):
        oa.step(torch.zeros(5, device=device))
    with pytest.raises(RuntimeError):
        GaussianParamStep({k: v.cpu() for k, v in a.items()}, lrs)


@pytest.mark.gpu
def test_render_loss_step_loop_reduces_the_loss(device):
    """The pieces chained as train.py:109-146 chains them — render (fwd), L1+SSIM loss, backward into the bucket, parameter
    step, re-activated inputs into the next render — on a small scene: the loss against a fixed target must go down."""
    from rain_b200 import multiview as mv
    from rain_b200.loss import loss_and_grad
    from rain_b200.optim import GaussianParamStep
    from rain_b200.synthetic import make_camera, make_scene
    P, W, H, D = 4000, 128, 96, 3
    target_sc = make_scene(P, D, seed=11).to(device)
    cams = [make_camera(v, 4, W, H).to(device) for v in range(4)]
    bg = torch.zeros(3, device=device)
    tparams = dict(means3D=target_sc.means3D, opacities=target_sc.opacities, shs=target_sc.shs, scales=target_sc.scales, rotations=target_sc.rotations)
    bucket = mv.GradientBucket(P, 16, device)
    targets = {}
    res = mv.render_views_fwd_bwd(tparams, cams, [0, 1, 2, 3], bucket, lambda v, c, d: targets.setdefault(v, c.clone()) * 0, bg=bg, sh_degree=D)
    # perturbed raw leaves (inverse activations of the target scene + noise on colour and opacity)
    gen = torch.Generator().manual_seed(3)
    noise = lambda t, s: t + s * torch.randn(t.shape, generator=gen).to(device)
    leaves = dict(xyz=target_sc.means3D.clone(), f_dc=noise(target_sc.shs[:, :1].contiguous(), 0.5), f_rest=target_sc.shs[:, 1:].contiguous(),
                  opacity=noise(torch.logit(target_sc.opacities.clamp(1e-4, 1 - 1e-4)), 1.0), scaling=torch.log(target_sc.scales), rotation=target_sc.rotations.clone())
    opt = GaussianParamStep(leaves, dict(xyz=0.0, f_dc=0.01, f_rest=0.0005, opacity=0.05, scaling=0.0, rotation=0.0))
    params = opt.step(torch.zeros(P * 59, device=device))           # zero gradient: parameters unchanged, activated inputs produced
    opt.steps = 0
    history = []

    def grad_fn(v, color, depth):
        out3, g = loss_and_grad(color, targets[v], 0.2)
        history[-1].append(out3)
        return g
    for it in range(25):
        history.append([])
        bucket.filled = False
        mv.render_views_fwd_bwd(params, cams, [0, 1, 2, 3], bucket, grad_fn, bg=bg, sh_degree=D, keep_outputs=False)
        params = opt.step(bucket.flat)
    loss = [float(torch.stack(h)[:, 2].mean().item()) for h in history]
    assert loss[-1] < 0.6 * loss[0], loss


def test_densify_stats_oracle_matches_torch_restatement_of_the_reference():
    from oracle.param_step_oracle import densify_stats
    g = torch.Generator().manual_seed(5)
    P = 501
    radii = torch.randint(-1, 40, (P,), generator=g).clamp(min=0).int()
    grad = torch.randn(P, 3, generator=g)
    mx, acc, den = torch.rand(P, generator=g) * 30, torch.rand(P, 1, generator=g), torch.randint(0, 5, (P, 1), generator=g).float()
    o = [t.numpy().copy().reshape(P) for t in (mx, acc, den)]
    densify_stats(radii.numpy(), grad.numpy(), *o)
    vis = radii > 0                                                      # gaussian_renderer/__init__.py: visibility_filter
    mx[vis] = torch.max(mx[vis], radii[vis].float())                     # train.py:133
    acc[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)          # gaussian_model.py:420
    den[vis] += 1                                                        # :421
    for a, b in zip(o, (mx, acc, den)):
        assert np.allclose(a, b.numpy().reshape(P), rtol=1e-6, atol=0)


@pytest.mark.gpu
def test_densify_stats_kernel_matches_oracle(device):
    from oracle.param_step_oracle import densify_stats
    from rain_b200.optim import add_densification_stats
    g = torch.Generator().manual_seed(6)
    for P in (1, 257, 100_003):
        radii = torch.randint(-1, 40, (P,), generator=g).clamp(min=0).int()
        grad = torch.randn(P, 3, generator=g)
        mx, acc, den = torch.rand(P, generator=g) * 30, torch.rand(P, 1, generator=g), torch.randint(0, 5, (P, 1), generator=g).float()
        o = [t.numpy().copy().reshape(P) for t in (mx, acc, den)]
        densify_stats(radii.numpy(), grad.numpy(), *o)
        d = [t.to(device) for t in (mx, acc, den)]
        add_densification_stats(radii.to(device), grad.to(device), *d)
        for a, b in zip(o, d):
            assert np.allclose(a, b.cpu().numpy().reshape(P), rtol=1e-6, atol=0)
    with pytest.raises(RuntimeError):
        add_densification_stats(radii, grad, mx, acc, den)


@pytest.mark.gpu
def test_leaf_step_matches_torch_adam(device):
    """rr_param_step_leaf (Adam over the six raw leaves from LEAF gradients in the leaf bucket layout) against torch.optim.Adam
    with the reference's parameter groups (scene/gaussian_model.py:139-157: per-group lr, eps 1e-15) on the same gradients."""
    from rain_b200.multiview import GradientBucket, LEAF_SEGMENTS
    from rain_b200.optim import GaussianParamStep
    gen = torch.Generator().manual_seed(7)
    P = 1234
    shapes = dict(xyz=(P, 3), f_dc=(P, 1, 3), f_rest=(P, 15, 3), opacity=(P, 1), scaling=(P, 3), rotation=(P, 4))
    init = {k: torch.randn(*shp, generator=gen) for k, shp in shapes.items()}
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)
    mine = {k: v.clone().to(device) for k, v in init.items()}
    ref = {k: torch.nn.Parameter(v.clone().to(device)) for k, v in init.items()}
    adam = torch.optim.Adam([{"params": [ref[k]], "lr": lrs[k]} for k in LEAF_SEGMENTS], lr=0.0, eps=1e-15)
    opt = GaussianParamStep(mine, lrs)
    bucket = GradientBucket(P, 16, device, layout="leaf")
    for step in range(3):
        for k in LEAF_SEGMENTS:
            g = torch.randn(*shapes[k], generator=gen).to(device) * (10.0 ** (step - 1))
            bucket.views[k].copy_(g)
            ref[k].grad = g.clone()
        adam.step()
        out = opt.step_leaf(bucket.flat)
        for k in LEAF_SEGMENTS:
            assert rel(out[k].cpu().numpy(), ref[k].detach().cpu().numpy()) <= 2e-6, (step, k)
            assert rel((out[k].cpu() - init[k]).numpy(), (ref[k].detach().cpu() - init[k]).numpy()) <= 2e-3, (step, k)
