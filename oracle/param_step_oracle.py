"""CPU ORACLE (test infrastructure, never on the product path) for the fused parameter step: activations of the reference
(scene/gaussian_model.py:85-105: exp, sigmoid, F.normalize, cat), their backward, and torch.optim.Adam's update rule —
the algorithm lives in a third-party dependency, torch (2.11.0 in this image; the reference's environment.yml pins
pytorch 1.12.1), torch/optim/adam.py::_single_tensor_adam with amsgrad=False, weight_decay=0, maximize=False:
    exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    step_size = lr / (1 - beta1 ** step); denom = exp_avg_sq.sqrt() / sqrt(1 - beta2 ** step) + eps
    param.addcdiv_(exp_avg, denom, value=-step_size)
Pinned by tests/golden/param_step.npz, produced by tests/golden/make_golden_param_step.py with the real torch.optim.Adam
and autograd through the activations."""
import numpy as np

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def activate(leaves):
    q = leaves["rotation"]
    n = np.maximum(np.sqrt((q * q).sum(1, keepdims=True)), 1e-12)
    return dict(means3D=leaves["xyz"], shs=np.concatenate([leaves["f_dc"], leaves["f_rest"]], axis=1), opacities=1.0 / (1.0 + np.exp(-leaves["opacity"])),
                scales=np.exp(leaves["scaling"]), rotations=q / n)


def leaf_grads(leaves, g):
    """g: dict(means3D, shs, opacities, scales, rotations) gradients w.r.t. the activated values -> gradients of the raw leaves."""
    s = 1.0 / (1.0 + np.exp(-leaves["opacity"]))
    q = leaves["rotation"]
    n = np.maximum(np.sqrt((q * q).sum(1, keepdims=True)), 1e-12)
    qh = q / n
    return dict(xyz=g["means3D"], f_dc=g["shs"][:, :1], f_rest=g["shs"][:, 1:], opacity=g["opacities"] * s * (1 - s), scaling=g["scales"] * np.exp(leaves["scaling"]),
                rotation=(g["rotations"] - qh * (qh * g["rotations"]).sum(1, keepdims=True)) / n)


def adam_step(leaves, m, v, g_act, lrs, step, betas=(0.9, 0.999), eps=1e-15):
    """in place on dicts of float64 arrays; returns the activated values after the update"""
    gl = leaf_grads(leaves, g_act)
    b1, b2 = betas
    for k in GROUPS:
        g = gl[k]
        m[k] += (g - m[k]) * (1 - b1)
        v[k] *= b2; v[k] += (1 - b2) * g * g
        leaves[k] -= lrs[k] / (1 - b1 ** step) * m[k] / (np.sqrt(v[k]) / np.sqrt(1 - b2 ** step) + eps)
    return activate(leaves)


def densify_stats(radii, dL_dmean2D, max_radii2D, xyz_gradient_accum, denom):
    """train.py:132-134 and scene/gaussian_model.py:419-421, in place on float arrays of shape [P] (radii: int [P])."""
    vis = radii > 0
    max_radii2D[vis] = np.maximum(max_radii2D[vis], radii[vis].astype(max_radii2D.dtype))
    xyz_gradient_accum[vis] += np.sqrt((dL_dmean2D[vis, :2].astype(np.float64) ** 2).sum(-1)).astype(xyz_gradient_accum.dtype)
    denom[vis] += 1
