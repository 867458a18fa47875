"""Generates tests/golden/loss_l1_ssim.npz by running the REFERENCE's own utils/loss_utils.py (torch CPU + autograd) on
seeded inputs.  Run in the build container (needs /root/reference):  python tests/golden/make_golden_loss.py"""
import importlib.util, os
import numpy as np, torch

spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/utils/loss_utils.py")
ref = importlib.util.module_from_spec(spec); spec.loader.exec_module(ref)
out = {}
for name, (C, H, W, lam, seed) in {"a": (3, 40, 52, 0.2, 0), "b": (3, 17, 9, 0.2, 1), "c": (1, 64, 64, 0.5, 2), "d": (3, 33, 70, 1.0, 3)}.items():
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g)
    img = (gt + 0.15 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
    if name == "b":
        img[:, :4] = gt[:, :4]                       # exact ties: |x - y| has zero gradient there
    x = img.clone().requires_grad_(True)
    Ll1 = ref.l1_loss(x, gt); s = ref.ssim(x, gt)
    loss = (1.0 - lam) * Ll1 + lam * (1.0 - s)       # train.py:113
    loss.backward()
    out.update({f"{name}_img": img.numpy(), f"{name}_gt": gt.numpy(), f"{name}_lambda": np.float32(lam),
                f"{name}_out3": np.array([Ll1.item(), s.item(), loss.item()], np.float32), f"{name}_grad": x.grad.numpy()})
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "loss_l1_ssim.npz"), **out)
print({k: v.shape for k, v in out.items() if k.endswith("grad")})
