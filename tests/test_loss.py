"""Image loss: the numpy oracle against golden vectors produced by the reference's own
utils/loss_utils.py (CPU); the sm_100a kernels against the oracle, the golden vectors and — at full resolution — a torch
restatement of loss_utils.py:23-51 on the same GPU."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.loss_oracle import loss_and_grad, window_1d  # noqa: E402

GOLDEN = np.load(os.path.join(ROOT, "tests", "golden", "loss_l1_ssim.npz"))
CASES = "abcd"


def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden_vectors(name):
    l1, s, loss, grad = loss_and_grad(GOLDEN[name + "_img"], GOLDEN[name + "_gt"], float(GOLDEN[name + "_lambda"]))
    o = GOLDEN[name + "_out3"]
    assert abs(l1 - o[0]) <= 1e-5 * abs(o[0]) and abs(s - o[1]) <= 1e-5 * abs(o[1]) and abs(loss - o[2]) <= 1e-5 * abs(o[2])
    assert rel(grad, GOLDEN[name + "_grad"]) <= 1e-5


def test_oracle_window_and_identities():
    g = window_1d()
    assert g.dtype == np.float32 and g.shape == (11,) and abs(float(g.sum()) - 1.0) < 1e-6 and np.allclose(g, g[::-1])
    x = np.random.default_rng(0).random((3, 20, 24)).astype(np.float32)
    l1, s, loss, grad = loss_and_grad(x, x, 0.2)
    assert l1 == 0.0 and abs(s - 1.0) < 1e-12 and abs(loss) < 1e-12
    # finite-difference check of the analytic gradient (float64)
    y = np.random.default_rng(1).random((3, 20, 24))
    x = (y + 0.1 * np.random.default_rng(2).standard_normal(y.shape)).clip(0.01, 0.99)
    _, _, f0, grad = loss_and_grad(x, y, 0.3)
    for (c, i, j) in [(0, 0, 0), (1, 7, 11), (2, 19, 23), (0, 10, 3)]:
        xp = x.copy(); xp[c, i, j] += 1e-6
        xm = x.copy(); xm[c, i, j] -= 1e-6
        fd = (loss_and_grad(xp, y, 0.3)[2] - loss_and_grad(xm, y, 0.3)[2]) / 2e-6
        assert abs(fd - grad[c, i, j]) <= 1e-6 * max(1.0, abs(grad[c, i, j]) * 1e3)


# ---------------------------------------------------------- GPU ----------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_kernels_match_golden_and_oracle(device, name):
    from rain_b200 import loss as L
    img = torch.from_numpy(GOLDEN[name + "_img"]).to(device); gt = torch.from_numpy(GOLDEN[name + "_gt"]).to(device)
    lam = float(GOLDEN[name + "_lambda"])
    x = img.clone().requires_grad_(True)
    loss = L.training_loss(x, gt, lam)
    loss.backward()
    o = GOLDEN[name + "_out3"]
    assert abs(loss.item() - o[2]) <= 1e-5 * abs(o[2])
    assert abs(L.l1_loss(img, gt).item() - o[0]) <= 1e-5 * abs(o[0]) and abs(L.ssim(img, gt).item() - o[1]) <= 1e-5 * abs(o[1])
    assert rel(x.grad.cpu().numpy(), GOLDEN[name + "_grad"]) <= 1e-4          # fp32 kernels vs the reference's fp32 autograd
    _, _, lo, go = loss_and_grad(GOLDEN[name + "_img"], GOLDEN[name + "_gt"], lam)
    assert abs(loss.item() - lo) <= 1e-5 * abs(lo) and rel(x.grad.cpu().numpy(), go) <= 1e-4


@pytest.mark.gpu
def test_full_resolution_against_torch_restatement_and_identities(device):
    import torch.nn.functional as F
    from rain_b200 import loss as L
    C, H, W, lam = 3, 1080, 1920, 0.2
    g = torch.Generator().manual_seed(7)
    gt = torch.rand(C, H, W, generator=g).to(device)
    img = (gt + 0.1 * torch.randn(C, H, W, generator=g).to(device)).clamp(0, 1)
    # torch restatement of loss_utils.py:12-51 (grouped conv2d with the 2-D window) + autograd, on the same device
    w1 = torch.from_numpy(window_1d()).to(device)
    win = (w1[:, None] @ w1[None, :]).expand(C, 1, 11, 11).contiguous()
    x = img.clone().requires_grad_(True)
    cv = lambda t: F.conv2d(t[None], win, padding=5, groups=C)[0]
    mu1, mu2 = cv(x), cv(gt)
    s1, s2, s12 = cv(x * x) - mu1 * mu1, cv(gt * gt) - mu2 * mu2, cv(x * gt) - mu1 * mu2
    S = ((2 * mu1 * mu2 + 0.01 ** 2) * (2 * s12 + 0.03 ** 2)) / ((mu1 * mu1 + mu2 * mu2 + 0.01 ** 2) * (s1 + s2 + 0.03 ** 2))
    ref = (1 - lam) * (x - gt).abs().mean() + lam * (1 - S.mean())
    ref.backward()
    y = img.clone().requires_grad_(True)
    mine = L.training_loss(y, gt, lam)
    mine.backward()
    assert abs(mine.item() - ref.item()) <= 1e-5 * abs(ref.item())
    assert rel(y.grad.cpu().numpy(), x.grad.cpu().numpy()) <= 1e-4
    # identities at full size: identical images -> L1 = 0, SSIM = 1, loss = 0
    assert L.l1_loss(gt, gt).item() == 0.0 and abs(L.ssim(gt, gt).item() - 1.0) <= 1e-6 and abs(L.training_loss(gt, gt, lam).item()) <= 1e-6


@pytest.mark.gpu
def test_l2_loss_grad_fused(device):
    from rain_b200 import loss as L
    for n in (1, 5, 4096 + 3, 3 * 1080 * 1920):
        a = torch.randn(n, device=device); b = torch.randn(n, device=device)
        acc = torch.zeros(1, device=device)
        d = L.l2_loss_grad(a, b, acc)
        assert torch.equal(d, a - b)
        ref = 0.5 * (a.double() - b.double()).pow(2).sum().item()
        assert abs(acc.item() - ref) <= 1e-5 * ref
        u = torch.randint(0, 256, (n,), dtype=torch.uint8, device=device)
        acc.zero_()
        d = L.l2_loss_grad(a, u, acc)
        assert torch.equal(d, a - u.float() * (1.0 / 255.0))
        ref = 0.5 * (a.double() - (u.float() * (1.0 / 255.0)).double()).pow(2).sum().item()
        assert abs(acc.item() - ref) <= 1e-5 * ref
    with pytest.raises(RuntimeError):
        L.l2_loss_grad(torch.zeros(4), torch.zeros(4))
