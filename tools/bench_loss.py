"""Image loss fwd+bwd at 3x1080x1920: rain_b200.loss.training_loss (csrc/loss.cu) against a torch restatement of the
reference's utils/loss_utils.py + autograd on the same GPU.  One JSON line.   python tools/bench_loss.py [reps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.nn.functional as F
from rain_b200 import loss as L
from oracle.loss_oracle import window_1d

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda:0")
C, H, W, lam = 3, 1080, 1920, 0.2
g = torch.Generator().manual_seed(0)
gt = torch.rand(C, H, W, generator=g).to(dev)
img = (gt + 0.1 * torch.randn(C, H, W, generator=g).to(dev)).clamp(0, 1)
w1 = torch.from_numpy(window_1d()).to(dev)
win = (w1[:, None] @ w1[None, :]).expand(C, 1, 11, 11).contiguous()


def ref_step():
    x = img.clone().requires_grad_(True)
    cv = lambda t: F.conv2d(t[None], win, padding=5, groups=C)[0]
    mu1, mu2 = cv(x), cv(gt)
    s1, s2, s12 = cv(x * x) - mu1 * mu1, cv(gt * gt) - mu2 * mu2, cv(x * gt) - mu1 * mu2
    S = ((2 * mu1 * mu2 + 0.01 ** 2) * (2 * s12 + 0.03 ** 2)) / ((mu1 * mu1 + mu2 * mu2 + 0.01 ** 2) * (s1 + s2 + 0.03 ** 2))
    loss = (1 - lam) * (x - gt).abs().mean() + lam * (1 - S.mean())
    loss.backward()
    return loss


def mine_step():
    x = img.clone().requires_grad_(True)
    loss = L.training_loss(x, gt, lam)
    loss.backward()
    return loss


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, float(out.item())


ms_m, lm = timed(mine_step)
ms_r, lr = timed(ref_step)
n = C * H * W
print(json.dumps({"metric": "L1+SSIM loss fwd+bwd time (3x1080x1920)", "unit": "ms", "higher_is_better": False, "value": round(ms_m, 4),
                  "torch_restatement_of_reference_ms": round(ms_r, 4), "loss": lm, "loss_reference": lr, "alg_bytes": 11 * 4 * n,
                  "gbs": round(11 * 4 * n / ms_m / 1e6, 1), "note": "includes the clone of the image and autograd bookkeeping on both sides"}))
