"""CPU ORACLE (test infrastructure, never on the product path) for the image loss of the reference:
utils/loss_utils.py:6-7 (l1_loss), :12-21 (gaussian / create_window), :23-51 (ssim / _ssim) and train.py:112-113,
restated in numpy with the analytic gradient w.r.t. the rendered image.  Pinned against the reference itself:
tests/golden/loss_l1_ssim.npz was produced by tests/golden/make_golden_loss.py, which imports the reference's own
utils/loss_utils.py (torch CPU, autograd) in the build container."""
import numpy as np

C1, C2 = 0.01 ** 2, 0.03 ** 2          # loss_utils.py:45-46


def window_1d(window_size=11, sigma=1.5):
    """loss_utils.py:12-14: float32 taps, normalised in float32."""
    g = np.array([np.exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)], dtype=np.float32)
    return g / g.sum(dtype=np.float32)


def _conv(img, w2d):
    """F.conv2d(img, window, padding=window_size // 2, groups=channel) (loss_utils.py:34): cross-correlation, zero padding."""
    k = w2d.shape[0]; h = k // 2
    C, H, W = img.shape
    pad = np.zeros((C, H + 2 * h, W + 2 * h), img.dtype)
    pad[:, h:h + H, h:h + W] = img
    out = np.zeros_like(img)
    for dy in range(k):
        for dx in range(k):
            out += w2d[dy, dx] * pad[:, dy:dy + H, dx:dx + W]
    return out


def loss_and_grad(img, gt, lambda_dssim, dtype=np.float64):
    """returns (Ll1, ssim, loss, dloss/dimg) for [C,H,W] arrays; arithmetic in `dtype`."""
    x = np.asarray(img, dtype=dtype); y = np.asarray(gt, dtype=dtype)
    g1 = window_1d().astype(dtype)
    w = np.outer(g1, g1).astype(dtype)                                   # loss_utils.py:18
    mu1, mu2 = _conv(x, w), _conv(y, w)                                  # :34-35
    mu1_sq, mu2_sq, mu12 = mu1 * mu1, mu2 * mu2, mu1 * mu2               # :37-39
    s1 = _conv(x * x, w) - mu1_sq; s2 = _conv(y * y, w) - mu2_sq; s12 = _conv(x * y, w) - mu12      # :41-43
    A1, A2, B1, B2 = 2 * mu12 + C1, 2 * s12 + C2, mu1_sq + mu2_sq + C1, s1 + s2 + C2
    S = A1 * A2 / (B1 * B2)                                              # :48
    n = x.size
    Ll1 = np.abs(x - y).mean()                                           # :6-7
    ssim = S.mean()                                                      # :50-51
    loss = (1.0 - lambda_dssim) * Ll1 + lambda_dssim * (1.0 - ssim)      # train.py:113
    # gradient: S depends on x through mu1, E[xx], E[xy], each a windowed sum (symmetric window)
    inv = 1.0 / (B1 * B2)
    dS_dmu1 = 2 * mu2 * (A2 - A1) * inv - S * 2 * mu1 * (B2 - B1) * inv
    dS_dxx = -S / B2
    dS_dxy = 2 * A1 * inv
    gS = -lambda_dssim / n
    grad = gS * (_conv(dS_dmu1, w) + 2 * x * _conv(dS_dxx, w) + y * _conv(dS_dxy, w)) + (1.0 - lambda_dssim) / n * np.sign(x - y)
    return float(Ll1), float(ssim), float(loss), grad
