"""Host side of the fused image loss: the names of the reference's utils/loss_utils.py over the
sm_100a kernels of csrc/loss.cu, through the C ABI (rr_loss_l1_ssim / rr_loss_l2, include/rain_rast.h).

    loss = training_loss(image, gt, lambda_dssim)          # == train.py:112-113, differentiable w.r.t. image
    Ll1, s = l1_loss(image, gt), ssim(image, gt)           # loss_utils.py:6-7, :23-51 (forward values, no graph)

There is no CPU path: CUDA tensors only."""
import torch

from . import _lib


def _check(image, gt):
    if not (image.is_cuda and gt.is_cuda):
        raise RuntimeError("rain_b200 has no CPU path: image and gt must be CUDA tensors")
    if image.shape != gt.shape or image.dim() != 3:
        raise RuntimeError(f"image and gt must both be [C,H,W]; got {tuple(image.shape)} and {tuple(gt.shape)}")
    return image.contiguous().float(), gt.contiguous().float()


def _run(image, gt, lambda_dssim, want_grad):
    lib = _lib.load()
    C, H, W = image.shape
    dev = image.device
    ws = torch.empty(lib.rr_loss_ws_bytes(C, H, W), dtype=torch.uint8, device=dev)
    out3 = torch.empty(3, dtype=torch.float32, device=dev)
    grad = torch.empty_like(image) if want_grad else None
    with torch.cuda.device(dev):
        _lib.check(lib.rr_loss_l1_ssim(torch.cuda.current_stream().cuda_stream, C, H, W, image.data_ptr(), gt.data_ptr(), float(lambda_dssim),
                                       ws.data_ptr(), ws.numel(), out3.data_ptr(), grad.data_ptr() if want_grad else None), "rr_loss_l1_ssim")
    return out3, grad


def loss_and_grad(image, gt, lambda_dssim=0.2):
    """({Ll1, ssim, loss} as a 3-float device tensor, d loss / d image) without autograd bookkeeping — what a grad_fn of
    rain_b200.multiview.render_views_fwd_bwd returns as dL/dcolor."""
    image, gt = _check(image, gt)
    return _run(image, gt, lambda_dssim, True)


class _TrainingLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        image, gt = _check(image, gt)
        out3, grad = _run(image, gt, lambda_dssim, ctx.needs_input_grad[0])
        ctx.save_for_backward(grad)
        return out3[2].clone()

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return (grad * g if grad is not None else None), None, None


def training_loss(image, gt, lambda_dssim=0.2):
    """(1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt)) (train.py:112-113); one forward + one gradient kernel."""
    return _TrainingLoss.apply(image, gt, lambda_dssim)


def l1_loss(network_output, gt):
    """utils/loss_utils.py:6-7 (value only)."""
    a, b = _check(network_output, gt)
    return _run(a, b, 0.0, False)[0][0]


def ssim(img1, img2, window_size=11, size_average=True):
    """utils/loss_utils.py:23-51 with the reference's only configuration (window 11, size_average=True); value only."""
    if window_size != 11 or not size_average:
        raise RuntimeError("only window_size=11, size_average=True (the configuration train.py uses) is implemented")
    a, b = _check(img1, img2)
    return _run(a, b, 0.0, False)[0][1]


def l2_loss_grad(image, gt, loss_accum=None, out=None, u8_scale=1.0 / 255.0):
    """dL/dimage = image - gt for L = 0.5 * sum (image - gt)^2, and loss_accum[0] += L, fused in one pass.  gt is float32,
    or uint8 (the image in its native 8-bit form, gt = byte * u8_scale, converted in the kernel)."""
    if not (image.is_cuda and gt.is_cuda):
        raise RuntimeError("rain_b200 has no CPU path: image and gt must be CUDA tensors")
    lib = _lib.load()
    image = image.contiguous(); gt = gt.contiguous()
    out = torch.empty_like(image) if out is None else out
    with torch.cuda.device(image.device):
        acc = loss_accum.data_ptr() if loss_accum is not None else None
        if gt.dtype == torch.uint8:
            _lib.check(lib.rr_loss_l2_u8(torch.cuda.current_stream().cuda_stream, image.numel(), image.data_ptr(), gt.data_ptr(), float(u8_scale),
                                         out.data_ptr(), acc), "rr_loss_l2_u8")
        else:
            _lib.check(lib.rr_loss_l2(torch.cuda.current_stream().cuda_stream, image.numel(), image.data_ptr(), gt.data_ptr(), out.data_ptr(), acc), "rr_loss_l2")
    return out
