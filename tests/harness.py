"""Shared helpers for the parity tests: run the library through its C-ABI-backed ops, pull the stage intermediates out
of the opaque blobs, and compare against the oracle / the rebuilt reference CUDA."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from rain_b200 import _C, _lib  # noqa: E402
from rain_b200.synthetic import make_camera, make_grad, make_scene  # noqa: E402


def _view(blob, off, count, dtype):
    nbytes = count * torch.empty((), dtype=dtype).element_size()
    return blob[off:off + nbytes].view(dtype)


def mine_forward(sc, cam, bg, *, colors_precomp=None, cov3D_precomp=None, use_sh=True, low_pass=0.3, scale_modifier=1.0, debug=False):
    """Run rain_b200._C.rasterize_gaussians and return outputs + intermediates (torch tensors on the device)."""
    dev = sc.means3D.device
    e = torch.empty(0, device=dev)
    sh = sc.shs if (use_sh and colors_precomp is None) else e
    args = (bg, sc.means3D, colors_precomp if colors_precomp is not None else e, sc.opacities,
            sc.scales if cov3D_precomp is None else e, sc.rotations if cov3D_precomp is None else e, scale_modifier,
            cov3D_precomp if cov3D_precomp is not None else e, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy,
            cam.height, cam.width, sh, sc.sh_degree, cam.campos, False, debug, low_pass)
    N, color, radii, depth, geom, binning, img = _C.rasterize_gaussians(*args)
    lib = _lib.load()
    P, W, H = sc.means3D.shape[0], cam.width, cam.height
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out = dict(N=N, color=color, radii=radii, depth=depth, geom=geom, binning=binning, img=img, args=args)
    if P == 0:
        return out
    go = (ctypes.c_int64 * 9)(); lib.rr_geom_layout(P, go)
    io = (ctypes.c_int64 * 4)(); lib.rr_image_layout(W, H, io)
    rec = _view(geom, go[0], P * 12, torch.float32).view(P, 12)
    out.update(rec=rec, means2D=rec[:, 0:2], conic_opacity=torch.cat([rec[:, 2:5], rec[:, 5:6]], 1), rgb=rec[:, 6:9], depths=rec[:, 9],
               cov3D=_view(geom, go[1], P * 6, torch.float32).view(P, 6), clamped=_view(geom, go[2], P, torch.uint8),
               tiles_touched=_view(geom, go[3], P, torch.int32), depth_order=_view(geom, go[5], P, torch.int32),
               final_T=_view(img, io[0], H * W, torch.float32), n_contrib=_view(img, io[1], H * W, torch.int32),
               ranges=_view(img, io[2], T * 2, torch.int32).view(T, 2), tile_max=_view(img, io[3], T, torch.int32))
    if N > 0:
        bo = (ctypes.c_int64 * 5)(); lib.rr_binning_layout(N, W, H, bo)
        out.update(tile_keys=_view(binning, bo[0], N, torch.int32), point_list=_view(binning, bo[3], N, torch.int32))
        keys = torch.empty(N, dtype=torch.int64, device=dev)
        _lib.check(lib.rr_export_sorted_keys(torch.cuda.current_stream().cuda_stream, P, N, W, H, geom.data_ptr(), binning.data_ptr(), img.data_ptr(), keys.data_ptr()), "export")
        out["keys"] = keys
    return out


def mine_backward(fw, sc, cam, bg, dL_dpix, *, low_pass=0.3, scale_modifier=1.0, debug=False):
    a = fw["args"]
    grads = _C.rasterize_gaussians_backward(a[0], a[1], fw["radii"], a[2], a[4], a[5], a[6], a[7], a[8], a[9], a[10], a[11], dL_dpix,
                                            a[14], a[15], a[16], fw["geom"], fw["N"], fw["binning"], fw["img"], debug, a[19], return_conic=True)
    names = ["dL_dmean2D", "dL_dcolor", "dL_dopacity", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot", "dL_dconic"]
    return dict(zip(names, grads))


def rel_err(a, b):
    """max |a-b| / max|b| and relative L2 over the tensor"""
    a = torch.as_tensor(a).double().flatten().cpu(); b = torch.as_tensor(b).double().flatten().cpu()
    if a.numel() == 0:
        return 0.0, 0.0
    scale = b.abs().max().item()
    n2 = b.norm().item()
    return ((a - b).abs().max().item() / scale if scale > 0 else (a - b).abs().max().item(),
            (a - b).norm().item() / n2 if n2 > 0 else (a - b).norm().item())


def elem_rel_err_frac(a, b, rtol, atol_frac=1e-6):
    """fraction of elements with |a-b| > rtol*|b| + atol, atol = atol_frac * max|b|"""
    a = torch.as_tensor(a).double().flatten().cpu(); b = torch.as_tensor(b).double().flatten().cpu()
    if a.numel() == 0:
        return 0.0
    atol = atol_frac * b.abs().max().item()
    return ((a - b).abs() > rtol * b.abs() + atol).double().mean().item()


def scene_on(device, P, D, seed=0, scale_mult=1.0):
    return make_scene(P, D, seed, scale_mult).to(device)
