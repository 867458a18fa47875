"""Debug-snapshot tuples as a fixture format: the reference's rasterizer, run with
``raster_settings.debug``, dumps the exact argument tuple of the failing binding call with ``torch.save`` —
``snapshot_fw.dump`` (20 forward arguments, reference diff_gaussian_rasterization/__init__.py:49-80) and
``snapshot_bw.dump`` (22 backward arguments, :91-120).  ``rain_b200.rasterizer`` writes the same files.  This module
saves such tuples on purpose and replays them through this library's binding, so that a scene captured from either
implementation can be run through the other.

    save_forward_snapshot(path, args) / save_backward_snapshot(path, args)
    replay_forward(path, device)  -> (num_rendered, color, radii, depth, geomBuffer, binningBuffer, imgBuffer)
    replay_backward(path, device) -> the 8 gradients of rasterize_gaussians_backward

The tuples hold CPU tensors and Python scalars only; the backward tuple carries the forward's scratch blobs, whose
layout is private to the implementation that wrote them: a backward snapshot replays only on the implementation it came
from, and ``replay_backward(..., rerun_forward=True)`` rebuilds the blobs from the inputs instead (the forward inputs it
needs — opacities — are not part of the backward tuple, so they are passed in).
"""
import torch

from . import _C

FORWARD_FIELDS = ("bg", "means3D", "colors_precomp", "opacities", "scales", "rotations", "scale_modifier", "cov3Ds_precomp", "viewmatrix",
                  "projmatrix", "tanfovx", "tanfovy", "image_height", "image_width", "sh", "sh_degree", "campos", "prefiltered", "debug", "low_pass")
BACKWARD_FIELDS = ("bg", "means3D", "radii", "colors_precomp", "scales", "rotations", "scale_modifier", "cov3Ds_precomp", "viewmatrix",
                   "projmatrix", "tanfovx", "tanfovy", "grad_out_color", "sh", "sh_degree", "campos", "geomBuffer", "num_rendered",
                   "binningBuffer", "imgBuffer", "debug", "low_pass")


def _cpu(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _check(args, fields, what):
    if not isinstance(args, (tuple, list)) or len(args) not in (len(fields), len(fields) - 1):
        raise RuntimeError(f"not a {what} snapshot: expected a tuple of {len(fields)} items (or {len(fields) - 1}: upstream builds without low_pass)")
    args = tuple(args)
    if len(args) == len(fields) - 1:
        args = args + (0.3,)            # upstream diff-gaussian-rasterization has no low_pass argument: its constant 0.3
    return args


def save_forward_snapshot(path, args):
    torch.save(_cpu(_check(args, FORWARD_FIELDS, "forward")), path)


def save_backward_snapshot(path, args):
    torch.save(_cpu(_check(args, BACKWARD_FIELDS, "backward")), path)


def load_snapshot(path):
    """The tuple as saved, plus its kind ("forward" / "backward") by length."""
    args = torch.load(path, map_location="cpu", weights_only=False)
    n = len(args)
    if n in (len(FORWARD_FIELDS), len(FORWARD_FIELDS) - 1):
        return "forward", _check(args, FORWARD_FIELDS, "forward")
    if n in (len(BACKWARD_FIELDS), len(BACKWARD_FIELDS) - 1):
        return "backward", _check(args, BACKWARD_FIELDS, "backward")
    raise RuntimeError(f"{path}: a tuple of {n} items is neither a forward ({len(FORWARD_FIELDS)}) nor a backward ({len(BACKWARD_FIELDS)}) snapshot")


def _to(args, device):
    return tuple(a.to(device) if isinstance(a, torch.Tensor) else a for a in args)


def replay_forward(path, device="cuda:0"):
    kind, args = load_snapshot(path)
    if kind != "forward":
        raise RuntimeError(f"{path} is a {kind} snapshot")
    return _C.rasterize_gaussians(*_to(args, device))


def replay_backward(path, device="cuda:0", rerun_forward=False, opacities=None):
    kind, args = load_snapshot(path)
    if kind != "backward":
        raise RuntimeError(f"{path} is a {kind} snapshot")
    a = dict(zip(BACKWARD_FIELDS, _to(args, device)))
    if rerun_forward:
        if opacities is None:
            raise RuntimeError("rerun_forward needs the opacities (they are not part of a backward snapshot)")
        H, W = a["grad_out_color"].shape[1:]
        N, _, radii, _, geom, binning, img = _C.rasterize_gaussians(
            a["bg"], a["means3D"], a["colors_precomp"], opacities.to(device), a["scales"], a["rotations"], a["scale_modifier"], a["cov3Ds_precomp"],
            a["viewmatrix"], a["projmatrix"], a["tanfovx"], a["tanfovy"], H, W, a["sh"], a["sh_degree"], a["campos"], False, a["debug"], a["low_pass"])
        a.update(num_rendered=N, radii=radii, geomBuffer=geom, binningBuffer=binning, imgBuffer=img)
    return _C.rasterize_gaussians_backward(*(a[f] for f in BACKWARD_FIELDS))
