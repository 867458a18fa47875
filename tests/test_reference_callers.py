"""Drop-in integration test (last row): the reference's UNMODIFIED Python callers of the path —
gaussian_renderer/__init__.py::render (:9-79) and GaussianModel.create_from_pcd (scene/gaussian_model.py:107-130) — run on top
of this repository's drop-in modules (submodules/diff_gaussian_rasterization/diff_gaussian_rasterization, simple_knn._C), and
their results are compared with the reference CUDA (oracle/_ref/libref_rast.so) driven with the same activated tensors.

The reference's Python files are staged verbatim by `make -f oracle/Makefile ref` into the git-ignored oracle/_ref/refpy/
(they travel to the GPU box with the snapshot; /root/reference does not exist there).  /root/repo comes first on sys.path so
`submodules.*` and `simple_knn` resolve to this repository; `plyfile` (not installed in this image) is tests/stubs/plyfile.py.
"""
import math
import os
import sys
import types

import numpy as np
import pytest
import torch

from harness import rel_err
from rain_b200.synthetic import make_camera, make_grad, make_scene

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFPY = os.path.join(ROOT, "oracle", "_ref", "refpy")
STUBS = os.path.join(ROOT, "tests", "stubs")
REF_TOP = ("gaussian_renderer", "scene", "utils", "arguments", "plyfile")


@pytest.fixture(scope="module")
def refmods(device):
    """import the reference's gaussian_renderer / scene.gaussian_model / scene.cameras / utils.graphics_utils, unmodified"""
    if not os.path.isdir(os.path.join(REFPY, "gaussian_renderer")):
        pytest.skip("oracle/_ref/refpy not staged (make -f oracle/Makefile ref, needs /root/reference at build time)")
    saved_path = list(sys.path)
    cwd = os.getcwd()
    os.chdir(ROOT)                                   # the callers assume cwd = repo root (gaussian_renderer/__init__.py:4)
    sys.path[:0] = [ROOT, REFPY, STUBS]
    try:
        import gaussian_renderer
        from scene.cameras import MiniCam
        from scene.gaussian_model import GaussianModel
        from utils.graphics_utils import BasicPointCloud
        from utils.general_utils import inverse_sigmoid
        import submodules.diff_gaussian_rasterization.diff_gaussian_rasterization as dgr
        import simple_knn._C as knn
        # the imports of the unmodified callers resolved to THIS repository's modules
        assert os.path.abspath(dgr.__file__).startswith(os.path.join(ROOT, "submodules"))
        assert os.path.abspath(knn.__file__).startswith(os.path.join(ROOT, "simple_knn"))
        assert os.path.abspath(gaussian_renderer.__file__).startswith(REFPY)
        yield types.SimpleNamespace(render=gaussian_renderer.render, MiniCam=MiniCam, GaussianModel=GaussianModel, BasicPointCloud=BasicPointCloud,
                                    inverse_sigmoid=inverse_sigmoid, dgr=dgr)
    finally:
        os.chdir(cwd)
        sys.path[:] = saved_path
        for name in list(sys.modules):
            if name.split(".")[0] in REF_TOP:
                del sys.modules[name]


def _model_from_scene(refmods, sc, device):
    """a reference GaussianModel whose raw leaves reproduce the synthetic scene's activated values"""
    pc = refmods.GaussianModel(3, 0.8)
    P = sc.means3D.shape[0]
    g = torch.Generator().manual_seed(17)
    raw_rot = sc.rotations * (0.5 + torch.rand(P, 1, generator=g))                     # un-normalised: the caller normalises
    leaf = lambda t: torch.nn.Parameter(t.to(device).contiguous().requires_grad_(True))
    pc._xyz = leaf(sc.means3D)
    pc._features_dc = leaf(sc.shs[:, 0:1, :])
    pc._features_rest = leaf(sc.shs[:, 1:, :])
    pc._scaling = leaf(torch.log(sc.scales))
    pc._rotation = leaf(raw_rot)
    pc._opacity = leaf(refmods.inverse_sigmoid(sc.opacities.clamp(1e-4, 1 - 1e-4)))
    pc.active_sh_degree = sc.sh_degree
    return pc


@pytest.mark.parametrize("convert_shs,compute_cov", [(False, False), (True, False), (False, True)])
def test_unmodified_render_matches_reference_cuda(device, refcuda, refmods, convert_shs, compute_cov):
    P, W, H, D = 20000, 208, 120, 3
    sc = make_scene(P, D, seed=21)
    cam0 = make_camera(1, 4, W, H)
    fovy = math.radians(50.0); fovx = 2.0 * math.atan(math.tan(fovy / 2) * W / H)
    cam = refmods.MiniCam(W, H, fovy, fovx, 0.01, 100.0, cam0.viewmatrix.to(device), cam0.projmatrix.to(device))
    pipe = types.SimpleNamespace(debug=False, compute_cov3D_python=compute_cov, convert_SHs_python=convert_shs)
    bg = torch.tensor([0.1, 0.2, 0.3], device=device)
    dpix = make_grad(W, H, 3).to(device)
    pc = _model_from_scene(refmods, sc, device)
    leaves = [pc._xyz, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling, pc._rotation]

    out = refmods.render(cam, pc, pipe, bg, low_pass=0.3)                               # the reference's own code, unchanged
    (out["render"] * dpix).sum().backward()
    mine = [t.grad.clone() for t in leaves]
    mine_2d = out["viewspace_points"].grad.clone()

    # the same activated tensors through the reference CUDA, then the same caller-side activations backward
    for t in leaves:
        t.grad = None
    tanfovx, tanfovy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
    act = dict(opacity=pc.get_opacity, scales=pc.get_scaling, rotations=pc.get_rotation, shs=pc.get_features)
    det = {k: v.detach().contiguous() for k, v in act.items()}
    kw = {}
    if convert_shs:
        # precomputed colours as the caller's python path forms them (gaussian_renderer/__init__.py:52-57)
        eval_sh = sys.modules["utils.sh_utils"].eval_sh
        shs_view = act["shs"].transpose(1, 2).view(-1, 3, 16)
        dir_pp = pc.get_xyz - cam.camera_center.repeat(P, 1)
        colors = torch.clamp_min(eval_sh(D, shs_view, dir_pp / dir_pp.norm(dim=1, keepdim=True)) + 0.5, 0.0)
        kw["colors_precomp"] = colors.detach().contiguous()
    else:
        kw["shs"] = det["shs"]
    if compute_cov:
        cov = pc.get_covariance(1.0)
        kw["cov3D_precomp"] = cov.detach().contiguous()
    # .contiguous() as the reference's own binding applies it (rasterize_points.cu:85-104): MiniCam.camera_center is a strided
    # slice of torch.inverse's column-major result
    N_r, color_r, radii_r, depth_r = refcuda.forward(bg, pc.get_xyz.detach(), det["opacity"], None if compute_cov else det["scales"],
                                                     None if compute_cov else det["rotations"], cam.world_view_transform.contiguous(), cam.full_proj_transform.contiguous(),
                                                     cam.camera_center.contiguous(), tanfovx, tanfovy, H, W, sh_degree=D, **kw)
    g_r = refcuda.backward(dpix)
    again = refcuda.backward(dpix)
    assert N_r > 0
    assert torch.equal(out["radii"], radii_r) and torch.equal(out["visibility_filter"], radii_r > 0)
    assert torch.equal(out["render"].detach(), color_r) and torch.equal(out["depth"].detach(), depth_r)      # bit-identical images
    # chain the reference's gradients through the caller-side graph (activations, eval_sh, covariance) exactly as autograd does
    outs, gouts = [pc.get_xyz], [g_r["dL_dmean3D"]]
    outs.append(act["opacity"]); gouts.append(g_r["dL_dopacity"])
    if convert_shs:
        outs.append(colors); gouts.append(g_r["dL_dcolor"])
    else:
        outs.append(act["shs"]); gouts.append(g_r["dL_dsh"])
    if compute_cov:
        outs.append(cov); gouts.append(g_r["dL_dcov3D"])
    else:
        outs += [act["scales"], act["rotations"]]; gouts += [g_r["dL_dscale"], g_r["dL_drot"]]
    torch.autograd.backward(outs, gouts)
    spread2d = rel_err(again["dL_dmean2D"], g_r["dL_dmean2D"])[0]
    assert rel_err(mine_2d, g_r["dL_dmean2D"])[0] <= 1e-4 + 2 * spread2d
    names = ["xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"]
    for name, t, m in zip(names, leaves, mine):
        # 1e-4 of max plus the reference's own run-to-run spread (its backward adds per pixel with unordered atomics)
        key = dict(xyz="dL_dmean3D", features_dc="dL_dsh", features_rest="dL_dsh", opacity="dL_dopacity", scaling="dL_dscale", rotation="dL_drot")[name]
        if convert_shs and key == "dL_dsh":
            key = "dL_dcolor"
        if compute_cov and key in ("dL_dscale", "dL_drot"):
            key = "dL_dcov3D"
        spread = rel_err(again[key], g_r[key])[0]
        assert t.grad is not None and rel_err(m, t.grad)[0] <= 1e-4 + 4 * spread, (name, rel_err(m, t.grad)[0], spread)


def test_unmodified_create_from_pcd_and_mark_visible(device, refcuda, refmods):
    """GaussianModel.create_from_pcd (scene/gaussian_model.py:107-130) -> simple_knn._C.distCUDA2 of this repository; the initial
    scales it derives are bit-identical to those from the reference's distCUDA2; then one unmodified render() and markVisible."""
    P = 30000
    g = torch.Generator().manual_seed(5)
    pts = (torch.rand(P, 3, generator=g) * 2 - 1).float()
    cols = torch.rand(P, 3, generator=g).float()
    pcd = refmods.BasicPointCloud(points=pts.numpy(), colors=cols.numpy(), normals=np.zeros((P, 3), np.float32))
    pc = refmods.GaussianModel(3, 0.8)
    pc.create_from_pcd(pcd, 1.0)
    d2 = torch.clamp_min(refcuda.dist2(pts.to(device)), 0.0000001)
    assert torch.equal(pc._scaling.detach(), torch.log(torch.sqrt(d2))[..., None].repeat(1, 3))
    W, H = 160, 96
    cam0 = make_camera(0, 3, W, H)
    fovy = math.radians(50.0); fovx = 2.0 * math.atan(math.tan(fovy / 2) * W / H)
    cam = refmods.MiniCam(W, H, fovy, fovx, 0.01, 100.0, cam0.viewmatrix.to(device), cam0.projmatrix.to(device))
    pipe = types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)
    bg = torch.zeros(3, device=device)
    out = refmods.render(cam, pc, pipe, bg)
    N_r, color_r, radii_r, depth_r = refcuda.forward(bg, pc.get_xyz.detach(), pc.get_opacity.detach(), pc.get_scaling.detach(), pc.get_rotation.detach(),
                                                     cam.world_view_transform.contiguous(), cam.full_proj_transform.contiguous(), cam.camera_center.contiguous(),
                                                     math.tan(fovx * 0.5), math.tan(fovy * 0.5), H, W, shs=pc.get_features.detach().contiguous(), sh_degree=0)
    assert torch.equal(out["radii"], radii_r) and torch.equal(out["render"].detach(), color_r)
    rs = refmods.dgr.GaussianRasterizationSettings(H, W, math.tan(fovx * 0.5), math.tan(fovy * 0.5), bg, 1.0, cam.world_view_transform, cam.full_proj_transform,
                                                   0, cam.camera_center, False, False, 0.3)
    vis = refmods.dgr.GaussianRasterizer(rs).markVisible(pc.get_xyz)
    assert torch.equal(vis, refcuda.mark_visible(pc.get_xyz.detach(), cam.world_view_transform.contiguous(), cam.full_proj_transform.contiguous()))
