"""Diagnostic for tests/test_reference_callers.py: where does the unmodified render() output differ from the reference CUDA?"""
import math, os, sys, types
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle", "_ref", "refpy"), os.path.join(ROOT, "tests", "stubs"), os.path.join(ROOT, "tests")]
os.chdir(ROOT)
import gaussian_renderer
from scene.cameras import MiniCam
from scene.gaussian_model import GaussianModel
from utils.general_utils import inverse_sigmoid
from harness import mine_forward
from oracle.oracle_py import RefCuda
from rain_b200.synthetic import make_camera, make_grad, make_scene, Scene, Camera

dev = torch.device("cuda:0")
ref = RefCuda()
P, W, H, D = 20000, 208, 120, 3
sc = make_scene(P, D, seed=21)
cam0 = make_camera(1, 4, W, H)
fovy = math.radians(50.0); fovx = 2.0 * math.atan(math.tan(fovy / 2) * W / H)
cam = MiniCam(W, H, fovy, fovx, 0.01, 100.0, cam0.viewmatrix.to(dev), cam0.projmatrix.to(dev))
print("campos mini", cam.camera_center.cpu().numpy(), "make_camera", cam0.campos.numpy(), "tanfov", math.tan(fovx * 0.5), cam0.tanfovx)
pc = GaussianModel(3, 0.8)
g = torch.Generator().manual_seed(17)
raw_rot = sc.rotations * (0.5 + torch.rand(P, 1, generator=g))
leaf = lambda t: torch.nn.Parameter(t.to(dev).contiguous().requires_grad_(True))
pc._xyz = leaf(sc.means3D); pc._features_dc = leaf(sc.shs[:, 0:1, :]); pc._features_rest = leaf(sc.shs[:, 1:, :])
pc._scaling = leaf(torch.log(sc.scales)); pc._rotation = leaf(raw_rot); pc._opacity = leaf(inverse_sigmoid(sc.opacities.clamp(1e-4, 1 - 1e-4)))
pc.active_sh_degree = 3
pipe = types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)
bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
out = gaussian_renderer.render(cam, pc, pipe, bg, low_pass=0.3)
tanfovx, tanfovy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
det = dict(opacity=pc.get_opacity.detach().contiguous(), scales=pc.get_scaling.detach().contiguous(), rotations=pc.get_rotation.detach().contiguous(), shs=pc.get_features.detach().contiguous())
N_r, color_r, radii_r, depth_r = ref.forward(bg, pc.get_xyz.detach(), det["opacity"], det["scales"], det["rotations"], cam.world_view_transform, cam.full_proj_transform,
                                             cam.camera_center, tanfovx, tanfovy, H, W, shs=det["shs"], sh_degree=D)
st = ref.state()
d = (out["render"].detach() != color_r)
print("render(): colour differing values", int(d.sum()), "of", d.numel(), "max abs", float((out["render"].detach() - color_r).abs().max()), "depth diff", int((out["depth"].detach() != depth_r).sum()))
# direct forward with the same activated tensors
sc2 = Scene(pc.get_xyz.detach(), det["scales"], det["rotations"], det["opacity"], det["shs"], D)
cam2 = Camera(W, H, tanfovx, tanfovy, cam.world_view_transform, cam.full_proj_transform, cam.camera_center)
fw = mine_forward(sc2, cam2, bg)
print("direct: N", fw["N"], N_r, "colour diff", int((fw["color"] != color_r).sum()), "depth diff", int((fw["depth"] != depth_r).sum()),
      "rgb diff", int((fw["rgb"][radii_r > 0] != st["rgb"][radii_r > 0]).sum()), "conic diff", int((fw["conic_opacity"][radii_r > 0] != st["conic_opacity"][radii_r > 0]).sum()),
      "n_contrib diff", int((fw["n_contrib"] != st["n_contrib"]).sum()), "final_T diff", int((fw["final_T"] != st["final_T"]).sum()))
print("render() vs direct colour diff", int((out["render"].detach() != fw["color"]).sum()))
bad = (fw["rgb"] != st["rgb"]).any(1) & (radii_r > 0)
if bad.any():
    i = int(bad.nonzero()[0])
    print("first rgb mismatch at", i, fw["rgb"][i].cpu().numpy(), st["rgb"][i].cpu().numpy(), "clamped", fw["clamped"][i].item(), st["clamped"][i].cpu().numpy())
print("---- more")
print("ref rgb[:3]", st["rgb"][:3].cpu().numpy())
print("mine rgb[:3]", fw["rgb"][:3].cpu().numpy())
for name, cp in (("view slice", cam.camera_center), ("clone", cam.camera_center.clone()), ("cpu-made", cam0.campos.to(dev))):
    N2, c2, r2, d2 = ref.forward(bg, pc.get_xyz.detach(), det["opacity"], det["scales"], det["rotations"], cam.world_view_transform, cam.full_proj_transform,
                                 cp, tanfovx, tanfovy, H, W, shs=det["shs"], sh_degree=D)
    s2 = ref.state()
    print(name, "ptr%256", cp.data_ptr() % 256, "ref rgb[0]", s2["rgb"][0].cpu().numpy(), "colour diff vs mine", int((c2 != fw["color"]).sum()))
sh2 = sc.shs.to(dev)
print("det shs == scene shs", bool(torch.equal(sh2, det["shs"])), det["shs"].shape, det["shs"].stride(), det["shs"].data_ptr() % 256)
N2, c2, r2, d2 = ref.forward(bg, pc.get_xyz.detach(), det["opacity"], det["scales"], det["rotations"], cam.world_view_transform, cam.full_proj_transform,
                             cam.camera_center.clone(), tanfovx, tanfovy, H, W, shs=sh2, sh_degree=D)
print("scene shs tensor: ref rgb[0]", ref.state()["rgb"][0].cpu().numpy(), "colour diff vs mine", int((c2 != fw["color"]).sum()))
for DD in range(4):
    N2, c2, r2, d2 = ref.forward(bg, pc.get_xyz.detach(), det["opacity"], det["scales"], det["rotations"], cam.world_view_transform, cam.full_proj_transform,
                                 cam.camera_center.clone(), tanfovx, tanfovy, H, W, shs=sh2, sh_degree=DD)
    print("D", DD, ref.state()["rgb"][0].cpu().numpy())
