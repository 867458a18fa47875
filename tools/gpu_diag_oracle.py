"""Diagnostic: gradient agreement of the CUDA path with the deterministic CPU oracle at C1 / C2 under several criteria
(max-norm relative error; fraction of elements outside rtol * |b| + atol_frac * max|b|).  Used to set the thresholds of
tests/test_gpu_parity.py::test_configs_match_cpu_oracle at what the code actually achieves.
    python tools/gpu_diag_oracle.py [C1 C2]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from harness import elem_rel_err_frac, mine_backward, mine_forward, rel_err  # noqa: E402
from oracle.oracle_py import Oracle  # noqa: E402
from rain_b200.synthetic import CONFIGS, make_camera, make_grad, make_scene  # noqa: E402

dev = torch.device("cuda:0")
orc = Oracle()
for name in (sys.argv[1:] or ["C1", "C2"]):
    cfg = CONFIGS[name]
    sc_h = make_scene(cfg["P"], cfg["D"]); cam_h = make_camera(0, 1, cfg["W"], cfg["H"])
    sc, cam = sc_h.to(dev), cam_h.to(dev)
    bg = torch.zeros(3, device=dev); dpix = make_grad(cfg["W"], cfg["H"])
    fw = mine_forward(sc, cam, bg)
    g = mine_backward(fw, sc, cam, bg, dpix.to(dev))
    g2 = mine_backward(fw, sc, cam, bg, dpix.to(dev))
    N_o, color_o, radii_o, depth_o = orc.forward(np.zeros(3, np.float32), sc_h.means3D.numpy(), sc_h.opacities.numpy(), sc_h.scales.numpy(), sc_h.rotations.numpy(),
                                                 cam_h.viewmatrix.numpy(), cam_h.projmatrix.numpy(), cam_h.campos.numpy(), cam_h.tanfovx, cam_h.tanfovy,
                                                 cam_h.height, cam_h.width, shs=sc_h.shs.numpy(), sh_degree=cfg["D"])
    g_o = orc.backward(dpix.numpy())
    print(f"== {name}: N {fw['N']} vs {N_o}; colour {rel_err(fw['color'], color_o)[0]:.2e} depth {rel_err(fw['depth'], depth_o)[0]:.2e}")
    for k in ["dL_dmean2D", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"]:
        b = g_o[k].reshape(g[k].shape)
        row = [f"{elem_rel_err_frac(g[k], b, 1e-4, af):.2e}" for af in (1e-7, 1e-6, 1e-5, 1e-4)]
        print(f"  {k:12s} max-norm {rel_err(g[k], b)[0]:.2e}  self run-to-run {rel_err(g[k], g2[k])[0]:.2e}  elem frac>1e-4 @atol 1e-7/1e-6/1e-5/1e-4 of max: {row}")
