"""GPU diagnostic for the backward render: run-to-run spread of this library's gradients, of the reference CUDA's, and the
distance between the two and of each to the deterministic CPU oracle (a third summation order), on the cases that sit
closest to the 1e-4 tolerance (dense tiles; mid-size repro case).
Usage: python tools/gpu_diag_bwd.py [runs]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from harness import mine_forward, mine_backward, rel_err
from rain_b200.synthetic import make_scene, make_camera, make_grad
from oracle.oracle_py import Oracle, RefCuda

runs = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = torch.device("cuda:0")
ref = RefCuda()
orc = Oracle()
KEYS = ["dL_dmean2D", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"]
CASES = [(50000, 48, 32, 0, 4, 0, 4, 4.0), (20000, 200, 120, 3, 5, 0, 2, 1.0), (5000, 100, 70, 3, 1, 1, 4, 1.0)]
for P, W, H, D, seed, view, nviews, smult in CASES:
    sc_h = make_scene(P, D, seed, smult); cam_h = make_camera(view, nviews, W, H)
    sc, cam = sc_h.to(dev), cam_h.to(dev)
    bg = torch.zeros(3, device=dev); dpix = make_grad(W, H, seed).to(dev)
    fw = mine_forward(sc, cam, bg)
    mine = [mine_backward(fw, sc, cam, bg, dpix) for _ in range(runs)]
    ref.forward(bg, sc.means3D, sc.opacities, sc.scales, sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos, cam.tanfovx, cam.tanfovy,
                cam.height, cam.width, shs=sc.shs, sh_degree=sc.sh_degree)
    refs = [ref.backward(dpix) for _ in range(runs)]
    orc.forward(np.zeros(3, np.float32), sc_h.means3D.numpy(), sc_h.opacities.numpy(), sc_h.scales.numpy(), sc_h.rotations.numpy(),
                cam_h.viewmatrix.numpy(), cam_h.projmatrix.numpy(), cam_h.campos.numpy(), cam_h.tanfovx, cam_h.tanfovy,
                cam_h.height, cam_h.width, shs=sc_h.shs.numpy(), sh_degree=D)
    g_o = orc.backward(dpix.cpu().numpy())
    print(f"== P={P} {W}x{H} D={D}: N={fw['N']}")
    for k in KEYS:
        mm = max(rel_err(mine[i][k], mine[0][k])[0] for i in range(1, runs))
        rr = max(rel_err(refs[i][k], refs[0][k])[0] for i in range(1, runs))
        mr = [rel_err(mine[i][k], refs[j][k])[0] for i in range(runs) for j in range(runs)]
        o = g_o[k].reshape(mine[0][k].shape)
        mo = rel_err(mine[0][k], o); ro = rel_err(refs[0][k], o)
        print(f"  {k:12s} mine-vs-mine {mm:.2e}   ref-vs-ref {rr:.2e}   mine-vs-ref min {min(mr):.2e} max {max(mr):.2e}   "
              f"vs oracle: mine {mo[0]:.2e} ref {ro[0]:.2e}   L2: mine-ref {rel_err(mine[0][k], refs[0][k])[1]:.2e} mine-oracle {mo[1]:.2e} ref-oracle {ro[1]:.2e}")
