"""GPU diagnostic for the backward render: run-to-run spread of this library's gradients, of the reference CUDA's, and the
distance between the two, on the two cases that sit closest to the 1e-4 tolerance (dense tiles; mid-size repro case).
Usage: RAIN_BWD_REDUCE=<s|2|3> python tools/gpu_diag_bwd.py [runs]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from harness import mine_forward, mine_backward, rel_err
from rain_b200.synthetic import make_scene, make_camera, make_grad
from oracle.oracle_py import RefCuda

runs = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = torch.device("cuda:0")
ref = RefCuda()
KEYS = ["dL_dmean2D", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"]
CASES = [(50000, 48, 32, 0, 4, 0, 4, 4.0), (20000, 200, 120, 3, 5, 0, 2, 1.0), (5000, 100, 70, 3, 1, 1, 4, 1.0)]
print("reduce variant:", os.environ.get("RAIN_BWD_REDUCE", "(default)"))
for P, W, H, D, seed, view, nviews, smult in CASES:
    sc = make_scene(P, D, seed, smult).to(dev); cam = make_camera(view, nviews, W, H).to(dev)
    bg = torch.zeros(3, device=dev); dpix = make_grad(W, H, seed).to(dev)
    fw = mine_forward(sc, cam, bg)
    mine = [mine_backward(fw, sc, cam, bg, dpix) for _ in range(runs)]
    ref.forward(bg, sc.means3D, sc.opacities, sc.scales, sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos, cam.tanfovx, cam.tanfovy,
                cam.height, cam.width, shs=sc.shs, sh_degree=sc.sh_degree)
    refs = [ref.backward(dpix) for _ in range(runs)]
    print(f"== P={P} {W}x{H} D={D}: N={fw['N']}")
    for k in KEYS:
        mm = max(rel_err(mine[i][k], mine[0][k])[0] for i in range(1, runs))
        rr = max(rel_err(refs[i][k], refs[0][k])[0] for i in range(1, runs))
        mr = [rel_err(mine[i][k], refs[j][k])[0] for i in range(runs) for j in range(runs)]
        print(f"  {k:12s} mine-vs-mine {mm:.2e}   ref-vs-ref {rr:.2e}   mine-vs-ref min {min(mr):.2e} max {max(mr):.2e}")
