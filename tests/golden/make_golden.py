"""TEST INFRASTRUCTURE — generates the golden vectors under tests/golden/ by running the UNMODIFIED reference CUDA
(oracle/_ref/libref_rast.so, compiled from /root/reference by oracle/Makefile `ref`) on a B200:

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'      # then copy gpurun_out/golden/*.npz here

Each fwd_bwd_*.npz holds the inputs of one small render (seeded rain_b200.synthetic scene + camera) and every output of
the reference: num_rendered, radii, tiles_touched, sorted keys, point_list, ranges, n_contrib, final_T, colour, depth and
the nine gradient tensors.  knn_*.npz holds points and the reference distCUDA2 result.  The reference's backward uses
float atomics, so its gradients vary in the last bits from run to run; the stored ones are one such run.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle_py import RefCuda  # noqa: E402
from rain_b200.synthetic import make_camera, make_grad, make_scene  # noqa: E402

CASES = [
    # name, P, W, H, D, seed, view, nviews, bg, low_pass, scale_modifier, scale_mult, mode
    ("sh3_small", 1500, 112, 80, 3, 0, 0, 4, (0.0, 0.0, 0.0), 0.3, 1.0, 1.5, "sh"),
    ("sh0_c1like", 4000, 128, 128, 0, 1, 1, 4, (0.0, 0.0, 0.0), 0.3, 1.0, 1.0, "sh"),
    ("sh2_lowpass_bg", 1200, 100, 70, 2, 2, 2, 4, (0.9, 0.2, 0.4), 4.0, 1.3, 1.0, "sh"),
    ("precomp", 1000, 96, 96, 0, 3, 3, 4, (0.1, 0.1, 0.1), 0.3, 1.0, 2.0, "precomp"),
]


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    dev = torch.device("cuda:0")
    ref = RefCuda()
    for name, P, W, H, D, seed, view, nviews, bg, low_pass, smod, smult, mode in CASES:
        sc_h, cam_h = make_scene(P, D, seed, smult), make_camera(view, nviews, W, H)
        sc, cam = sc_h.to(dev), cam_h.to(dev)
        bg_t = torch.tensor(bg, dtype=torch.float32, device=dev)
        dpix = make_grad(W, H, seed + 100).to(dev)
        kw = dict(scale_modifier=smod, low_pass=low_pass)
        colors = cov = None
        if mode == "precomp":
            g = torch.Generator().manual_seed(seed)
            colors = torch.rand(P, 3, generator=g).to(dev)
            A = torch.randn(P, 3, 3, generator=g) * 0.1
            S = A @ A.transpose(1, 2) + 0.001 * torch.eye(3)
            cov = torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).contiguous().to(dev)
            N, color, radii, depth = ref.forward(bg_t, sc.means3D, sc.opacities, None, None, cam.viewmatrix, cam.projmatrix, cam.campos,
                                                 cam.tanfovx, cam.tanfovy, H, W, sh_degree=D, colors_precomp=colors, cov3D_precomp=cov, **kw)
        else:
            N, color, radii, depth = ref.forward(bg_t, sc.means3D, sc.opacities, sc.scales, sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos,
                                                 cam.tanfovx, cam.tanfovy, H, W, shs=sc.shs, sh_degree=D, **kw)
        st = ref.state()
        g = ref.backward(dpix)
        torch.cuda.synchronize()
        c = lambda t: t.detach().cpu().numpy()
        e = np.zeros((0,), np.float32)
        np.savez_compressed(
            os.path.join(out_dir, f"fwd_bwd_{name}.npz"),
            W=W, H=H, D=D, tanfovx=cam_h.tanfovx, tanfovy=cam_h.tanfovy, scale_modifier=smod, low_pass=low_pass, bg=np.float32(bg),
            means3D=c(sc.means3D), opacities=c(sc.opacities), scales=c(sc.scales) if mode == "sh" else e, rotations=c(sc.rotations) if mode == "sh" else e,
            shs=c(sc.shs) if mode == "sh" else e, colors_precomp=c(colors) if colors is not None else e, cov3D_precomp=c(cov) if cov is not None else e,
            viewmatrix=c(cam.viewmatrix), projmatrix=c(cam.projmatrix), campos=c(cam.campos), dL_dpix=c(dpix),
            num_rendered=N, radii=c(radii), color=c(color), depth=c(depth), tiles_touched=c(st["tiles_touched"]),
            keys=c(st["keys"]) if N else np.zeros((0,), np.int64), point_list=c(st["point_list"]) if N else np.zeros((0,), np.int32),
            ranges=c(st["ranges"]), n_contrib=c(st["n_contrib"]), final_T=c(st["final_T"]),
            means2D=c(st["means2D"]), conic_opacity=c(st["conic_opacity"]), rgb=c(st["rgb"]), depths=c(st["depths"]), cov3D=c(st["cov3D"]),
            **{k: c(v) for k, v in g.items()})
        print(f"{name}: P={P} {W}x{H} D={D} N={N}")
    for name, P, seed in [("uniform", 4000, 0), ("clustered", 3000, 1)]:
        gen = torch.Generator().manual_seed(seed)
        pts = (torch.rand(P, 3, generator=gen) * 2 - 1).float()
        if name == "clustered":
            pts[: P // 2] = torch.randn(P // 2, 3, generator=gen) * 0.01 + 0.5
            pts[10] = pts[11]
        d = ref.dist2(pts.to(dev))
        np.savez_compressed(os.path.join(out_dir, f"knn_{name}.npz"), points=pts.numpy(), dist2=d.cpu().numpy())
        print(f"knn_{name}: P={P}")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
