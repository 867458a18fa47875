"""GPU diagnostic: stage-by-stage comparison of rain_b200 against the rebuilt reference CUDA and the CPU oracle, with
mismatch statistics (not just pass/fail), plus rough per-call timings.  Usage: python tools/gpu_diag.py [C1 C2 C3] """
import sys, os, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
from harness import mine_forward, mine_backward, rel_err, elem_rel_err_frac
from rain_b200.synthetic import make_scene, make_camera, make_grad, CONFIGS
from oracle.oracle_py import Oracle, RefCuda

dev = torch.device("cuda:0")
names = sys.argv[1:] or ["C1", "C2"]
ref = RefCuda()
orc = Oracle()
report = {}

def cmp_bits(name, a, b, rep):
    a = a.cpu(); b = b.cpu()
    if a.dtype.is_floating_point:
        neq = (a != b) & ~(torch.isnan(a) & torch.isnan(b))
    else:
        neq = a != b
    rep[name] = dict(mismatch=int(neq.sum()), total=int(neq.numel()))
    print(f"   {name:16s} mismatches {int(neq.sum()):>10d} / {neq.numel()}")

for name in names:
    cfg = CONFIGS[name]
    rep = report[name] = {}
    sc_h = make_scene(cfg["P"], cfg["D"]); cam_h = make_camera(0, 1, cfg["W"], cfg["H"])
    sc = sc_h.to(dev); cam = cam_h.to(dev)
    bg = torch.zeros(3, device=dev); dpix = make_grad(cfg["W"], cfg["H"]).to(dev)
    print(f"== {name}: P={cfg['P']} {cfg['W']}x{cfg['H']} D={cfg['D']}")
    for variant in ("1", "0"):
        os.environ["RAIN_RAST_VARIANT"] = variant
        fw = mine_forward(sc, cam, bg); torch.cuda.synchronize()
        N_r, color_r, radii_r, depth_r = ref.forward(bg, sc.means3D, sc.opacities, sc.scales, sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos,
                                                     cam.tanfovx, cam.tanfovy, cam.height, cam.width, shs=sc.shs, sh_degree=sc.sh_degree)
        st = ref.state()
        print(f" variant {variant}: N mine {fw['N']} ref {N_r}")
        rep[f"v{variant}_N"] = [fw["N"], N_r]
        vis = radii_r > 0
        cmp_bits("radii", fw["radii"], radii_r, rep)
        cmp_bits("tiles_touched", fw["tiles_touched"], st["tiles_touched"], rep)
        cmp_bits("depths(vis)", fw["depths"][vis], st["depths"][vis], rep)
        cmp_bits("means2D(vis)", fw["means2D"][vis], st["means2D"][vis], rep)
        cmp_bits("conic_op(vis)", fw["conic_opacity"][vis], st["conic_opacity"][vis], rep)
        cmp_bits("rgb(vis)", fw["rgb"][vis], st["rgb"][vis], rep)
        cmp_bits("cov3D(vis)", fw["cov3D"][vis], st["cov3D"][vis], rep)
        if fw["N"] == N_r and N_r > 0:
            cmp_bits("point_list", fw["point_list"], st["point_list"], rep)
            cmp_bits("keys", fw["keys"], st["keys"], rep)
        cmp_bits("ranges", fw["ranges"], st["ranges"], rep)
        cmp_bits("n_contrib", fw["n_contrib"], st["n_contrib"], rep)
        cmp_bits("final_T", fw["final_T"], st["final_T"], rep)
        cmp_bits("color", fw["color"], color_r, rep)
        cmp_bits("depth", fw["depth"], depth_r, rep)
        print("   color rel(max,l2)", rel_err(fw["color"], color_r), "depth", rel_err(fw["depth"], depth_r))
        g = mine_backward(fw, sc, cam, bg, dpix); torch.cuda.synchronize()
        g2 = mine_backward(fw, sc, cam, bg, dpix); torch.cuda.synchronize()
        gr = ref.backward(dpix)
        gr2 = ref.backward(dpix)
        for k in ["dL_dmean2D", "dL_dconic", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"]:
            a, b = g[k], gr[k]
            if k == "dL_dconic":
                a = a.reshape(-1, 4)[:, [0, 1, 3]]; b = b.reshape(-1, 4)[:, [0, 1, 3]]
            e = rel_err(a, b); fr = elem_rel_err_frac(a, b, 1e-4)
            e_rr = rel_err(gr2[k], gr[k]); det = bool((g2[k] == g[k]).all())
            rep[f"v{variant}_{k}"] = dict(max_rel=e[0], l2_rel=e[1], frac_gt_1e_4=fr, ref_vs_ref_max=e_rr[0], deterministic=det)
            print(f"   {k:12s} mine-vs-ref max {e[0]:.2e} l2 {e[1]:.2e} frac>1e-4 {fr:.2e} | ref-vs-ref max {e_rr[0]:.2e} | mine deterministic {det}")
    # oracle comparison (CPU)
    if cfg["P"] <= 300_000:
        N_o, color_o, radii_o, depth_o = orc.forward(np.zeros(3, np.float32), sc_h.means3D.numpy(), sc_h.opacities.numpy(), sc_h.scales.numpy(), sc_h.rotations.numpy(),
            cam_h.viewmatrix.numpy(), cam_h.projmatrix.numpy(), cam_h.campos.numpy(), cam_h.tanfovx, cam_h.tanfovy, cam_h.height, cam_h.width, shs=sc_h.shs.numpy(), sh_degree=sc_h.sh_degree)
        print(f" oracle: N {N_o}")
        cmp_bits("orc radii", torch.from_numpy(radii_o), radii_r, rep)
        vis_c = vis.cpu()
        cmp_bits("orc tiles", torch.from_numpy(orc.get("tiles_touched").astype(np.int32)), st["tiles_touched"], rep)
        cmp_bits("orc depths", torch.from_numpy(orc.get("depths"))[vis_c], st["depths"][vis].cpu(), rep)
        cmp_bits("orc means2D", torch.from_numpy(orc.get("means2D")).view(-1, 2)[vis_c], st["means2D"][vis].cpu(), rep)
        cmp_bits("orc conic", torch.from_numpy(orc.get("conic_opacity")).view(-1, 4)[vis_c], st["conic_opacity"][vis].cpu(), rep)
        cmp_bits("orc rgb", torch.from_numpy(orc.get("rgb")).view(-1, 3)[vis_c], st["rgb"][vis].cpu(), rep)
        cmp_bits("orc cov3D", torch.from_numpy(orc.get("cov3D")).view(-1, 6)[vis_c], st["cov3D"][vis].cpu(), rep)
        if N_o == N_r:
            cmp_bits("orc point_list", torch.from_numpy(orc.get("point_list").astype(np.int32)), st["point_list"], rep)
            cmp_bits("orc keys", torch.from_numpy(orc.get("keys").astype(np.int64)), st["keys"], rep)
        cmp_bits("orc ranges", torch.from_numpy(orc.get("ranges").astype(np.int32)).view(-1, 2), st["ranges"], rep)
        cmp_bits("orc n_contrib", torch.from_numpy(orc.get("n_contrib").astype(np.int32)), st["n_contrib"], rep)
        print("   orc color rel", rel_err(color_o, color_r), "depth", rel_err(depth_o, depth_r))
        go = orc.backward(make_grad(cfg["W"], cfg["H"]).numpy())
        for k in ["dL_dmean2D", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"]:
            print(f"   orc {k:12s} oracle-vs-ref {rel_err(go[k], gr[k])}  mine-vs-oracle {rel_err(g[k], go[k])}")
    # timing (CUDA events, 10 reps)
    def timeit(fn, reps=10):
        fn(); torch.cuda.synchronize()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(reps): fn()
        e.record(); torch.cuda.synchronize()
        return s.elapsed_time(e) / reps
    os.environ["RAIN_RAST_VARIANT"] = "1"
    def mine_fb():
        fw = mine_forward.__wrapped__(sc, cam, bg) if hasattr(mine_forward, "__wrapped__") else None
    from rain_b200 import _C
    e_ = torch.empty(0, device=dev)
    args = (bg, sc.means3D, e_, sc.opacities, sc.scales, sc.rotations, 1.0, e_, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.height, cam.width, sc.shs, sc.sh_degree, cam.campos, False, False, 0.3)
    state = {}
    def m_f():
        state["o"] = _C.rasterize_gaussians(*args)
    def m_b():
        N, color, radii, depth, geom, binning, img = state["o"]
        _C.rasterize_gaussians_backward(bg, sc.means3D, radii, e_, sc.scales, sc.rotations, 1.0, e_, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, dpix, sc.shs, sc.sh_degree, cam.campos, geom, N, binning, img, False, 0.3)
    def r_f():
        ref.forward(bg, sc.means3D, sc.opacities, sc.scales, sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos, cam.tanfovx, cam.tanfovy, cam.height, cam.width, shs=sc.shs, sh_degree=sc.sh_degree)
    def r_b():
        ref.backward(dpix)
    for v in ("1", "0"):
        os.environ["RAIN_RAST_VARIANT"] = v
        tf, tb = timeit(m_f), timeit(m_b)
        print(f" timing mine v{v}: fwd {tf:.3f} ms bwd {tb:.3f} ms total {tf+tb:.3f} ms")
        rep[f"time_mine_v{v}"] = [tf, tb]
    tf, tb = timeit(r_f), timeit(r_b)
    print(f" timing ref : fwd {tf:.3f} ms bwd {tb:.3f} ms total {tf+tb:.3f} ms")
    rep["time_ref"] = [tf, tb]
os.makedirs("gpurun_out", exist_ok=True)
json.dump(report, open("gpurun_out/diag.json", "w"), indent=1)
