"""GPU parity tests (run with -m gpu on a B200).  Everything goes through the C ABI of librainrast.so (ctypes), and is
compared with
  * the UNMODIFIED reference CUDA rebuilt for sm_100a (oracle/_ref/libref_rast.so, fixture `refcuda`) on the same
    seeded inputs — bit-exact for integers/indices (num_rendered, radii, tiles_touched, sorted keys, point_list, ranges,
    n_contrib), 1e-5 relative for colour/depth (in practice bit-identical), 1e-4 relative for gradients (the
    reference's own backward is order-nondeterministic because of its float atomics);
  * the CPU oracle (oracle/liboracle.so) at sizes it finishes in seconds;
  * the golden vectors committed under tests/golden/ (reference CUDA outputs);
  * size-independent properties at BASELINE.json's full sizes (linearity, determinism, sortedness, range partition).
"""
import glob
import os

import numpy as np
import pytest
import torch

from harness import elem_rel_err_frac, mine_backward, mine_forward, rel_err
from rain_b200 import _C, _lib
from rain_b200.synthetic import CONFIGS, make_camera, make_grad, make_scene

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GRAD_KEYS = ["dL_dmean2D", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"]
RTOL_COLOR = 1e-5     # north_star: rendered colour and depth within 1e-5 relative
RTOL_GRAD = 1e-4      # north_star: input gradients within 1e-4 relative
# RAIN_TILE_CULL=1 (opt-in, csrc/preprocess.cu): tiles in which the reference's alpha test rejects a Gaussian at every pixel
# are not binned.  The internal lists then are sub-lists of the reference's; images must still be bit-identical.
TILE_CULL = os.environ.get("RAIN_TILE_CULL", "0") == "1"
XC_INNER = "RAIN_TEST_LIB" in os.environ          # this process runs against the cross-check build (test_crosscheck_build_with_tma_staging)


def eq(a, b):
    return bool(torch.equal(torch.as_tensor(a).cpu(), torch.as_tensor(b).cpu()))


def ref_forward(ref, sc, cam, bg, **kw):
    return ref.forward(bg, sc.means3D, sc.opacities, sc.scales, sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos, cam.tanfovx, cam.tanfovy,
                       cam.height, cam.width, shs=sc.shs, sh_degree=sc.sh_degree, **kw)


def oracle_grads_of(oracle, sc, cam, bg, dpix, **kw):
    """the deterministic CPU oracle's gradients for the same inputs, computed only if asked for (check_against_ref)"""
    def run():
        c = lambda t: t.detach().cpu().numpy()
        oracle.forward(c(bg), c(sc.means3D), c(sc.opacities), c(sc.scales), c(sc.rotations), c(cam.viewmatrix), c(cam.projmatrix), c(cam.campos),
                       cam.tanfovx, cam.tanfovy, cam.height, cam.width, shs=c(sc.shs), sh_degree=sc.sh_degree, **kw)
        return oracle.backward(c(dpix))
    return run


def check_against_ref(fw, g, ref, N_r, color_r, radii_r, depth_r, g_r, dpix=None, oracle_grads=None):
    """Integer outputs bit-exact, colour/depth 1e-5, gradients 1e-4 PLUS the reference's own run-to-run spread: its
    backward adds per pixel with unordered atomics, and on gradients with heavy cancellation (dL_drot, dL_dscale) two of
    its own runs differ by 1e-5 .. 3e-4 of max (tools/gpu_diag_bwd.py), so a fixed 1e-4 against ONE of its runs is a coin
    flip there.  The spread is measured here from three more reference backward runs on the same state.
    oracle_grads (optional, a callable): the spread of a few runs does not show the part of the reference's rounding error
    that repeats from run to run (its atomics arrive in much the same order every time: on the dense-tile case its dL_dcov3D
    sits 1e-5 from the deterministic oracle while two of its runs differ by 2e-6, and dL_drot amplifies that tenfold —
    tools/gpu_diag_bwd.py).  A gradient that misses the reference run by more than the tolerance must then be within a plain
    1e-4 of the ORACLE, the deterministic restatement of the same algorithm; without oracle_grads it fails."""
    st = ref.state()
    spread = {}
    g_o = None
    if g is not None and dpix is not None:
        again = [ref.backward(dpix) for _ in range(3)]
        spread = {k: max(rel_err(x[k], g_r[k])[0] for x in again) for k in GRAD_KEYS}
    assert eq(fw["radii"], radii_r)
    if TILE_CULL:
        assert fw["N"] <= N_r and (fw["tiles_touched"].long().cpu() <= torch.as_tensor(st["tiles_touched"]).long().cpu()).all()
        if fw["N"]:
            # every (key, Gaussian) pair of ours is one of the reference's, in the same order
            mine = torch.stack([fw["keys"].long(), fw["point_list"].long()], 1).cpu()
            ref_pairs = torch.stack([torch.as_tensor(st["keys"]).long(), torch.as_tensor(st["point_list"]).long()], 1).cpu()
            assert (mine[1:, 0] >= mine[:-1, 0]).all()
            # pairs are unique (one per tile and Gaussian): membership through a combined 64-bit code
            code = lambda t: (t[:, 0] >> 32) * (1 << 24) + t[:, 1]            # tile id * 2^24 + Gaussian id (P < 2^24 here)
            if int(ref_pairs[:, 1].max()) < (1 << 24):
                assert torch.isin(code(mine), code(ref_pairs)).all()
        # the images are bit-identical, not just within tolerance
        assert eq(fw["final_T"], st["final_T"]) and eq(fw["color"], color_r) and eq(fw["depth"], depth_r)
    else:
        assert fw["N"] == N_r
        assert eq(fw["tiles_touched"], st["tiles_touched"])
        if N_r:
            assert eq(fw["keys"], st["keys"])
            assert eq(fw["point_list"], st["point_list"])
        assert eq(fw["ranges"], st["ranges"])
        assert eq(fw["n_contrib"], st["n_contrib"])
    assert rel_err(fw["final_T"], st["final_T"])[0] <= RTOL_COLOR
    assert rel_err(fw["color"], color_r)[0] <= RTOL_COLOR
    assert rel_err(fw["depth"], depth_r)[0] <= RTOL_COLOR
    if g is not None:
        for k in GRAD_KEYS:
            err = rel_err(g[k], g_r[k])[0]
            if err > RTOL_GRAD + 2.0 * spread.get(k, 0.0) and oracle_grads is not None:
                g_o = g_o if g_o is not None else oracle_grads()
                err_o = rel_err(g[k], g_o[k].reshape(g[k].shape))[0]
                assert err_o <= RTOL_GRAD, f"{k}: {err:.3e} vs the reference run (its spread {spread.get(k, 0.0):.3e}) and {err_o:.3e} vs the oracle"
                continue
            assert err <= RTOL_GRAD + 2.0 * spread.get(k, 0.0), f"{k}: {err:.3e} vs the reference run; the reference's own run-to-run spread here is {spread.get(k, 0.0):.3e}"
        a = g["dL_dconic"].reshape(-1, 4)[:, [0, 1, 3]]; b = g_r["dL_dconic"].reshape(-1, 4)[:, [0, 1, 3]]
        assert rel_err(a, b)[0] <= RTOL_GRAD


@pytest.mark.parametrize("name", ["C1", "C2", "C3", "C4"])
def test_configs_match_reference_cuda(device, refcuda, name):
    """BASELINE.json configs 1-4 at full size (C4: the 6M-Gaussian scene, one of its views) against the reference CUDA on the
    same GPU: integers bit-exact, colour / depth bit-identical, gradients within tolerance."""
    cfg = CONFIGS[name]
    sc = make_scene(cfg["P"], cfg["D"]).to(device); cam = make_camera(0, 1, cfg["W"], cfg["H"]).to(device)
    bg = torch.zeros(3, device=device); dpix = make_grad(cfg["W"], cfg["H"]).to(device)
    fw = mine_forward(sc, cam, bg)
    g = mine_backward(fw, sc, cam, bg, dpix)
    N_r, color_r, radii_r, depth_r = ref_forward(refcuda, sc, cam, bg)
    g_r = refcuda.backward(dpix)
    assert N_r > cfg["P"]
    check_against_ref(fw, g, refcuda, N_r, color_r, radii_r, depth_r, g_r, dpix)
    if not TILE_CULL:
        # beyond the 1e-5 tolerance of the north star: the forward images are the reference's bits
        assert eq(fw["color"], color_r) and eq(fw["depth"], depth_r) and eq(fw["final_T"], refcuda.state()["final_T"])


@pytest.mark.parametrize("P,W,H,D,seed,view,bgc,low_pass,smod,smult", [
    (5000, 100, 70, 3, 1, 1, (0.3, 0.6, 0.9), 0.3, 1.0, 1.0),        # image not a multiple of 16, non-zero background
    (3000, 64, 64, 2, 2, 2, (0.0, 0.0, 0.0), 30.0, 1.0, 1.0),         # large coarse-to-fine low_pass: huge rects, long tile lists
    (2000, 160, 96, 1, 3, 3, (1.0, 1.0, 1.0), 0.3, 2.5, 3.0),         # scale_modifier, fat opaque splats: early termination
    (50000, 48, 32, 0, 4, 0, (0.0, 0.0, 0.0), 0.3, 1.0, 4.0),         # > 128 entries in every tile: multi-batch staging, saturation
    (1, 32, 32, 3, 5, 0, (0.5, 0.5, 0.5), 0.3, 1.0, 1.0),             # single Gaussian
    (257, 17, 15, 3, 6, 1, (0.2, 0.2, 0.2), 0.3, 1.0, 2.0),           # image smaller than one tile row/column multiple
    (6000, 100, 52, 3, 7, 1, (0.1, 0.2, 0.3), 0.3, 1.0, 2.0),         # H % 16 == 4: the lower 8x4 half of the last row's upper 8x8 blocks is outside
    (6000, 90, 41, 2, 8, 2, (0.0, 0.0, 0.0), 0.3, 1.0, 3.0),          # H % 16 == 9: one pixel row of the lower 8x8 blocks is inside
    (30000, 72, 36, 1, 9, 0, (0.3, 0.3, 0.3), 0.3, 1.0, 5.0),         # H % 16 == 4 with long, early-saturating lists (flag words of a half that stops early)
])
def test_shapes_and_settings_match_reference_cuda(device, refcuda, oracle, P, W, H, D, seed, view, bgc, low_pass, smod, smult):
    sc = make_scene(P, D, seed, smult).to(device); cam = make_camera(view, 4, W, H).to(device)
    if P == 1:
        sc.means3D.zero_()
    bg = torch.tensor(bgc, device=device); dpix = make_grad(W, H, seed).to(device)
    fw = mine_forward(sc, cam, bg, low_pass=low_pass, scale_modifier=smod)
    g = mine_backward(fw, sc, cam, bg, dpix)
    N_r, color_r, radii_r, depth_r = ref_forward(refcuda, sc, cam, bg, low_pass=low_pass, scale_modifier=smod)
    g_r = refcuda.backward(dpix)
    check_against_ref(fw, g, refcuda, N_r, color_r, radii_r, depth_r, g_r, dpix,
                      oracle_grads=oracle_grads_of(oracle, sc, cam, bg, dpix, low_pass=low_pass, scale_modifier=smod))


@pytest.mark.parametrize("P,W,H,opaque", [(200_000, 40, 24, True), (200_000, 40, 24, False), (60_000, 33, 50, True), (700, 40, 24, True),
                                          (200_000, 40, 20, True), (120_000, 40, 25, False)])
def test_forward_ring_handshake_long_lists_early_saturation(device, refcuda, P, W, H, opaque):
    """The forward's stage ring (csrc/render_fwd.cu: full / empty mbarriers, ghost warps, the producer's early stop): tile
    lists of hundreds of batches whose pixels saturate within the first ones (opaque), or never (translucent: every batch is
    walked), on images whose last tile row / column is partly outside (warps that are ghosts from the start), and lists that
    end inside the first ring round (fewer than FWD_STAGES batches).  Everything against the reference CUDA."""
    sc = make_scene(P, 1, seed=11, scale_mult=3.0).to(device); cam = make_camera(1, 4, W, H).to(device)
    if opaque:
        sc.opacities.fill_(0.999)
    else:
        sc.opacities.mul_(0.02)
    bg = torch.tensor((0.1, 0.2, 0.3), device=device); dpix = make_grad(W, H, 5).to(device)
    fw = mine_forward(sc, cam, bg)
    g = mine_backward(fw, sc, cam, bg, dpix)
    N_r, color_r, radii_r, depth_r = ref_forward(refcuda, sc, cam, bg)
    g_r = refcuda.backward(dpix)
    if P >= 60_000:
        assert N_r > 128 * 8 * ((W + 15) // 16) * ((H + 15) // 16)      # more than two ring rounds per tile on average
    check_against_ref(fw, g, refcuda, N_r, color_r, radii_r, depth_r, g_r, dpix)


@pytest.mark.parametrize("opaque", [True, False])
def test_forward_is_bit_reproducible_over_many_runs(device, opaque):
    """The forward has no atomics, so every run must give the same bits: 60 runs of the long-list scenes of the ring
    handshake test (a missed wait, a stage refilled too early or a ghost warp released at the wrong batch would show up as a
    run that differs), interleaved with a second scene so that the shared-memory contents the CTAs find vary."""
    W, H = 40, 24
    sc = make_scene(200_000, 1, seed=11, scale_mult=3.0).to(device); cam = make_camera(1, 4, W, H).to(device)
    other = make_scene(30_000, 1, seed=12, scale_mult=2.0).to(device)
    if opaque:
        sc.opacities.fill_(0.999)
    else:
        sc.opacities.mul_(0.02)
    bg = torch.tensor((0.1, 0.2, 0.3), device=device)
    first = mine_forward(sc, cam, bg)
    keep = {k: first[k].clone() for k in ("color", "depth", "final_T", "n_contrib")}
    for it in range(60):
        if it % 3 == 0:
            mine_forward(other, cam, bg)
        fw = mine_forward(sc, cam, bg)
        for k, v in keep.items():
            assert eq(fw[k], v), (it, k)


def test_precomputed_colour_and_covariance_modes(device, refcuda):
    P, W, H = 4000, 96, 80
    sc = make_scene(P, 0, 9, 1.5).to(device); cam = make_camera(1, 3, W, H).to(device)
    gen = torch.Generator().manual_seed(1)
    colors = torch.rand(P, 3, generator=gen).to(device)
    A = torch.randn(P, 3, 3, generator=gen) * 0.05
    S = A @ A.transpose(1, 2) + 0.0005 * torch.eye(3)
    cov = torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).contiguous().to(device)
    bg = torch.tensor([0.1, 0.0, 0.2], device=device); dpix = make_grad(W, H, 2).to(device)
    for cp, cv in ((colors, None), (None, cov), (colors, cov)):
        fw = mine_forward(sc, cam, bg, colors_precomp=cp, cov3D_precomp=cv)
        g = mine_backward(fw, sc, cam, bg, dpix)
        N_r, color_r, radii_r, depth_r = refcuda.forward(bg, sc.means3D, sc.opacities, None if cv is not None else sc.scales,
                                                         None if cv is not None else sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos,
                                                         cam.tanfovx, cam.tanfovy, H, W, shs=None if cp is not None else sc.shs, sh_degree=0,
                                                         colors_precomp=cp, cov3D_precomp=cv)
        g_r = refcuda.backward(dpix)
        if cp is not None:
            g_r["dL_dsh"] = g["dL_dsh"]          # M = 0: nothing to compare
        if cv is not None:
            g_r["dL_dscale"], g_r["dL_drot"] = g["dL_dscale"], g["dL_drot"]   # the reference leaves them untouched (zeros)
            assert not g["dL_dscale"].any() and not g["dL_drot"].any()
        check_against_ref(fw, g, refcuda, N_r, color_r, radii_r, depth_r, g_r, dpix)


# elementwise criterion against the deterministic oracle: an element passes if |a - b| <= 1e-4 |b| + ELEM_ATOL_FRAC * max|b|
# (the absolute term covers elements that are sums with heavy cancellation: both sides accumulate in fp32, in different
# orders); at most ELEM_FRAC_MAX of the elements of any gradient tensor may fail it
# (measured, profiles/r02_oracle_agreement_v1.log: at most 1.0e-4 of the elements of any tensor at C1 / C2 with atol = 1e-6 max|b|,
# none with 1e-5; max-norm errors 6e-7 .. 2.2e-5)
ELEM_ATOL_FRAC = 1e-6
ELEM_FRAC_MAX = 5e-4


@pytest.mark.skipif(TILE_CULL, reason="compares the internal lists with the oracle's")
@pytest.mark.parametrize("name", ["C1", "C2"])
def test_configs_match_cpu_oracle(device, oracle, name):
    """BASELINE.json configs 1 (10k Gaussians, 256x256, SH degree 0) and 2 (300k, 800x800, SH degree 3) against the
    DETERMINISTIC host restatement: integers bit-exact, colour / depth 1e-5, and every gradient within a plain 1e-4 of max —
    no allowance for run-to-run spread, the oracle has none — plus the elementwise criterion above."""
    cfg = CONFIGS[name]
    sc_h = make_scene(cfg["P"], cfg["D"]); cam_h = make_camera(0, 1, cfg["W"], cfg["H"])
    sc, cam = sc_h.to(device), cam_h.to(device)
    bg = torch.zeros(3, device=device); dpix = make_grad(cfg["W"], cfg["H"])
    fw = mine_forward(sc, cam, bg)
    g = mine_backward(fw, sc, cam, bg, dpix.to(device))
    N_o, color_o, radii_o, depth_o = oracle.forward(np.zeros(3, np.float32), sc_h.means3D.numpy(), sc_h.opacities.numpy(), sc_h.scales.numpy(),
                                                    sc_h.rotations.numpy(), cam_h.viewmatrix.numpy(), cam_h.projmatrix.numpy(), cam_h.campos.numpy(),
                                                    cam_h.tanfovx, cam_h.tanfovy, cam_h.height, cam_h.width, shs=sc_h.shs.numpy(), sh_degree=cfg["D"])
    g_o = oracle.backward(dpix.numpy())
    assert fw["N"] == N_o
    assert np.array_equal(fw["radii"].cpu().numpy(), radii_o)
    assert np.array_equal(fw["keys"].cpu().numpy().astype(np.uint64), oracle.get("keys"))
    assert np.array_equal(fw["point_list"].cpu().numpy().astype(np.uint32), oracle.get("point_list"))
    assert np.array_equal(fw["ranges"].cpu().numpy().astype(np.uint32).ravel(), oracle.get("ranges"))
    # the blend's alpha / T thresholds see exp() — libdevice's on the GPU (and in the reference), glibc's in the oracle (the one
    # arithmetic difference between them) — so a pixel in a few hundred thousand may stop one entry apart
    nc_bad = float((fw["n_contrib"].cpu().numpy().astype(np.uint32) != oracle.get("n_contrib")).mean())
    assert nc_bad <= 2e-5, f"n_contrib differs from the oracle on {nc_bad:.2e} of the pixels"
    assert rel_err(fw["color"], color_o)[0] <= (RTOL_COLOR if nc_bad == 0 else 1e-4) and rel_err(fw["depth"], depth_o)[0] <= (RTOL_COLOR if nc_bad == 0 else 1e-4)
    for k in GRAD_KEYS:
        ref_k = g_o[k].reshape(g[k].shape)
        assert rel_err(g[k], ref_k)[0] <= RTOL_GRAD, k
        frac = elem_rel_err_frac(g[k], ref_k, RTOL_GRAD, ELEM_ATOL_FRAC)
        assert frac <= ELEM_FRAC_MAX, f"{k}: {frac:.2e} of the elements differ from the oracle by more than 1e-4 relative"


GOLDEN_FILES = sorted(glob.glob(os.path.join(GOLDEN, "fwd_bwd_*.npz")))


@pytest.mark.parametrize("path", GOLDEN_FILES or [None])
def test_golden_vectors(device, path):
    if path is None:
        pytest.skip("no golden vectors committed yet")
    z = np.load(path)
    t = lambda k: torch.from_numpy(z[k]).to(device)
    has = lambda k: z[k].size > 0
    e = torch.empty(0, device=device)
    args = (t("bg"), t("means3D"), t("colors_precomp") if has("colors_precomp") else e, t("opacities"), t("scales") if has("scales") else e,
            t("rotations") if has("rotations") else e, float(z["scale_modifier"]), t("cov3D_precomp") if has("cov3D_precomp") else e, t("viewmatrix"),
            t("projmatrix"), float(z["tanfovx"]), float(z["tanfovy"]), int(z["H"]), int(z["W"]), t("shs") if has("shs") else e, int(z["D"]), t("campos"),
            False, False, float(z["low_pass"]))
    N, color, radii, depth, geom, binning, img = _C.rasterize_gaussians(*args)
    assert N <= int(z["num_rendered"]) if TILE_CULL else N == int(z["num_rendered"])
    assert np.array_equal(radii.cpu().numpy(), z["radii"])
    assert rel_err(color, z["color"])[0] <= RTOL_COLOR and rel_err(depth, z["depth"])[0] <= RTOL_COLOR
    grads = _C.rasterize_gaussians_backward(args[0], args[1], radii, args[2], args[4], args[5], args[6], args[7], args[8], args[9], args[10], args[11],
                                            t("dL_dpix"), args[14], args[15], args[16], geom, N, binning, img, False, args[19])
    for k, gr in zip(["dL_dmean2D", "dL_dcolor", "dL_dopacity", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"], grads):
        if z[k].size and gr.numel():
            assert rel_err(gr, z[k].reshape(gr.shape))[0] <= RTOL_GRAD, k


def test_empty_and_fully_culled_inputs(device):
    cam = make_camera(0, 1, 64, 48).to(device)
    bg = torch.tensor([0.25, 0.5, 0.75], device=device)
    e = torch.empty(0, device=device)
    z3 = torch.zeros(0, 3, device=device)
    N, color, radii, depth, *_ = _C.rasterize_gaussians(bg, z3, e, torch.zeros(0, 1, device=device), z3, torch.zeros(0, 4, device=device), 1.0, e, cam.viewmatrix,
                                                        cam.projmatrix, cam.tanfovx, cam.tanfovy, 48, 64, torch.zeros(0, 16, 3, device=device), 3, cam.campos, False, False, 0.3)
    assert N == 0 and radii.numel() == 0 and color.shape == (3, 48, 64) and not color.any()        # rasterize_points.cu:58-72: zeros, not bg
    sc = make_scene(100, 3, 0).to(device)
    sc.means3D[:] = cam.campos * 2.0                                                                # everything behind the camera
    fw = mine_forward(sc, cam, bg)
    assert fw["N"] == 0 and not fw["radii"].any() and not fw["depth"].any()
    assert torch.allclose(fw["color"], bg[:, None, None].expand(3, 48, 64))
    g = mine_backward(fw, sc, cam, bg, make_grad(64, 48).to(device))
    assert all(not v.any() for v in g.values())
    assert not _C.mark_visible(sc.means3D, cam.viewmatrix, cam.projmatrix).any()


def test_mark_visible_matches_oracle_and_reference(device, oracle, refcuda):
    sc_h = make_scene(20000, 0, 3, 1.0); cam_h = make_camera(2, 5, 64, 64, radius=1.0)               # camera inside the cloud
    m, v, pj = sc_h.means3D.to(device), cam_h.viewmatrix.to(device), cam_h.projmatrix.to(device)
    got = _C.mark_visible(m, v, pj)
    exp = oracle.mark_visible(sc_h.means3D.numpy(), cam_h.viewmatrix.numpy(), cam_h.projmatrix.numpy())
    assert 0 < exp.sum() < exp.size and np.array_equal(got.cpu().numpy(), exp)
    assert eq(got, refcuda.mark_visible(m, v, pj))                                                   # Rasterizer::markVisible itself


def test_autograd_api_equals_direct_backward_and_is_reproducible(device):
    from rain_b200 import GaussianRasterizationSettings, GaussianRasterizer
    P, W, H, D = 20000, 200, 120, 3
    sc = make_scene(P, D, 5).to(device); cam = make_camera(0, 2, W, H).to(device)
    bg = torch.tensor([0.1, 0.2, 0.3], device=device); dpix = make_grad(W, H, 5).to(device)
    rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, bg, 1.0, cam.viewmatrix, cam.projmatrix, D, cam.campos, False, True, 0.3)
    outs = []
    for _ in range(2):
        leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.shs, sc.scales, sc.rotations)]
        m2d = torch.zeros_like(sc.means3D, requires_grad=True)
        color, radii, depth = GaussianRasterizer(rs)(means3D=leaves[0], means2D=m2d, opacities=leaves[1], shs=leaves[2], scales=leaves[3], rotations=leaves[4])
        assert not radii.requires_grad and color.requires_grad
        (color * dpix).sum().backward()
        outs.append([color.detach(), radii, depth.detach(), m2d.grad] + [t.grad for t in leaves])
    for a, b in zip(outs[0][:3], outs[1][:3]):
        assert eq(a, b)                                                    # forward: bit-reproducible
    for a, b in zip(outs[0][3:], outs[1][3:]):
        # backward: one RED.ADD per (warp, Gaussian) visit, so only the order of a handful of float adds per Gaussian
        # varies run to run: measured <= 1e-6 of max except on dL_drot (heavy cancellation: up to 2.6e-5; the reference,
        # with one atomic per pixel, differs from itself by 3.4e-4 of max on this very case, tools/gpu_diag_bwd.py)
        assert rel_err(a, b)[0] < 5e-5
    fw = mine_forward(sc, cam, bg)
    g = mine_backward(fw, sc, cam, bg, dpix)
    for a, k in zip(outs[0][3:], ["dL_dmean2D", "dL_dmean3D", "dL_dopacity", "dL_dsh", "dL_dscale", "dL_drot"]):
        assert rel_err(a, g[k])[0] < 5e-5, k
    # two backward passes of ONE forward (retain_graph): the first accumulates in the geometry blob's region that the forward
    # cleared while it rendered, the second must not trust that region again (it fills its own scratch) — same gradients both times
    leaves = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.shs, sc.scales, sc.rotations)]
    m2d = torch.zeros_like(sc.means3D, requires_grad=True)
    color, radii, depth = GaussianRasterizer(rs)(means3D=leaves[0], means2D=m2d, opacities=leaves[1], shs=leaves[2], scales=leaves[3], rotations=leaves[4])
    loss = (color * dpix).sum()
    loss.backward(retain_graph=True)
    first = [m2d.grad.clone()] + [t.grad.clone() for t in leaves]
    for t in [m2d] + leaves:
        t.grad = None
    loss.backward()
    for a, b, c in zip(first, [m2d.grad] + [t.grad for t in leaves], outs[0][3:]):
        assert rel_err(a, b)[0] < 5e-5 and rel_err(a, c)[0] < 5e-5


def test_full_size_properties_config3(device):
    """Size-independent properties at BASELINE.json's headline size (1M Gaussians, 1920x1080, SH degree 3)."""
    cfg = CONFIGS["C3"]
    sc = make_scene(cfg["P"], cfg["D"]).to(device); cam = make_camera(0, 1, cfg["W"], cfg["H"]).to(device)
    bg = torch.tensor([0.3, 0.3, 0.3], device=device)
    fw = mine_forward(sc, cam, bg)
    N, keys, pl, ranges = fw["N"], fw["keys"], fw["point_list"].long(), fw["ranges"].long()
    assert N == int(fw["tiles_touched"].long().sum()) and (TILE_CULL or ((fw["radii"] > 0) == (fw["tiles_touched"] > 0)).all())
    d = keys[1:] - keys[:-1]
    assert (d >= 0).all()                                                  # sorted by (tile, depth)
    assert (pl[1:][d == 0] > pl[:-1][d == 0]).all()                        # stable: ties keep ascending Gaussian index
    assert eq(keys & 0xFFFFFFFF, fw["depths"].contiguous().view(torch.int32)[pl].long() & 0xFFFFFFFF)
    cnt = torch.bincount(keys >> 32, minlength=ranges.shape[0])
    assert eq(ranges[:, 1] - ranges[:, 0], cnt)
    nz = cnt > 0
    assert eq(ranges[nz, 0], (torch.cumsum(cnt, 0) - cnt)[nz]) and not ranges[~nz].any()
    assert eq(torch.bincount(pl, minlength=cfg["P"]), fw["tiles_touched"].long())
    assert (fw["final_T"] >= 0).all() and (fw["final_T"] <= 1).all() and torch.isfinite(fw["color"]).all()
    # linearity of the backward in the upstream gradient
    d1, d2 = make_grad(cfg["W"], cfg["H"], 1).to(device), make_grad(cfg["W"], cfg["H"], 2).to(device)
    g1, g2, g12 = mine_backward(fw, sc, cam, bg, d1), mine_backward(fw, sc, cam, bg, d2), mine_backward(fw, sc, cam, bg, 2.0 * d1 - 0.5 * d2)
    for k in GRAD_KEYS:
        assert rel_err(g12[k], 2.0 * g1[k] - 0.5 * g2[k])[0] < 1e-4, k
    # a zero upstream gradient gives exactly zero everywhere (every output element is written)
    g0 = mine_backward(fw, sc, cam, bg, torch.zeros_like(d1))
    assert all(not v.any() for v in g0.values())


@pytest.mark.parametrize("pipeline,P,W,H,V,batch,fbatch", [
    ("two_stream", 30000, 160, 96, 5, 8, 8), ("one_stream", 30000, 160, 96, 5, 8, 8), ("two_stream", 30000, 160, 96, 5, 2, 2), ("two_stream", 30000, 160, 96, 5, 0, 0),
    ("two_stream", 30000, 160, 96, 5, 0, 3), ("two_stream", 30000, 160, 96, 5, 8, 0), ("one_stream", 30000, 160, 96, 19, 19, 16), ("two_stream", 30000, 160, 96, 19, 8, 8),
    ("two_stream", 30001, 160, 96, 3, 8, 8), ("two_stream", 30002, 100, 52, 3, 2, 1), ("one_stream", 29999, 160, 96, 2, 0, 0), ("two_stream", 30001, 160, 96, 3, 0, 8),
    ("two_stream", 1_000_000, 1920, 1080, 3, 8, 8), ("two_stream", 1_000_000, 1920, 1080, 3, 0, 0)])
def test_accumulate_mode_and_multiview_bucket(device, pipeline, P, W, H, V, batch, fbatch):
    """The view-batch driver (both pipelines; the view-batched geometry pass in chunks of `fbatch` views, 0 = per view; the
    view-batched backward in groups of `batch` views — 19 views in one call span two kernel launches — and the per-view
    accumulate path, batch = 0) against per-view forward + backward through the drop-in
    ops: images bit-identical, bucket = sum of the per-view gradients, per-view viewspace gradients equal.  P % 4 != 0 puts the
    segments of the flat bucket at addresses that are not 16-byte aligned; the last cases are BASELINE.json config 3's size."""
    from rain_b200 import multiview as mv
    D = 3
    sc = make_scene(P, D, 8).to(device)
    cams = [make_camera(v, V, W, H).to(device) for v in range(V)]
    bg = torch.zeros(3, device=device)
    dp = [make_grad(W, H, 20 + v).to(device) for v in range(V)]
    params = dict(means3D=sc.means3D, opacities=sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    bucket = mv.GradientBucket(P, 16, device)
    bucket.flat.fill_(float("nan"))                                        # the first view must overwrite everything
    from rain_b200.optim import DensificationStats
    stats = DensificationStats(P, device)
    res = mv.train_step_views(params, cams, V, bucket, lambda v, c, d: dp[v], bg=bg, sh_degree=D, pipeline=pipeline, batch_backward=batch, batch_forward=fbatch, stats=stats)
    stats.sync()
    ref_stats = DensificationStats(P, device)
    tot = None
    for v in range(V):
        fw = mine_forward(sc, cams[v], bg)
        g = mine_backward(fw, sc, cams[v], bg, dp[v])
        assert eq(res[v].color, fw["color"]) and eq(res[v].radii, fw["radii"]) and res[v].num_rendered == fw["N"]
        # two runs of the same backward tile renderer: its per-Gaussian sums arrive through unordered fp32 atomics, and over a
        # million Gaussians the largest of the differences reaches a few 1e-6 of max (3.0e-6 seen); small scenes stay below 2e-6
        assert rel_err(res[v].dL_dmeans2D, g["dL_dmean2D"])[0] < (2e-6 if P < 100_000 else 1e-5)
        ref_stats.add_view(fw["radii"], g["dL_dmean2D"])                  # rr_densify_stats per view (train.py:132-134)
        cur = dict(means3D=g["dL_dmean3D"], sh=g["dL_dsh"], opacity=g["dL_dopacity"], scales=g["dL_dscale"], rotations=g["dL_drot"])
        tot = cur if tot is None else {k: tot[k] + cur[k] for k in cur}
    for k in mv.SEGMENTS:
        assert torch.isfinite(bucket.views[k]).all()
        # summation order differs (the bucket adds view by view or sums in registers, `tot` adds tensors): 1e-5 at the small
        # sizes; at 1M Gaussians dL_dscale / dL_drot (sums with heavy cancellation; their run-to-run spread alone is 5e-6 /
        # 1.6e-5 at C2, profiles/r02_oracle_agreement_v1.log) get 6e-5 — still under the north star's 1e-4
        assert rel_err(bucket.views[k], tot[k])[0] < (1e-5 if P < 100_000 else 6e-5), k
    # densification statistics folded into the batched backward == the per-view kernel's
    ref_stats.sync()
    assert eq(stats.max_radii2D, ref_stats.max_radii2D) and eq(stats.denom, ref_stats.denom)
    assert rel_err(stats.xyz_gradient_accum, ref_stats.xyz_gradient_accum)[0] < (5e-6 if P < 100_000 else 2e-5)      # sums of the norms bounded above


@pytest.mark.skipif(TILE_CULL, reason="the view-batched geometry pass has no tile culling")
@pytest.mark.parametrize("P,W,H,V", [(30001, 160, 96, 5), (20000, 100, 52, 16), (1_000_000, 1920, 1080, 3)])
def test_batched_geometry_pass_is_bit_identical_to_the_per_view_one(device, P, W, H, V):
    """rr_forward_geom_views (one pass over the Gaussians for all cameras, segmented depth sort, segmented scan) against
    rr_forward_geom per view: radii, records, clamp flags, tile counts, depth order, offsets, num_rendered and — after the
    binning / render half — images, tile lists and ranges, all bit for bit."""
    import ctypes
    lib = _lib.load()
    D = 3
    sc = make_scene(P, D, 13).to(device)
    cams = [make_camera(v, V, W, H).to(device) for v in range(V)]
    bg = torch.tensor([0.1, 0.2, 0.3], device=device)
    sts = _C.forward_geom_views(bg, sc.means3D, sc.opacities, sc.scales, sc.rotations, 1.0, cams, H, W, sc.shs, D, False, 0.3, keep_cov3D=True)
    go = (ctypes.c_int64 * 9)(); lib.rr_geom_layout(P, go)
    view = lambda blob, off, cnt, dt: blob[off:off + cnt * torch.empty((), dtype=dt).element_size()].view(dt)
    for v in range(V):
        fw = mine_forward(sc, cams[v], bg)
        N, color, radii, depth, geom, binning, img = _C.forward_raster(sts[v])
        assert N == fw["N"] and eq(radii, fw["radii"]) and eq(color, fw["color"]) and eq(depth, fw["depth"])
        vis = fw["radii"] > 0
        assert eq(view(geom, go[0], P * 12, torch.int32).view(P, 12)[vis], fw["rec"].contiguous().view(torch.int32).view(P, 12)[vis])
        assert eq(view(geom, go[2], P, torch.uint8)[vis], fw["clamped"][vis])
        assert eq(view(geom, go[3], P, torch.int32), fw["tiles_touched"]) and eq(view(geom, go[5], P, torch.int32), fw["depth_order"])
        assert eq(view(geom, go[6], P, torch.int32), view(fw["geom"], go[6], P, torch.int32))
        assert eq(view(geom, go[1], P * 6, torch.float32).view(P, 6)[vis], fw["cov3D"][vis])
        bo = (ctypes.c_int64 * 5)(); lib.rr_binning_layout(N, W, H, bo)
        if N:
            assert eq(view(binning, bo[1], N, torch.int32), fw["point_list"])
        io = (ctypes.c_int64 * 4)(); lib.rr_image_layout(W, H, io)
        T = ((W + 15) // 16) * ((H + 15) // 16)
        assert eq(view(img, io[2], T * 2, torch.int32), fw["ranges"].reshape(-1)) and eq(view(img, io[1], H * W, torch.int32), fw["n_contrib"])


def test_backward_views_leaf_mode_applies_the_activation_chain_rule(device):
    """rr_backward_views with RR_BWD_LEAF takes the six raw leaf tensors of scene/gaussian_model.py:131-136 and writes leaf
    gradients: against the activated-mode result pushed through torch's own backward of exp / sigmoid / F.normalize / cat
    (scene/gaussian_model.py:85-105)."""
    from rain_b200 import multiview as mv
    P, W, H, D, V = 30001, 160, 96, 3, 4
    sc = make_scene(P, D, 8)
    gen = torch.Generator().manual_seed(3)
    leaves = dict(xyz=sc.means3D.clone(), f_dc=sc.shs[:, 0:1, :].contiguous(), f_rest=sc.shs[:, 1:, :].contiguous(),
                  opacity=torch.log(sc.opacities.clamp(1e-4, 1 - 1e-4) / (1 - sc.opacities.clamp(1e-4, 1 - 1e-4))), scaling=torch.log(sc.scales),
                  rotation=sc.rotations * (0.5 + torch.rand(P, 1, generator=gen)))
    leaves = {k: v.to(device).contiguous().requires_grad_(True) for k, v in leaves.items()}
    act = dict(means3D=leaves["xyz"], shs=torch.cat((leaves["f_dc"], leaves["f_rest"]), dim=1), opacities=torch.sigmoid(leaves["opacity"]),
               scales=torch.exp(leaves["scaling"]), rotations=torch.nn.functional.normalize(leaves["rotation"]))
    det = {k: v.detach().contiguous() for k, v in act.items()}
    cams = [make_camera(v, V, W, H).to(device) for v in range(V)]
    bg = torch.zeros(3, device=device)
    e = torch.empty(0, device=device)
    group = []
    for v in range(V):
        cam = cams[v]
        N, color, radii, depth, geom, binning, img = _C.rasterize_gaussians(bg, det["means3D"], e, det["opacities"], det["scales"], det["rotations"], 1.0, e, cam.viewmatrix,
                                                                            cam.projmatrix, cam.tanfovx, cam.tanfovy, H, W, det["shs"], D, cam.campos, False, False, 0.3)
        acc = _C.backward_render(bg, make_grad(W, H, 40 + v).to(device), geom, N, binning, img, P)
        group.append(dict(viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, radii=radii, geom=geom, acc=acc))
    A = mv.GradientBucket(P, 16, device); A.flat.fill_(float("nan"))
    _C.backward_views(group, det["means3D"], det["shs"], det["scales"], det["rotations"], 1.0, 0.3, D, H, W, A.grads_out())
    L = mv.GradientBucket(P, 16, device, layout="leaf"); L.flat.fill_(float("nan"))
    raw = {k: v.detach() for k, v in leaves.items()}
    _C.backward_views(group, raw["xyz"], raw["f_dc"], raw["scaling"], raw["rotation"], 1.0, 0.3, D, H, W, L.grads_out(), leaf=True, f_rest=raw["f_rest"])
    torch.autograd.backward([act["means3D"], act["shs"], act["opacities"], act["scales"], act["rotations"]],
                            [A.views["means3D"], A.views["sh"], A.views["opacity"], A.views["scales"], A.views["rotations"]])
    for k in mv.LEAF_SEGMENTS:
        assert torch.isfinite(L.views[k]).all(), k
        # leaf mode re-activates the leaves itself (splat_math.cuh: expf / sigmoid / normalise, an ulp or two from torch's) and
        # recomputes Sigma from them, so the cancellation-heavy scale / rotation sums move by a few 1e-5 of max
        assert rel_err(L.views[k], leaves[k].grad.reshape(L.views[k].shape))[0] < 5e-5, k


def test_leaf_fast_path_forward_and_step(device):
    """The opt-in leaf mode end to end: rr_forward_geom_views with RR_GEOM_LEAF on the raw leaves gives,
    bit for bit, what the activated path gives on rr_activate's outputs (one set of activation device functions); rr_activate
    agrees with torch's sigmoid / exp / F.normalize to a few ulp; and a whole multi-view step on the leaves fills a leaf-layout
    bucket equal to torch's chain rule applied to the activated-path bucket."""
    from rain_b200 import multiview as mv
    P, W, H, D, V = 30001, 160, 96, 3, 4
    sc = make_scene(P, D, 8)
    gen = torch.Generator().manual_seed(3)
    op = sc.opacities.clamp(1e-4, 1 - 1e-4)
    leaves = dict(xyz=sc.means3D.clone(), f_dc=sc.shs[:, 0:1, :].contiguous(), f_rest=sc.shs[:, 1:, :].contiguous(), opacity=torch.log(op / (1 - op)),
                  scaling=torch.log(sc.scales), rotation=sc.rotations * (0.5 + torch.rand(P, 1, generator=gen)))
    leaves = {k: v.to(device).contiguous() for k, v in leaves.items()}
    o_a, s_a, r_a = _C.activate(leaves["opacity"], leaves["scaling"], leaves["rotation"])
    assert rel_err(o_a, torch.sigmoid(leaves["opacity"]))[0] < 1e-6 and rel_err(s_a, torch.exp(leaves["scaling"]))[0] < 1e-6
    assert rel_err(r_a, torch.nn.functional.normalize(leaves["rotation"]))[0] < 1e-6
    shs = torch.cat((leaves["f_dc"], leaves["f_rest"]), dim=1).contiguous()
    cams = [make_camera(v, V, W, H).to(device) for v in range(V)]
    bg = torch.zeros(3, device=device)
    A = _C.forward_geom_views(bg, leaves["xyz"], o_a, s_a, r_a, 1.0, cams, H, W, shs, D, False, 0.3)
    L = _C.forward_geom_views(bg, leaves["xyz"], leaves["opacity"], leaves["scaling"], leaves["rotation"], 1.0, cams, H, W, leaves["f_dc"], D, False, 0.3,
                              leaf=True, f_rest=leaves["f_rest"])
    for a, l in zip(A, L):
        Na, ca, ra, da, *_ = _C.forward_raster(a)
        Nl, cl, rl, dl, *_ = _C.forward_raster(l)
        assert Na == Nl and eq(ra, rl) and eq(ca, cl) and eq(da, dl)
    # a whole step on the leaves against the activated path + torch's chain rule
    dp = [make_grad(W, H, 60 + v).to(device) for v in range(V)]
    bl = mv.GradientBucket(P, 16, device, layout="leaf"); bl.flat.fill_(float("nan"))
    res_l = mv.train_step_views(None, cams, V, bl, lambda v, c, d: dp[v], bg=bg, sh_degree=D, leaves=leaves)
    req = {k: v.clone().requires_grad_(True) for k, v in leaves.items()}
    act = dict(means3D=req["xyz"], shs=torch.cat((req["f_dc"], req["f_rest"]), dim=1), opacities=torch.sigmoid(req["opacity"]),
               scales=torch.exp(req["scaling"]), rotations=torch.nn.functional.normalize(req["rotation"]))
    ba = mv.GradientBucket(P, 16, device)
    res_a = mv.train_step_views({k: v.detach().contiguous() for k, v in act.items()}, cams, V, ba, lambda v, c, d: dp[v], bg=bg, sh_degree=D)
    torch.autograd.backward([act["means3D"], act["shs"], act["opacities"], act["scales"], act["rotations"]],
                            [ba.views["means3D"], ba.views["sh"], ba.views["opacity"], ba.views["scales"], ba.views["rotations"]])
    for a, l in zip(res_a, res_l):
        assert rel_err(l.color, a.color)[0] < 1e-5 and (l.radii != a.radii).float().mean().item() < 1e-4      # torch's activations are an ulp or two away
    for k in mv.LEAF_SEGMENTS:
        assert torch.isfinite(bl.views[k]).all(), k
        assert rel_err(bl.views[k], req[k].grad.reshape(bl.views[k].shape))[0] < 1e-4, k


def _sort_case(device, n, key_bytes, b0, b1, seed, skew):
    lib = _lib.load()
    g = torch.Generator().manual_seed(seed)
    if key_bytes == 4:
        keys = torch.randint(0, 2 ** 31 - 1, (n,), generator=g, dtype=torch.int64)
        keys = keys % 7 if skew else keys
        kd = keys.to(torch.int32).to(device)
    else:
        hi = torch.randint(0, 8160, (n,), generator=g, dtype=torch.int64)
        lo = torch.randint(0, 2 ** 31 - 1, (n,), generator=g, dtype=torch.int64)
        keys = (hi << 32) | (lo % 3 if skew else lo)
        kd = keys.to(device)
    mask = ((1 << b1) - 1) ^ ((1 << b0) - 1)
    perm = torch.sort(keys & mask, stable=True).indices
    ws = torch.empty(lib.rr_sort_ws_bytes(n, key_bytes, b0, b1), dtype=torch.uint8, device=device)
    kout = torch.empty_like(kd); vin = torch.arange(n, dtype=torch.int32, device=device); vout = torch.empty_like(vin)
    fn = lib.rr_sort_pairs_u32 if key_bytes == 4 else lib.rr_sort_pairs_u64
    _lib.check(fn(torch.cuda.current_stream().cuda_stream, n, kd.data_ptr(), kout.data_ptr(), vin.data_ptr(), vout.data_ptr(), b0, b1, ws.data_ptr()), "sort")
    torch.cuda.synchronize()
    assert eq(vout.long(), perm) and eq(kout.long(), keys[perm])


@pytest.mark.parametrize("n,key_bytes,b0,b1,skew", [
    (1, 4, 0, 32, False), (2, 8, 0, 45, False), (1000, 4, 0, 32, False), (4096, 4, 0, 13, False), (4097, 4, 0, 13, True), (100_003, 4, 0, 32, False),
    (3073, 8, 0, 45, False), (3071, 8, 0, 41, True), (1_000_000, 4, 0, 30, False), (8_400_000, 4, 0, 13, False), (2_000_000, 8, 0, 45, False),
    (2_000_000, 8, 0, 45, True), (50_000, 4, 8, 24, False)])
def test_onesweep_sort_is_a_stable_sort(device, n, key_bytes, b0, b1, skew):
    """the hand-written onesweep (the replacement for cub::DeviceRadixSort::SortPairs, rasterizer_impl.cu:295-300 and
    simple_knn.cu:199-202) against torch's stable sort — bit-exact keys and permutation"""
    _sort_case(device, n, key_bytes, b0, b1, 1234 + n, skew)


@pytest.mark.parametrize("P", [4, 7, 1000, 100_000])
def test_dist2_exact_against_oracle_and_reference(device, oracle, refcuda, P):
    g = torch.Generator().manual_seed(P)
    pts = (torch.rand(P, 3, generator=g) * 2 - 1).float()
    if P >= 1000:
        pts[10] = pts[11]; pts[12] = pts[11]                               # coincident points count with distance 0
        pts[: P // 4] = torch.randn(P // 4, 3, generator=g) * 0.01 + 0.3   # a dense cluster
    got = _C.distCUDA2(pts.to(device)).cpu()
    assert np.array_equal(got.numpy(), oracle.dist2(pts.numpy(), brute=P <= 1000))
    assert eq(got, refcuda.dist2(pts.to(device)))


def test_dist2_five_million_points_sampled_queries(device, oracle):
    """BASELINE.json config 5: 5M uniform points; exactness checked on 10 000 sampled queries."""
    P = CONFIGS["C5"]["P"]
    pts = (torch.rand(P, 3, generator=torch.Generator().manual_seed(0)) * 2 - 1).float()
    got = _C.distCUDA2(pts.to(device)).cpu().numpy()
    idx = np.random.default_rng(0).choice(P, 10_000, replace=False).astype(np.int32)
    assert np.array_equal(got[idx], oracle.dist2_queries(pts.numpy(), idx))
    assert np.isfinite(got).all() and (got > 0).all()


def test_golden_knn_vectors(device):
    files = sorted(glob.glob(os.path.join(GOLDEN, "knn_*.npz")))
    if not files:
        pytest.skip("no golden vectors committed yet")
    for f in files:
        z = np.load(f)
        assert np.array_equal(_C.distCUDA2(torch.from_numpy(z["points"]).to(device)).cpu().numpy(), z["dist2"]), f


def test_packed_exp_is_bit_identical_to_libdevice_expf(device):
    """exp2x (csrc/render_common.cuh) restates the instruction sequence nvcc emits for the reference's expf
    (forward.cu:336, backward.cu:488) with its packable steps packed; the forward's bit-exact n_contrib / final_T /
    colour parity needs it to agree with libdevice expf in every bit for every x <= 0: checked exhaustively."""
    import ctypes
    lib = _lib.load()
    bad = torch.zeros(1, dtype=torch.int64, device=device)
    stream = torch.cuda.current_stream().cuda_stream
    # +0, then every negative float from -0 through -inf (0x80000000 .. 0xFF800000)
    _lib.check(lib.rr_selftest_exp(stream, 0, 1, bad.data_ptr()), "selftest")
    _lib.check(lib.rr_selftest_exp(stream, 0x80000000, 0x7F800001, bad.data_ptr()), "selftest")
    torch.cuda.synchronize()
    assert int(bad.item()) == 0


@pytest.mark.skipif(TILE_CULL or XC_INNER, reason="this is the inner run")
def test_crosscheck_build_with_tma_staging():
    """The cross-check build of the library (make crosscheck: librainrast_xc.so, -DRR_CROSSCHECK) stages the tile renderers with
    one 48-byte TMA bulk copy per record (cp.async.bulk onto an expect_tx mbarrier) instead of cp.async: the reference-parity
    tests again in a process that loads that build (conftest.py: RAIN_TEST_LIB) with RAIN_XC_STAGING=tma."""
    import subprocess, sys
    xc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rain_b200", "librainrast_xc.so")
    if not os.path.exists(xc):
        pytest.skip("librainrast_xc.so not built (make -C rain_b200/csrc crosscheck)")
    env = dict(os.environ, RAIN_TEST_LIB=xc, RAIN_XC_STAGING="tma")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "(configs_match_reference and (C1 or C2)) or shapes_and_settings or ring_handshake or precomputed or golden_vectors"],
                       env=env, capture_output=True, text=True, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]


@pytest.mark.skipif(TILE_CULL or XC_INNER, reason="this is the inner run")
def test_tile_cull_mode_keeps_images_bit_identical():
    """The opt-in tile culling (RAIN_TILE_CULL=1 is read once per process): the reference-parity tests of this file again in a
    process with it on — shorter lists, bit-identical colour / depth / final_T, gradients within tolerance."""
    import subprocess, sys
    env = dict(os.environ, RAIN_TILE_CULL="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "(configs_match_reference and not C4) or shapes_and_settings or precomputed or golden_vectors or (accumulate_mode and not 1000000) or full_size or autograd"],
                       env=env, capture_output=True, text=True, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
