"""Live-lane fraction of the two tile renderers: of the 64 pixel evaluations (two per lane) a warp
issues per list entry it works on, how many pass the reference's own pair test.  Uses the work counters of the
CROSS-CHECK build of the library (rain_b200/librainrast_xc.so, -DRR_CROSSCHECK; the product library has none) on a few views
of a workload and writes profiles/live_lanes.json, which bench.py quotes in stages[] / roofline.

    python tools/live_lanes.py [C3] [views]
"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rain_b200 import _lib  # noqa: E402

_lib.LIB_PATH = os.path.join(ROOT, "rain_b200", "librainrast_xc.so")          # before anything loads the library
import torch  # noqa: E402

from rain_b200 import _C  # noqa: E402
from rain_b200.synthetic import CONFIGS, make_camera, make_grad, make_scene  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
views = int(sys.argv[2]) if len(sys.argv) > 2 else 4
cfg = CONFIGS[wl]
dev = torch.device("cuda:0")
lib = _lib.load()
lib.rr_xc_counters.argtypes = [ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]
sc = make_scene(cfg["P"], cfg["D"], 0).to(dev)
bg = torch.zeros(3, device=dev)
e = torch.empty(0, device=dev)
out = (ctypes.c_uint64 * 16)()
lib.rr_xc_counters(out, 1)
N_tot = 0
for v in range(views):
    cam = make_camera(v, 64, cfg["W"], cfg["H"]).to(dev)
    dpix = make_grad(cfg["W"], cfg["H"], 100 + v).to(dev)
    args = (bg, sc.means3D, e, sc.opacities, sc.scales, sc.rotations, 1.0, e, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.height, cam.width,
            sc.shs, cfg["D"], cam.campos, False, False, 0.3)
    N, color, radii, depth, geom, binning, img = _C.rasterize_gaussians(*args)
    _C.rasterize_gaussians_backward(bg, sc.means3D, radii, e, sc.scales, sc.rotations, 1.0, e, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy,
                                    dpix, sc.shs, cfg["D"], cam.campos, geom, N, binning, img, False, 0.3)
    N_tot += N
_lib.check(lib.rr_xc_counters(out, 1), "rr_xc_counters")
c = [int(x) for x in out]
per = lambda x: x / views
rec = {
    "render_fwd": {
        "live_lane_frac": round(c[4] / max(1, 64 * c[3]), 4), "blended_lane_frac": round(c[5] / max(1, 64 * c[3]), 4),
        "warp_entries_examined_per_view": per(c[1]), "after_aabb_cull_per_view": per(c[2]), "survivors_blended_per_view": per(c[3]),
        "live_lane_evals_per_view": per(c[4]), "blended_lane_evals_per_view": per(c[5]),
        "survivor_frac_of_examined": round(c[3] / max(1, c[1]), 4), "list_entries_per_view": N_tot / views},
    "render_bwd": {
        "live_lane_frac": round(c[9] / max(1, 64 * c[8]), 4), "visits_per_view": per(c[8]), "live_pixel_evals_per_view": per(c[9])},
    "_source": f"tools/live_lanes.py {wl} {views}: work counters of librainrast_xc.so (cross-check build) over {views} views",
}
print(json.dumps(rec, indent=1))
p = os.path.join(ROOT, "profiles", "live_lanes.json")
allj = json.load(open(p)) if os.path.exists(p) else {}
allj[wl] = rec
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(allj, open(os.path.join(ROOT, "gpurun_out", "live_lanes.json"), "w"), indent=1)
