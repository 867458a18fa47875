"""Design study (CPU, uses the oracle as a data source — tooling, not product): for a workload, how many
(pixel-block, list entry) visits does the backward tile walk need for different pixel-block shapes, and how many lanes
of a visit contribute?  Drives the choice of the sub-warp group footprint in render_{fwd,bwd}.cu.

    python tools/visit_stats.py [C3] [tile_stride]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.oracle_py import Oracle  # noqa: E402
from rain_b200.synthetic import CONFIGS, make_camera, make_scene  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
stride = int(sys.argv[2]) if len(sys.argv) > 2 else 16
cfg = CONFIGS[wl]
P, W, H, D = cfg["P"], cfg["W"], cfg["H"], cfg["D"]
sc, cam = make_scene(P, D, seed=0), make_camera(5, 64, W, H)
orc = Oracle()
N, color, radii, depth = orc.forward(np.zeros(3, np.float32), sc.means3D.numpy(), sc.opacities.numpy(), sc.scales.numpy(), sc.rotations.numpy(),
                                     cam.viewmatrix.numpy(), cam.projmatrix.numpy(), cam.campos.numpy(), cam.tanfovx, cam.tanfovy, H, W,
                                     shs=sc.shs.numpy(), sh_degree=D)
m2 = orc.get("means2D").reshape(-1, 2)
co = orc.get("conic_opacity").reshape(-1, 4)
pl = orc.get("point_list")
rg = orc.get("ranges").reshape(-1, 2)
nc = orc.get("n_contrib").reshape(H, W)
gx, gy = (W + 15) // 16, (H + 15) // 16
print(f"{wl}: P={P} N={N} tiles={gx * gy} visible={(radii > 0).sum()}")

shapes = {"16x16": (16, 16), "8x4(warp)": (4, 8), "4x4": (4, 4), "4x2": (2, 4), "2x4": (4, 2), "2x2": (2, 2), "8x1": (1, 8), "1x1": (1, 1)}   # (rows, cols)
visits = {k: 0 for k in shapes}
steps_lock = {"4x4": 0, "4x2": 0, "2x4": 0, "2x2": 0}      # lock-step steps of a warp whose groups walk their own lists per 64-entry batch
pairs_total = pairs_any = contrib_total = 0
BATCH = 64
ys, xs = np.mgrid[0:16, 0:16]
for t in range(0, gx * gy, stride):
    ty, tx = divmod(t, gx)
    lo, hi = rg[t]
    n = int(hi - lo)
    if n == 0:
        continue
    ids = pl[lo:hi]
    px = (tx * 16 + xs).astype(np.float32)[None]
    py = (ty * 16 + ys).astype(np.float32)[None]
    inside = ((tx * 16 + xs) < W) & ((ty * 16 + ys) < H)
    ncl = np.zeros((16, 16), np.int64)
    yy, xx = np.minimum(ty * 16 + ys, H - 1), np.minimum(tx * 16 + xs, W - 1)
    ncl[:] = nc[yy, xx]
    ncl[~inside] = 0
    dx = m2[ids, 0][:, None, None] - px
    dy = m2[ids, 1][:, None, None] - py
    A, B, Cc, op = (co[ids, k][:, None, None] for k in range(4))
    power = -0.5 * (A * dx * dx + Cc * dy * dy) - B * dx * dy
    alpha = np.minimum(0.99, op * np.exp(np.minimum(power, 0)))
    contrib = (power <= 0) & (alpha >= 1.0 / 255.0) & (np.arange(n)[:, None, None] < ncl[None])
    pairs_total += n
    pairs_any += int(contrib.any(axis=(1, 2)).sum())
    contrib_total += int(contrib.sum())
    for k, (r, c) in shapes.items():
        blk = contrib.reshape(n, 16 // r, r, 16 // c, c).any(axis=(2, 4))       # [n, by, bx]
        visits[k] += int(blk.sum())
        if k in steps_lock:
            # groups of one warp = the blocks inside each 4-row x 8-col warp footprint
            g = blk.reshape(n, 4, 4 // r, 2, 8 // c).transpose(0, 1, 3, 2, 4).reshape(n, 8, -1)     # [n, warp, group]
            nb = (n + BATCH - 1) // BATCH
            pad = nb * BATCH - n
            gp = np.concatenate([g, np.zeros((pad,) + g.shape[1:], bool)]) if pad else g
            per_batch = gp.reshape(nb, BATCH, 8, -1).sum(axis=1)                  # [nb, warp, group] visits
            steps_lock[k] += int(per_batch.max(axis=2).sum())
print(f"sampled tiles 1/{stride}: list entries {pairs_total}, with >=1 contributing pixel {pairs_any} ({100 * pairs_any / pairs_total:.1f}%), contributing (pixel,entry) {contrib_total}")
base = visits["8x4(warp)"]
for k, v in visits.items():
    r, c = shapes[k]
    lanes = r * c
    print(f"  block {k:10s}: visits {v:10d}  ({v / base:5.2f}x warp visits)  lane utilisation {contrib_total / (v * lanes):.3f}  warp-steps if perfectly packed {v * lanes / 32 / base:5.2f}x")
for k, v in steps_lock.items():
    print(f"  lock-step warp steps with {k} groups (own lists, batch {BATCH}): {v / base:5.2f}x of today's warp visits")
