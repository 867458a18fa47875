"""How many (tile, Gaussian) pairs of a render never pass the alpha test in any pixel of their tile?  Reads the forward's
blend-flag words (binning blob, kernels.cuh) of one C3 view.  Diagnostic only: python tools/gpu_diag_dead_pairs.py [workload]"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rain_b200 import _C, _lib, synthetic as syn

wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
cfg = syn.CONFIGS[wl]
P, W, H, D = cfg["P"], cfg["W"], cfg["H"], cfg["D"]
dev = torch.device("cuda:0")
sc = syn.make_scene(P, D, 0).to(dev)
cam = syn.make_camera(0, 64, W, H).to(dev)
bg = torch.zeros(3, device=dev)
e = torch.empty(0, device=dev)
N, color, radii, depth, geom, binning, img = _C.rasterize_gaussians(bg, sc.means3D, e, sc.opacities, sc.scales, sc.rotations, 1.0, e, cam.viewmatrix, cam.projmatrix,
                                                                     cam.tanfovx, cam.tanfovy, H, W, sc.shs, D, cam.campos, False, False, 0.3)
torch.cuda.synchronize()
lib = _lib.load()
bo = (ctypes.c_int64 * 5)(); lib.rr_binning_layout(N, W, H, bo)
io = (ctypes.c_int64 * 4)(); lib.rr_image_layout(W, H, io)
tiles = ((W + 15) // 16) * ((H + 15) // 16)
ranges = img[io[2]:io[2] + tiles * 8].view(torch.int32).view(tiles, 2).cpu().numpy().astype(np.int64)
flags = binning[bo[2]:].view(torch.int32).cpu().numpy().view(np.uint32)
ncon = img[io[1]:io[1] + H * W * 4].view(torch.int32).cpu().numpy()
alive = 0; reached = 0
gx = (W + 15) // 16
nc2 = np.zeros((((H + 15) // 16) * 16, gx * 16), dtype=np.int64); nc2[:H, :W] = ncon.reshape(H, W)
tile_reach = nc2.reshape(-1, 16, gx, 16).max(axis=(1, 3)).reshape(-1)       # list prefix the tile's forward consumed
for t in range(tiles):
    x, y = ranges[t]
    n = y - x
    if n == 0: continue
    ng = (n + 31) >> 5
    base = 8 * ((x >> 5) + t)
    w = flags[base:base + 8 * ng].reshape(8, ng)
    alive += int(np.unpackbits(np.bitwise_or.reduce(w, axis=0).view(np.uint8)).sum())
reached = int(np.minimum(tile_reach, ranges[:, 1] - ranges[:, 0]).sum())
print(f"{wl}: list prefixes up to the last contributor of each tile: {reached} ({reached / N:.3f} of N)")
print(f"{wl}: N = {N} pairs, blended into at least one pixel: {alive} ({alive / N:.3f}); opacity mean {float(sc.opacities.mean()):.3f}")
