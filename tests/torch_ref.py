"""TEST INFRASTRUCTURE — an independent float64 PyTorch re-derivation of the rasterizer's mathematics, differentiated by
torch.autograd.  It pins the CPU oracle (oracle/rast_oracle.cpp) from a second direction: the oracle's backward is a
restatement of the reference's hand-derived backward (backward.cu), this file never looks at it.

Dense formulation: every pixel sees every Gaussian in global (depth, index) order, masked by the Gaussian's tile
rectangle (the per-tile lists of the reference are exactly the sub-sequences this mask selects,
rasterizer_impl.cu:59-127).  Only usable for a few hundred Gaussians on a small image.

Two places where the reference's backward deliberately differs from the true derivative are reproduced with
detach() so that autograd yields the reference's definition:
  * the 0.99 alpha clamp is not masked in the backward (backward.cu:488-544): straight-through;
  * the clamped view-space x/y of computeCov2D is treated as an independent variable (backward.cu:165-166,252-258).
"""
import math

import torch

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
         1.445305721320277, -0.5900435899266435]


def sh_basis(d, degree):
    """[P,16] basis values (auxiliary.h:11-28, forward.cu:9-60); entries above the active degree are zero."""
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    one = torch.ones_like(x)
    zero = torch.zeros_like(x)
    b = [SH_C0 * one]
    if degree > 0:
        b += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if degree > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2 * zz - xx - yy), SH_C2[3] * xz, SH_C2[4] * (xx - yy)]
    if degree > 2:
        b += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
              SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy), SH_C3[5] * z * (xx - yy),
              SH_C3[6] * x * (xx - 3 * yy)]
    b += [zero] * (16 - len(b))
    return torch.stack(b, 1)


def quat_to_rot(q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([
        torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)], 1),
        torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)], 1),
        torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1)], 1)


def render(means3D, means2D, opacities, shs, scales, rotations, *, viewmatrix, projmatrix, campos, tanfovx, tanfovy, W, H, bg,
           sh_degree, scale_modifier=1.0, low_pass=0.3, colors_precomp=None, cov3D_precomp=None):
    """Returns dict(color[3,H,W], depth[1,H,W], radii[P], tiles_touched[P], final_T[H*W], n_contrib[H*W], order[P]).
    All float inputs are float64 tensors; gradients flow to means3D, means2D (the reference's NDC-scaled screen-space
    gradient receptacle), opacities, shs/colors_precomp, scales, rotations/cov3D_precomp."""
    f64 = torch.float64
    V, Pm = viewmatrix.to(f64), projmatrix.to(f64)
    P = means3D.shape[0]
    fx, fy = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    hom = torch.cat([means3D, torch.ones(P, 1, dtype=f64)], 1)
    t = hom @ V[:, :3]                                   # view space (transformPoint4x3, auxiliary.h:47-55)
    ph = hom @ Pm
    p_w = 1.0 / (ph[:, 3] + 0.0000001)                   # forward.cu:189
    ndc = ph[:, :2] * p_w[:, None]
    in_front = t[:, 2] > 0.2                             # auxiliary.h:137-143

    if cov3D_precomp is None:
        R = quat_to_rot(rotations)
        S2 = (scale_modifier * scales) ** 2
        Sigma = R @ torch.diag_embed(S2) @ R.transpose(1, 2)
    else:
        c = cov3D_precomp
        Sigma = torch.stack([torch.stack([c[:, 0], c[:, 1], c[:, 2]], 1), torch.stack([c[:, 1], c[:, 3], c[:, 4]], 1),
                             torch.stack([c[:, 2], c[:, 4], c[:, 5]], 1)], 1)

    tz = t[:, 2]
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    txtz, tytz = t[:, 0] / tz, t[:, 1] / tz
    cl_x, cl_y = (txtz < -limx) | (txtz > limx), (tytz < -limy) | (tytz > limy)
    tx = torch.where(cl_x, (txtz.clamp(-limx, limx) * tz).detach(), t[:, 0])
    ty = torch.where(cl_y, (tytz.clamp(-limy, limy) * tz).detach(), t[:, 1])
    zero = torch.zeros_like(tz)
    J = torch.stack([torch.stack([fx / tz, zero, -(fx * tx) / (tz * tz)], 1),
                     torch.stack([zero, fy / tz, -(fy * ty) / (tz * tz)], 1)], 1)          # [P,2,3]
    Wm = V[:3, :3].t()                                   # standard world->view rotation
    Tm = J @ Wm                                          # [P,2,3]
    cov = Tm @ Sigma @ Tm.transpose(1, 2)
    a = cov[:, 0, 0] + low_pass
    b = cov[:, 0, 1]
    cc = cov[:, 1, 1] + low_pass
    det = a * cc - b * b
    det_ok = det != 0
    det_s = torch.where(det_ok, det, torch.ones_like(det))
    conic = torch.stack([cc / det_s, -b / det_s, a / det_s], 1)
    mid = 0.5 * (a + cc)
    lam = mid + torch.sqrt(torch.clamp_min(mid * mid - det, 0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam)).detach()
    pix = torch.stack([((ndc[:, 0] + means2D[:, 0] + 1.0) * W - 1.0) * 0.5, ((ndc[:, 1] + means2D[:, 1] + 1.0) * H - 1.0) * 0.5], 1)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    pd = pix.detach()

    def tile(v, g):
        return torch.clamp(torch.trunc(v / 16.0), 0, g).long()
    x0, y0 = tile(pd[:, 0] - radius, gx), tile(pd[:, 1] - radius, gy)
    x1, y1 = tile(pd[:, 0] + radius + 15.0, gx), tile(pd[:, 1] + radius + 15.0, gy)
    tiles = (x1 - x0) * (y1 - y0)
    visible = in_front & det_ok & (tiles > 0)
    radii = torch.where(visible, radius.long(), torch.zeros_like(tiles))
    tiles = torch.where(visible, tiles, torch.zeros_like(tiles))

    if colors_precomp is None:
        d = means3D - campos.to(f64)[None]
        d = d / d.norm(dim=1, keepdim=True)
        rgb = torch.einsum("pk,pkc->pc", sh_basis(d, sh_degree), shs[:, :16]) + 0.5
        rgb = torch.clamp_min(rgb, 0.0)
    else:
        rgb = colors_precomp

    # global (depth, index) order — the stable sort of rasterizer_impl.cu:295-300 restricted to any tile
    depth32 = t[:, 2].detach().float()
    order = torch.argsort(depth32.double(), stable=True)
    order = order[visible[order]]
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    px, py = xs.reshape(-1).to(f64), ys.reshape(-1).to(f64)
    ptx, pty = (xs.reshape(-1) // 16), (ys.reshape(-1) // 16)
    o = order
    in_rect = (ptx[None] >= x0[o, None]) & (ptx[None] < x1[o, None]) & (pty[None] >= y0[o, None]) & (pty[None] < y1[o, None])   # [V,X]
    dx = pix[o, 0:1] - px[None]
    dy = pix[o, 1:2] - py[None]
    power = -0.5 * (conic[o, 0:1] * dx * dx + conic[o, 2:3] * dy * dy) - conic[o, 1:2] * dx * dy
    G = torch.exp(torch.clamp_max(power, 0.0))
    alpha_raw = opacities[o].reshape(-1, 1) * G
    alpha = alpha_raw + (torch.clamp_max(alpha_raw, 0.99) - alpha_raw).detach()
    valid = in_rect & (power <= 0) & (alpha.detach() >= 1.0 / 255.0)
    a_eff = torch.where(valid, alpha, torch.zeros_like(alpha))
    T_incl = torch.cumprod(1.0 - a_eff.detach(), 0)
    incl = valid & (T_incl >= 0.0001)
    # everything after the first rejected-by-T entry is dropped (forward.cu:339-344)
    stopped = torch.cumsum((valid & ~incl).long(), 0) > 0
    incl = incl & ~stopped
    a_fin = torch.where(incl, alpha, torch.zeros_like(alpha))
    one_m = 1.0 - a_fin
    T_excl = torch.cat([torch.ones(1, one_m.shape[1], dtype=f64), torch.cumprod(one_m, 0)[:-1]], 0)
    w = a_fin * T_excl
    final_T = torch.prod(one_m, 0)
    color = (w[:, :, None] * rgb[o][:, None, :]).sum(0) + final_T[:, None] * bg.to(f64)[None]
    depth = (w * t[o, 2:3]).sum(0)
    # n_contrib = 1-based position, in the pixel's TILE list, of the last contributor
    pos_in_tile = torch.cumsum(in_rect.long(), 0)
    n_contrib = torch.where(incl, pos_in_tile, torch.zeros_like(pos_in_tile)).max(0).values if o.numel() else torch.zeros(H * W, dtype=torch.long)
    return dict(color=color.t().reshape(3, H, W), depth=depth.reshape(1, H, W), radii=radii, tiles_touched=tiles,
                final_T=final_T.detach(), n_contrib=n_contrib, order=order)


def dist2_brute(points):
    """mean of the 3 smallest squared distances to other points (simple_knn.cu:136-172), float64"""
    p = points.double()
    d = ((p[:, None, :] - p[None, :, :]) ** 2).sum(-1)
    d.fill_diagonal_(math.inf)
    return d.topk(3, dim=1, largest=False).values.mean(1)
