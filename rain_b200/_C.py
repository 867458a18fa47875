"""Drop-in for the reference's pybind module ``diff_gaussian_rasterization._C`` (ext.cpp:4-8): the three ops
``rasterize_gaussians``, ``rasterize_gaussians_backward`` and ``mark_visible`` with the argument order, return order,
shapes and dtypes of rasterize_points.h:7-58, implemented on top of the C ABI of librainrast.so.

PyTorch is only plumbing here: it owns device memory (outputs, gradients, the three opaque scratch blobs) and the
current stream.  All compute happens in the hand-written sm_100a kernels.
"""
import torch

from . import _lib

NUM_CHANNELS = 3  # config.h:4

_pinned = {}


def _ptr(t):
    """device pointer or NULL for the empty tensors the reference API uses for 'not provided' (__init__.py:189-199)"""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t, dev):
    if t is None or t.numel() == 0:
        return None
    if t.device != dev:
        raise RuntimeError(f"expected a tensor on {dev}, got {t.device}")
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected float32, got {t.dtype}")
    return t.contiguous()


class _GeomPass:
    """State between the two halves of the forward (rr_forward_geom / rr_forward_raster)."""
    __slots__ = ("P", "H", "W", "dev", "bg", "debug", "out_color", "out_depth", "radii", "geom", "img", "nr", "event", "stream", "keep", "N", "binning", "bin_event", "bin_stream")


def _take_slot(dev):
    """A pinned int64 word for num_rendered, owned by ONE forward pass from forward_geom until its count has been read
    (_read_count), then returned to the per-device pool: any number of first halves may be in flight."""
    pool = _pinned.setdefault(dev.index, [])
    return pool.pop() if pool else torch.zeros(1, dtype=torch.int64).pin_memory()


def _read_count(st):
    """num_rendered of a pass whose first half has completed (the caller synchronised st.event); releases the pinned word"""
    N = int(st.nr.item())
    _pinned.setdefault(st.dev.index, []).append(st.nr)
    st.nr = None
    return N


def forward_geom(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                 viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                 prefiltered, debug, low_pass, stream=None):
    """First half of RasterizeGaussiansCUDA (preprocess, depth order, prefix sum; Rasterizer::forward up to the num_rendered
    read-back, rasterizer_impl.cu:187-273), enqueued without waiting: returns the state ``forward_raster`` completes.
    Splitting the forward lets a caller enqueue the next view's first half before the current view's backward, so the
    device never idles on the num_rendered round trip (rain_b200.multiview).  ``stream``: launch on this stream instead of the
    current one (buffers still come from the current stream's pool: the caller orders ``stream`` after the current stream
    first, and forward_raster orders its stream after this half)."""
    lib = _lib.load()
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:47-49
    if not means3D.is_cuda:
        raise RuntimeError("rain_b200 has no CPU path: means3D must be a CUDA tensor")
    dev = means3D.device
    st = _GeomPass()
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    st.P, st.H, st.W, st.dev, st.debug = P, H, W, dev, bool(debug)
    byte = dict(dtype=torch.uint8, device=dev)
    if P == 0:  # rasterize_points.cu:72
        return st
    means3D_c = _f32c(means3D, dev)
    bg, colors_c, opac, scales_c, rots, cov_c = (_f32c(t, dev) for t in (background, colors, opacity, scales, rotations, cov3D_precomp))
    view, proj, cam, sh_c = (_f32c(t, dev) for t in (viewmatrix, projmatrix, campos, sh))
    M = sh.size(1) if (sh is not None and sh.numel() != 0) else 0  # rasterize_points.cu:75-78
    st.bg = bg
    st.out_color = torch.empty((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
    st.out_depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    st.radii = torch.empty((P,), dtype=torch.int32, device=dev)
    st.geom = torch.empty(lib.rr_geom_bytes(P), **byte)
    st.img = torch.empty(lib.rr_image_bytes(W, H), **byte)
    st.nr = _take_slot(dev)
    st.keep = (means3D_c, colors_c, opac, scales_c, rots, cov_c, view, proj, cam, sh_c)   # inputs stay alive until the kernels ran
    with torch.cuda.device(dev):
        st.stream = torch.cuda.current_stream() if stream is None else stream
        _lib.check(lib.rr_forward_geom(st.stream.cuda_stream, P, int(degree), M, W, H,
                                       _ptr(means3D_c), _ptr(sh_c), _ptr(colors_c), _ptr(opac),
                                       _ptr(scales_c), float(scale_modifier), _ptr(rots), _ptr(cov_c),
                                       _ptr(view), _ptr(proj), _ptr(cam),
                                       float(tan_fovx), float(tan_fovy), float(low_pass), int(bool(prefiltered)), int(bool(debug)),
                                       st.radii.data_ptr(), st.geom.data_ptr(), st.geom.numel(), st.img.data_ptr(), st.img.numel(), st.nr.data_ptr()), "rr_forward_geom")
        st.event = torch.cuda.Event()
        st.event.record(st.stream)
    return st


def activate(opacity, scaling, rotation):
    """sigmoid / exp / F.normalize of the raw leaf tensors (scene/gaussian_model.py:85-105) with the library's own device
    functions (rr_activate): returns (opacities[P,1], scales[P,3], rotations[P,4])."""
    lib = _lib.load()
    dev = opacity.device
    P = opacity.shape[0]
    o, s, r = (_f32c(t, dev) if P else t for t in (opacity, scaling, rotation))
    out = (torch.empty((P, 1), dtype=torch.float32, device=dev), torch.empty((P, 3), dtype=torch.float32, device=dev), torch.empty((P, 4), dtype=torch.float32, device=dev))
    if P:
        with torch.cuda.device(dev):
            _lib.check(lib.rr_activate(torch.cuda.current_stream().cuda_stream, P, o.data_ptr(), s.data_ptr(), r.data_ptr(), out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr()), "rr_activate")
    return out


def forward_geom_views(background, means3D, opacity, scales, rotations, scale_modifier, cams, image_height, image_width, sh, degree,
                       debug, low_pass, stream=None, keep_cov3D=False, leaf=False, f_rest=None):
    """forward_geom for ALL views of a batch in one set of launches (rr_forward_geom_views: one pass over the Gaussians for
    every camera, one segmented depth sort, one segmented scan).  cams: list (at most 16) of objects with viewmatrix,
    projmatrix, campos, tanfovx, tanfovy.  Returns one forward-pass state per view, each to be completed with forward_bin /
    forward_render (or forward_raster) exactly like forward_geom's; their outputs are bit-identical to the per-view call's.
    SH colours + scales / rotations only (M = 16).  keep_cov3D: also store the 3D covariance per view (needed only if the
    per-view rasterize_gaussians_backward will be used on these states).  leaf=True: means3D / sh / opacity / scales / rotations
    are the RAW leaves _xyz / _features_dc / _opacity / _scaling / _rotation and f_rest is _features_rest
    (scene/gaussian_model.py:131-136); their activations are applied inside the kernel."""
    import ctypes
    lib = _lib.load()
    if not means3D.is_cuda:
        raise RuntimeError("rain_b200 has no CPU path: means3D must be a CUDA tensor")
    dev = means3D.device
    P, H, W, n = means3D.size(0), int(image_height), int(image_width), len(cams)
    byte = dict(dtype=torch.uint8, device=dev)
    means3D_c, bg, opac, scales_c, rots, sh_c = (_f32c(t, dev) for t in (means3D, background, opacity, scales, rotations, sh))
    frest_c = _f32c(f_rest, dev) if leaf else None
    if leaf and (sh.numel() != P * 3 or f_rest is None or f_rest.numel() != P * 45):
        raise RuntimeError("leaf mode needs _features_dc [P,1,3] and _features_rest [P,15,3]")
    M = 16 if leaf else (sh.size(1) if (sh is not None and sh.numel() != 0) else 0)
    states = []
    for cam in cams:
        st = _GeomPass()
        st.P, st.H, st.W, st.dev, st.debug, st.bg = P, H, W, dev, bool(debug), bg
        states.append(st)
    if P == 0 or n == 0:
        return states
    mats = [[_f32c(getattr(c, k), dev) for k in ("viewmatrix", "projmatrix", "campos")] for c in cams]
    gbytes = lib.rr_geom_bytes(P)
    for st in states:
        st.out_color = torch.empty((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
        st.out_depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
        st.radii = torch.empty((P,), dtype=torch.int32, device=dev)
        st.geom = torch.empty(gbytes, **byte)
        st.img = torch.empty(lib.rr_image_bytes(W, H), **byte)
        st.nr = _take_slot(dev)
        st.keep = (means3D_c, opac, scales_c, rots, sh_c, frest_c, mats)
    vp = ctypes.c_void_p
    arr = lambda xs: (vp * n)(*xs)
    with torch.cuda.device(dev):
        s = torch.cuda.current_stream() if stream is None else stream
        _lib.check(lib.rr_forward_geom_views(
            s.cuda_stream, P, int(degree), M, W, H, n,
            arr([m[0].data_ptr() for m in mats]), arr([m[1].data_ptr() for m in mats]), arr([m[2].data_ptr() for m in mats]),
            (ctypes.c_float * n)(*[float(c.tanfovx) for c in cams]), (ctypes.c_float * n)(*[float(c.tanfovy) for c in cams]),
            means3D_c.data_ptr(), sh_c.data_ptr(), frest_c.data_ptr() if leaf else None, opac.data_ptr(), scales_c.data_ptr(), float(scale_modifier), rots.data_ptr(), float(low_pass),
            arr([st.radii.data_ptr() for st in states]), arr([st.geom.data_ptr() for st in states]), gbytes,
            arr([st.img.data_ptr() for st in states]), states[0].img.numel(), arr([st.nr.data_ptr() for st in states]), (1 if keep_cov3D else 0) | (2 if leaf else 0), int(bool(debug))), "rr_forward_geom_views")
        ev = torch.cuda.Event()
        ev.record(s)
    for st in states:
        st.stream, st.event = s, ev
    return states


def forward_raster(st):
    """Second half of RasterizeGaussiansCUDA: waits for num_rendered (it sizes the binning blob; the reference blocks on a
    cudaMemcpy here, rasterizer_impl.cu:273), then duplicate / tile sort / ranges / render.
    Returns (num_rendered, out_color[3,H,W], radii i32[P], out_depth[1,H,W], geomBuffer, binningBuffer, imgBuffer)."""
    lib = _lib.load()
    dev, P, H, W = st.dev, st.P, st.H, st.W
    byte = dict(dtype=torch.uint8, device=dev)
    if P == 0:  # rasterize_points.cu:72
        z = torch.zeros
        return (0, z((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev), z((0,), dtype=torch.int32, device=dev),
                z((1, H, W), dtype=torch.float32, device=dev), torch.empty(0, **byte), torch.empty(0, **byte), torch.empty(0, **byte))
    st.event.synchronize()
    N = _read_count(st)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream()
        if stream != st.stream:                  # first half ran on another stream (rain_b200.multiview's geometry stream)
            stream.wait_event(st.event)
    binning = torch.empty(lib.rr_binning_bytes(N, W, H), **byte)
    with torch.cuda.device(dev):
        _lib.check(lib.rr_forward_raster(stream.cuda_stream, P, W, H, N, _ptr(st.bg),
                                         st.geom.data_ptr(), st.geom.numel(), binning.data_ptr(), binning.numel(), st.img.data_ptr(), st.img.numel(),
                                         st.out_color.data_ptr(), st.out_depth.data_ptr(), int(st.debug)), "rr_forward_raster")
    st.keep = None
    return N, st.out_color, st.radii, st.out_depth, st.geom, binning, st.img


def forward_bin(st, stream=None):
    """forward_raster without the render kernel: waits for num_rendered, then duplicate / tile sort / ranges (rr_forward_bin),
    on ``stream`` if given (same ordering contract as forward_geom's).  ``forward_render(st)`` completes the forward."""
    lib = _lib.load()
    if st.P == 0:
        return st
    st.event.synchronize()
    st.N = _read_count(st)
    st.binning = torch.empty(lib.rr_binning_bytes(st.N, st.W, st.H), dtype=torch.uint8, device=st.dev)
    with torch.cuda.device(st.dev):
        st.bin_stream = torch.cuda.current_stream() if stream is None else stream
        if st.bin_stream != st.stream:
            st.bin_stream.wait_event(st.event)
        _lib.check(lib.rr_forward_bin(st.bin_stream.cuda_stream, st.P, st.W, st.H, st.N, st.geom.data_ptr(), st.geom.numel(),
                                      st.binning.data_ptr(), st.binning.numel(), st.img.data_ptr(), st.img.numel(), int(st.debug)), "rr_forward_bin")
        st.bin_event = torch.cuda.Event()
        st.bin_event.record(st.bin_stream)
    return st


def forward_render(st):
    """The render kernel of a view whose forward_geom and forward_bin were enqueued (possibly on other streams); same return
    value as forward_raster."""
    lib = _lib.load()
    if st.P == 0:
        return forward_raster(st)
    with torch.cuda.device(st.dev):
        stream = torch.cuda.current_stream()
        if stream != st.bin_stream:
            stream.wait_event(st.bin_event)
        _lib.check(lib.rr_forward_render(stream.cuda_stream, st.P, st.W, st.H, st.N, _ptr(st.bg),
                                         st.geom.data_ptr(), st.geom.numel(), st.binning.data_ptr(), st.binning.numel(), st.img.data_ptr(), st.img.numel(),
                                         st.out_color.data_ptr(), st.out_depth.data_ptr(), int(st.debug)), "rr_forward_render")
    st.keep = None
    return st.N, st.out_color, st.radii, st.out_depth, st.geom, st.binning, st.img


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug, low_pass):
    """RasterizeGaussiansCUDA (rasterize_points.cu:24-108).
    Returns (num_rendered, out_color[3,H,W], radii i32[P], out_depth[1,H,W], geomBuffer, binningBuffer, imgBuffer)."""
    return forward_raster(forward_geom(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                                       viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                                       prefiltered, debug, low_pass))


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug, low_pass, return_conic=False,
                                 grads_out=None, accumulate=False, omit=(), acc_in_geom=False):
    """RasterizeGaussiansBackwardCUDA (rasterize_points.cu:110-191).
    Returns (dL_dmeans2D[P,3], dL_dcolors[P,3], dL_dopacity[P,1], dL_dmeans3D[P,3], dL_dcov3D[P,6], dL_dsh[P,M,3],
    dL_dscales[P,3], dL_drotations[P,4]).

    Extensions used by the view-sharded batch path (rain_b200.multiview), absent from the reference: ``grads_out`` maps
    any of "means3D", "sh", "colors", "opacity", "scales", "rotations", "cov3D" to caller-owned contiguous f32 tensors
    the kernel writes into (e.g. views of one flat allreduce bucket); ``accumulate`` adds to their contents instead
    of overwriting (RR_BWD_ACCUMULATE); ``omit`` names intermediates not to produce at all — "colors" (when ``sh`` is used)
    and/or "cov3D" (when ``cov3D_precomp`` is not) — whose entries of the returned tuple are then None.
    ``acc_in_geom``: the backward tile renderer accumulates into the region of the geometry blob that the forward cleared while
    it rendered (no fill pass); only the FIRST backward after a forward may ask for it (the autograd function does)."""
    lib = _lib.load()
    dev = means3D.device
    P, H, W = means3D.size(0), dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if (sh is not None and sh.numel() != 0) else 0
    f32 = dict(dtype=torch.float32, device=dev)
    new = torch.zeros if P == 0 else torch.empty
    dL_dmeans3D, dL_dmeans2D, dL_dcolors = new((P, 3), **f32), new((P, 3), **f32), new((P, NUM_CHANNELS), **f32)
    dL_dconic, dL_dopacity, dL_dcov3D = new((P, 2, 2), **f32), new((P, 1), **f32), new((P, 6), **f32)
    dL_dsh, dL_dscales, dL_drotations = new((P, M, 3), **f32), new((P, 3), **f32), new((P, 4), **f32)
    has_sh = sh is not None and sh.numel() != 0 and (colors is None or colors.numel() == 0)
    has_cov = cov3D_precomp is not None and cov3D_precomp.numel() != 0
    for name in omit:
        if name not in ("colors", "cov3D") or (name == "colors" and not has_sh) or (name == "cov3D" and has_cov):
            raise RuntimeError(f"cannot omit {name!r}: it is the gradient of an input of this call")
    if grads_out is not None or accumulate:
        go = grads_out or {}
        own = dict(means3D=dL_dmeans3D, sh=dL_dsh, colors=dL_dcolors, opacity=dL_dopacity, scales=dL_dscales, rotations=dL_drotations,
                   cov3D=dL_dcov3D)
        for k, t in go.items():
            if k not in own:
                raise RuntimeError(f"unknown gradient slot {k!r}")
            if t.dtype != torch.float32 or t.device != dev or not t.is_contiguous() or t.numel() != own[k].numel():
                raise RuntimeError(f"grads_out[{k!r}] must be a contiguous float32 tensor of {own[k].numel()} elements on {dev}")
            own[k] = t
        if accumulate:
            for k, t in own.items():
                if k not in go and k not in omit and t.numel():
                    t.zero_()        # slots the caller did not supply start from zero
        dL_dmeans3D, dL_dsh, dL_dcolors, dL_dopacity = own["means3D"], own["sh"], own["colors"], own["opacity"]
        dL_dscales, dL_drotations, dL_dcov3D = own["scales"], own["rotations"], own["cov3D"]
    if P != 0:
        means3D_c = _f32c(means3D, dev)
        bg, colors_c, scales_c, rots, cov_c = (_f32c(t, dev) for t in (background, colors, scales, rotations, cov3D_precomp))
        view, proj, cam, sh_c, dpix = (_f32c(t, dev) for t in (viewmatrix, projmatrix, campos, sh, dL_dout_color))
        radii_c = radii.contiguous()
        ws = None if acc_in_geom else torch.empty(lib.rr_backward_ws_bytes(P), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream()
            _lib.check(lib.rr_backward(stream.cuda_stream, P, int(degree), M, W, H, int(R), _ptr(bg),
                                       _ptr(means3D_c), _ptr(sh_c), _ptr(colors_c),
                                       _ptr(scales_c), float(scale_modifier), _ptr(rots), _ptr(cov_c),
                                       _ptr(view), _ptr(proj), _ptr(cam),
                                       float(tan_fovx), float(tan_fovy), float(low_pass), radii_c.data_ptr(),
                                       geomBuffer.data_ptr(), geomBuffer.numel(), binningBuffer.data_ptr(), binningBuffer.numel(),
                                       imageBuffer.data_ptr(), imageBuffer.numel(),
                                       dpix.data_ptr(), None if ws is None else ws.data_ptr(), 0 if ws is None else ws.numel(),
                                       dL_dmeans2D.data_ptr(), dL_dconic.data_ptr(), dL_dopacity.data_ptr(), None if "colors" in omit else dL_dcolors.data_ptr(),
                                       dL_dmeans3D.data_ptr(), None if "cov3D" in omit else dL_dcov3D.data_ptr(), _ptr(dL_dsh), dL_dscales.data_ptr(), dL_drotations.data_ptr(),
                                       int(bool(debug)), 1 if accumulate else 0), "rr_backward")
    out = (dL_dmeans2D, None if "colors" in omit else dL_dcolors, dL_dopacity, dL_dmeans3D, None if "cov3D" in omit else dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)
    return out + (dL_dconic,) if return_conic else out


def backward_render(background, dL_dout_color, geomBuffer, R, binningBuffer, imageBuffer, P, debug=False, in_geom=False):
    """First half of the backward alone (rr_backward_render: BACKWARD::render, rasterizer_impl.cu:386-404): returns the
    per-Gaussian accumulator blob of this view, which ``backward_views`` consumes together with those of the step's other
    views.  No reference counterpart as a separate op (the reference runs its whole backward per view).
    in_geom: keep the accumulators in the geometry blob (see below) instead of a fresh scratch tensor."""
    lib = _lib.load()
    dev = dL_dout_color.device
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    if in_geom:
        # the accumulator region of the view's geometry blob, which the forward cleared while it rendered: no fill pass here.
        # Valid for ONE backward_render per forward of the view (rain_b200.multiview calls it exactly once).
        off = lib.rr_geom_acc_offset(P)
        ws = geomBuffer[off:off + lib.rr_backward_ws_bytes(P) - 256]
        ws_ptr, ws_len = None, 0
    else:
        ws = torch.empty(lib.rr_backward_ws_bytes(P), dtype=torch.uint8, device=dev)
        ws_ptr, ws_len = ws.data_ptr(), ws.numel()
    if P != 0:
        bg, dpix = _f32c(background, dev), _f32c(dL_dout_color, dev)
        with torch.cuda.device(dev):
            _lib.check(lib.rr_backward_render(torch.cuda.current_stream().cuda_stream, P, W, H, int(R), _ptr(bg),
                                              geomBuffer.data_ptr(), geomBuffer.numel(), binningBuffer.data_ptr(), binningBuffer.numel(),
                                              imageBuffer.data_ptr(), imageBuffer.numel(), dpix.data_ptr(), ws_ptr, ws_len, int(bool(debug))), "rr_backward_render")
    return ws


def backward_views(views, means3D, sh, scales, rotations, scale_modifier, low_pass, degree, image_height, image_width, grads_out,
                   accumulate=False, stats=None, leaf=False, f_rest=None, debug=False, first=0, count=None, leave_room=False):
    """Second half of the backward for ALL views of a batch in one pass over the Gaussians (rr_backward_views).
    views: list of dicts with viewmatrix, projmatrix, campos (device tensors), tanfovx, tanfovy (floats), radii (i32[P]), geom (the
    view's geometry blob), acc (the blob ``backward_render`` returned) and optionally dL_dmean2D (f32[P,3], written).
    grads_out: the gradient bucket's segments — activated layout dict(means3D, sh, opacity, scales, rotations), or with
    ``leaf=True`` dict(xyz, f_dc, f_rest, opacity, scaling, rotation); then means3D / sh / scales / rotations are the raw leaves
    _xyz / _features_dc / _scaling / _rotation and ``f_rest`` is _features_rest (scene/gaussian_model.py:131-136).
    stats: optional (max_radii2D, xyz_gradient_accum, denom) f32[P] tensors updated in place.
    first / count: the range of Gaussians this call covers (default all); leave_room: cap the kernel's occupancy so that a
    concurrently running all-reduce kernel fits on every SM (rain_b200.multiview overlaps the two chunk by chunk)."""
    import ctypes
    lib = _lib.load()
    dev = means3D.device
    P, n = means3D.size(0), len(views)
    if P == 0 or n == 0:
        return
    M = 16
    if (leaf and (sh.numel() != P * 3 or f_rest is None or f_rest.numel() != P * 45)) or (not leaf and sh.numel() != P * 48):
        raise RuntimeError("backward_views needs 16 SH coefficients per Gaussian")
    keep = [_f32c(t, dev) for t in (means3D, sh, scales, rotations)] + ([_f32c(f_rest, dev)] if leaf else [])
    names = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation") if leaf else ("means3D", "sh", "opacity", "scales", "rotations")
    sizes = dict(xyz=3, means3D=3, f_dc=3, f_rest=45, sh=48, opacity=1, scaling=3, scales=3, rotation=4, rotations=4)
    for k in names:
        t = grads_out.get(k)
        if t is None or t.dtype != torch.float32 or t.device != dev or not t.is_contiguous() or t.numel() != P * sizes[k]:
            raise RuntimeError(f"grads_out[{k!r}] must be a contiguous float32 tensor of {P * sizes[k]} elements on {dev}")
    vp = ctypes.c_void_p
    arr = lambda xs: (vp * n)(*xs)
    mats = [[_f32c(v[k], dev) for k in ("viewmatrix", "projmatrix", "campos")] for v in views]
    radii = [v["radii"].contiguous() for v in views]
    d2 = [v.get("dL_dmean2D") for v in views]
    for t in d2:
        if t is not None and (t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != 3 * P):
            raise RuntimeError("dL_dmean2D must be a contiguous float32 [P,3] tensor")
    st = stats or (None, None, None)
    g = grads_out
    with torch.cuda.device(dev):
        _lib.check(lib.rr_backward_views(
            torch.cuda.current_stream().cuda_stream, P, int(degree), M, int(image_width), int(image_height), n,
            arr([m[0].data_ptr() for m in mats]), arr([m[1].data_ptr() for m in mats]), arr([m[2].data_ptr() for m in mats]),
            (ctypes.c_float * n)(*[float(v["tanfovx"]) for v in views]), (ctypes.c_float * n)(*[float(v["tanfovy"]) for v in views]),
            arr([r.data_ptr() for r in radii]), arr([v["geom"].data_ptr() for v in views]), arr([v["acc"].data_ptr() for v in views]),
            arr([t.data_ptr() if t is not None else None for t in d2]),
            keep[0].data_ptr(), keep[1].data_ptr(), keep[4].data_ptr() if leaf else None, keep[2].data_ptr(), float(scale_modifier), keep[3].data_ptr(), float(low_pass),
            g[names[0]].data_ptr(), g[names[1]].data_ptr(), g["f_rest"].data_ptr() if leaf else None,
            g["opacity"].data_ptr(), g["scaling" if leaf else "scales"].data_ptr(), g["rotation" if leaf else "rotations"].data_ptr(),
            _ptr(st[0]), _ptr(st[1]), _ptr(st[2]), int(first), int(P - first if count is None else count),
            (1 if accumulate else 0) | (2 if leaf else 0) | (4 if leave_room else 0), int(bool(debug))), "rr_backward_views")


def tile_cull_enabled():
    """True if this process runs with the opt-in exact tile culling (RAIN_TILE_CULL=1, read once by the library)"""
    return bool(_lib.load().rr_tile_cull_enabled())


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible (rasterize_points.cu:193-212): bool[P], view-space z > 0.2."""
    lib = _lib.load()
    if not means3D.is_cuda:
        raise RuntimeError("rain_b200 has no CPU path: means3D must be a CUDA tensor")
    dev = means3D.device
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    if P != 0:
        m, v, p = _f32c(means3D, dev), _f32c(viewmatrix, dev), _f32c(projmatrix, dev)
        with torch.cuda.device(dev):
            _lib.check(lib.rr_mark_visible(torch.cuda.current_stream().cuda_stream, P, m.data_ptr(), v.data_ptr(), p.data_ptr(),
                                           present.data_ptr()), "rr_mark_visible")
    return present


def distCUDA2(points):
    """simple_knn._C.distCUDA2 (simple-knn/spatial.cu:4-15): f32[P,3] cuda -> f32[P] mean squared distance to the 3 nearest
    other points."""
    lib = _lib.load()
    if not points.is_cuda:
        raise RuntimeError("rain_b200 has no CPU path: points must be a CUDA tensor")
    dev = points.device
    P = points.size(0)
    pts = _f32c(points, dev) if P else points
    means = torch.zeros((P,), dtype=torch.float32, device=dev)
    if P != 0:
        ws = torch.empty(lib.rr_dist2_ws_bytes(P), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.check(lib.rr_dist2(torch.cuda.current_stream().cuda_stream, P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(), ws.numel()), "rr_dist2")
    return means
