"""View-sharded data parallelism for the rasterizer — the one component with no reference
counterpart: sharonal10/rain is single-process, single-GPU (utils/general_utils.py:90).

The path shards across camera views: every rank (one process per GPU) holds a full replica of the Gaussian parameters
and renders its slice of the view batch.  The backward-preprocess kernel ADDS each view's parameter gradients straight
into one flat per-rank bucket (RR_BWD_ACCUMULATE, no zero-fill pass, Gaussians a view does not see are not touched),
and ONE all-reduce (NCCL over NVLink/NVSwitch, sum, f32, 59 floats per Gaussian with 16 SH coefficients) combines the
ranks per step.  No other exchange is needed; Gaussians and tiles are not partitioned.

Bucket layout (floats, contiguous, so that each segment is directly usable as the .grad of the matching tensor):
    means3D [P,3] | sh [P,M,3] | opacity [P,1] | scales [P,3] | rotations [P,4]
"""
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import torch

from . import _C as _native

SEGMENTS = ("means3D", "sh", "opacity", "scales", "rotations")
# the leaf layout: gradients w.r.t. the six raw leaf tensors, in the order of scene/gaussian_model.py:131-136
LEAF_SEGMENTS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def shard_views(num_views: int, rank: int, world_size: int) -> List[int]:
    """Contiguous block partition of view indices 0..num_views-1; block sizes differ by at most one and every view
    belongs to exactly one rank."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    base, extra = divmod(num_views, world_size)
    start = rank * base + min(rank, extra)
    return list(range(start, start + base + (1 if rank < extra else 0)))


class GradientBucket:
    """One flat f32 buffer holding the per-Gaussian parameter gradients of a rank, plus typed views of its segments."""

    def __init__(self, P: int, M: int, device, layout: str = "activated", symmetric: bool = False, group=None):
        """symmetric=True (collective: every rank of ``group`` constructs it): the buffer lives in symmetric memory and
        ``allreduce_bucket`` runs the library's one-kernel NVLink all-reduce (rain_b200.collective) instead of ncclAllReduce."""
        self.P, self.M, self.layout = int(P), int(M), layout
        self.floats_per_gaussian = 3 + 3 * self.M + 1 + 3 + 4
        self.collective = None
        if symmetric:
            from .collective import SymmetricAllReduce
            self.collective = SymmetricAllReduce(self.P * self.floats_per_gaussian, device, group)
            self.flat = self.collective.flat
            self.flat.zero_()
        else:
            self.flat = torch.zeros(self.P * self.floats_per_gaussian, dtype=torch.float32, device=device)
        if layout == "activated":      # gradients w.r.t. the activated tensors the drop-in API takes
            shapes = dict(means3D=(self.P, 3), sh=(self.P, self.M, 3), opacity=(self.P, 1), scales=(self.P, 3), rotations=(self.P, 4))
        elif layout == "leaf":         # gradients w.r.t. the raw leaves (view-batched backward in leaf mode)
            shapes = dict(xyz=(self.P, 3), f_dc=(self.P, 1, 3), f_rest=(self.P, self.M - 1, 3), opacity=(self.P, 1), scaling=(self.P, 3), rotation=(self.P, 4))
        else:
            raise ValueError(f"unknown bucket layout {layout!r}")
        self.segments = SEGMENTS if layout == "activated" else LEAF_SEGMENTS
        self.reduced = False         # the step's all-reduce already happened inside render_views_fwd_bwd (overlapped with the backward)
        self.views, off = {}, 0
        for name in self.segments:
            n = 1
            for d in shapes[name]:
                n *= d
            self.views[name] = self.flat[off:off + n].view(shapes[name])
            self.seg_offsets = getattr(self, "seg_offsets", {})
            self.seg_offsets[name] = (off, n // self.P if self.P else 0)      # (float offset of the segment, floats per Gaussian)
            off += n
        assert off == self.flat.numel()
        self.filled = False

    def grads_out(self):
        return dict(self.views)

    def zero_(self):
        self.flat.zero_()
        self.filled = False
        self.reduced = False

    def chunk_ranges(self, g0, g1):
        """(float offset, length) of the rows of Gaussians [g0, g1) in every segment of the flat buffer"""
        return [(off + g0 * per, (g1 - g0) * per) for off, per in (self.seg_offsets[k] for k in self.segments)]

    def nbytes(self):
        return self.flat.numel() * 4


@dataclass
class ViewResult:
    view: int
    color: torch.Tensor
    depth: torch.Tensor
    radii: torch.Tensor
    num_rendered: int
    dL_dmeans2D: Optional[torch.Tensor]


def render_views_fwd_bwd(params: dict, cameras: Sequence, view_ids: Sequence[int], bucket: GradientBucket,
                         grad_fn: Callable[[int, torch.Tensor, torch.Tensor], torch.Tensor], *, bg: torch.Tensor, sh_degree: int,
                         scale_modifier: float = 1.0, low_pass: float = 0.3, debug: bool = False, keep_outputs: bool = True,
                         pipeline: str = "two_stream", batch_backward: int = 8, batch_forward: int = 8, stats=None, leaves: Optional[dict] = None, overlap_allreduce: int = 0, overlap_leave_room: bool = True,
                         backend=_native) -> List[ViewResult]:
    """Forward + backward of this rank's views; parameter gradients of all views are summed into ``bucket``.

    params: dict(means3D[P,3], opacities[P,1], shs[P,M,3], scales[P,3], rotations[P,4]) — activated values, as
    gaussian_renderer/__init__.py:37-60 passes them.  cameras[v] has width/height/tanfovx/tanfovy/viewmatrix/projmatrix/
    campos (rain_b200.synthetic.Camera or the reference's scene.cameras.Camera fields).  grad_fn(view, color, depth)
    returns dL/dcolor [3,H,W] (the loss backward of train.py:113-115).  The first view overwrites the bucket (no
    zero-fill pass), later views accumulate.  pipeline: "two_stream" (default: the front of the next view on a
    high-priority side stream beside the render kernels of the current one) or "one_stream" (same work, serial order; used
    for per-stage timing, where a kernel's own duration is wanted).

    batch_backward = B > 0 (default 8): the per-Gaussian half of the backward runs ONCE per B views (rr_backward_views: parameters
    read once, bucket row written once, the views' accumulator records summed in registers) instead of once per view; per view
    only the backward tile renderer runs (rr_backward_render).  0 restores the per-view rr_backward with RR_BWD_ACCUMULATE.
    batch_forward = B > 0 (default 8): the per-Gaussian half of the forward — preprocess, depth sort, prefix sum — runs ONCE per
    B views (rr_forward_geom_views: parameters read once and projected through every camera, one segmented sort, one segmented
    scan; per-view outputs bit-identical to the per-view call's).  0 restores one rr_forward_geom per view.
    leaves (opt-in fast path next to the unchanged API): dict(xyz, f_dc[P,1,3], f_rest[P,15,3], opacity,
    scaling, rotation) — the six RAW leaf tensors of scene/gaussian_model.py:131-136.  ``params`` may then be None: the batched
    geometry pass applies sigmoid / exp / normalise / cat itself, the batched backward applies their chain rule and writes LEAF
    gradients into ``bucket`` (which must have layout="leaf"); no activated copy of the parameters exists anywhere.
    overlap_allreduce = K > 1 (N > 1, symmetric bucket; default off): the step's all-reduce runs in K chunks of Gaussians beside the
    last batched backward and ``bucket.reduced`` is set, so the ``allreduce_bucket`` call that follows is a no-op.  Measured
    SLOWER than the plain all-reduce after the backward on B200 (both kernels are HBM bound: 8.33 vs 8.08 ms per step at 2 GPUs,
    profiles/r02_n2_overlap.txt), hence off.
    stats: optional rain_b200.optim.DensificationStats; the batched backward folds the views' densification statistics into
    its per-step delta in the same pass (otherwise call stats.add_view per view).
    Host side: every view's forward blocks the host until its num_rendered has arrived, so host stalls drain the GPU.  A long-running
    loop should call ``gc.collect(); gc.freeze()`` once after set-up: a full pass of Python's cyclic collector over torch's objects
    takes tens of ms and was measured as one 8-40 ms step in a five-step-old process.
    """
    if pipeline not in ("two_stream", "one_stream"):
        raise ValueError(f"unknown pipeline {pipeline!r}")
    leaf = leaves is not None
    if leaf:
        if bucket.layout != "leaf" or batch_backward <= 0 or batch_forward <= 0 or backend is not _native or _native.tile_cull_enabled():
            raise RuntimeError("the leaf fast path needs a layout='leaf' bucket and the view-batched forward and backward (batch sizes > 0, no tile culling)")
        params = dict(means3D=leaves["xyz"], shs=leaves["f_dc"], opacities=leaves["opacity"], scales=leaves["scaling"], rotations=leaves["rotation"])
    batched = batch_backward > 0 and hasattr(backend, "backward_views") and bucket.M == 16 and bucket.layout == ("leaf" if leaf else "activated")
    fbatch = min(int(batch_forward), 16) if (batch_forward > 0 and hasattr(backend, "forward_geom_views") and bucket.M == 16) else 0
    if fbatch and backend is _native and _native.tile_cull_enabled():
        fbatch = 0               # the opt-in tile culling lives in the per-view preprocess kernel only
    dev = params["means3D"].device
    e = torch.empty(0, device=dev)
    results = []
    has_split = hasattr(backend, "forward_geom")
    side = _geom_stream(dev) if (pipeline == "two_stream" and hasattr(backend, "forward_bin")) else None
    view_ids = list(view_ids)
    n = len(view_ids)
    main = torch.cuda.current_stream(dev) if dev.type == "cuda" else None

    def raster_args(v):
        cam = cameras[v]
        return (bg, params["means3D"], e, params["opacities"], params["scales"], params["rotations"], scale_modifier, e,
                cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.height, cam.width, params["shs"], sh_degree, cam.campos,
                False, debug, low_pass)

    # ---- geometry half (preprocess, depth sort, prefix sum): per view, or per chunk of fbatch views -------------------------
    states = {}                  # view position -> forward state whose geometry half is enqueued

    def ensure_geom(pos, stream=None):
        if pos >= n or pos in states:
            return
        if side is not None:
            stream = side if stream is None else stream
            stream.wait_stream(main)             # buffers come from the compute stream's pool: order after their previous users
        if fbatch:
            lo = (pos // fbatch) * fbatch
            hi = min(n, lo + fbatch)
            sts = backend.forward_geom_views(bg, params["means3D"], params["opacities"], params["scales"], params["rotations"], scale_modifier,
                                             [cameras[view_ids[j]] for j in range(lo, hi)], cameras[view_ids[lo]].height, cameras[view_ids[lo]].width,
                                             params["shs"], sh_degree, debug, low_pass, stream=stream, keep_cov3D=not batched,
                                             **(dict(leaf=True, f_rest=leaves["f_rest"]) if leaf else {}))
            states.update(zip(range(lo, hi), sts))
        elif has_split:
            states[pos] = backend.forward_geom(*raster_args(view_ids[pos]), **({"stream": stream} if stream is not None else {}))
        else:
            states[pos] = raster_args(view_ids[pos])

    def bin_pass(pos):
        side.wait_stream(main)
        return backend.forward_bin(states[pos], stream=side)

    if side is not None and n:
        # two-stream pipeline: the HBM-bound front of the forward (preprocess, depth sort, scan, duplicate, tile sort, ranges)
        # runs on a high-priority side stream ahead of the issue-bound render / backward kernels on the compute stream, so its
        # CTAs fill the SM slots those kernels release.
        ensure_geom(0)
        if n > 1 and not fbatch:
            # step start: nothing runs beside view 0's front yet, so view 1's geometry half goes on a second side stream
            # next to it (two latency-bound fronts overlap well); enqueued before the host blocks on view 0's num_rendered
            ensure_geom(1, stream=_geom_stream(dev, 1))
        bin_pass(0)
    elif n:
        # one stream: the geometry half of view i+1 is enqueued BEFORE the backward of view i, so its num_rendered is on the
        # host long before it is needed and the device never waits for the round trip that sizes the binning buffer
        ensure_geom(0)
    # overlap the step's all-reduce with its last batched backward (N > 1, symmetric bucket, P a multiple of 4 so that every chunk's
    # rows are 16-byte aligned in every segment)
    P = params["means3D"].shape[0]
    overlap_chunks = 0
    if batched and overlap_allreduce > 1 and bucket.collective is not None and bucket.collective.world > 1 and P % 4 == 0 and P >= 4096:
        overlap_chunks = int(overlap_allreduce)
    bucket.reduced = False
    group = []
    for i, v in enumerate(view_ids):
        cam = cameras[v]
        st = states.pop(i)
        if side is not None:
            N, color, radii, depth, geom, binning, img = backend.forward_render(st)
        else:
            N, color, radii, depth, geom, binning, img = backend.forward_raster(st) if has_split else backend.rasterize_gaussians(*st)
            ensure_geom(i + 1)
        dpix = grad_fn(v, color, depth)
        if side is not None:
            # enqueued here, the side-stream passes wait for this view's render_fwd and loss and run beside its backward
            # (measured: before render_fwd 865, after it 871, after the loss 874 renders/s)
            if i + 1 < n:
                ensure_geom(i + 1)
                bin_pass(i + 1)
            ensure_geom(i + 2)
        if batched:
            # per view only the backward tile renderer; its accumulator blob, radii and geometry blob stay alive until the batch's
            # single per-Gaussian pass
            # (native backend: the accumulators live in the view's geometry blob, cleared by its forward render — no fill pass)
            acc = backend.backward_render(bg, dpix, geom, N, binning, img, P, debug, **({"in_geom": True} if backend is _native else {}))
            d2 = torch.empty((P, 3), dtype=torch.float32, device=dev)
            group.append(dict(viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, campos=cam.campos, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                              radii=radii, geom=geom, acc=acc, dL_dmean2D=d2))
            if len(group) == batch_backward or i + 1 == n:
                bv = lambda **kx: backend.backward_views(group, params["means3D"], params["shs"], params["scales"], params["rotations"], scale_modifier, low_pass,
                                                         sh_degree, cam.height, cam.width, bucket.grads_out(), accumulate=bucket.filled,
                                                         stats=stats.delta_tensors() if stats is not None else None, debug=debug,
                                                         **(dict(leaf=True, f_rest=leaves["f_rest"]) if leaf else {}), **kx)
                if i + 1 == n and overlap_chunks > 1:
                    # last pass of the step on N > 1 GPUs: the Gaussians are walked in chunks, and while chunk k + 1 is computed the
                    # finished bucket rows of chunk k are all-reduced by the one-kernel NVLink all-reduce on a second stream (the
                    # backward is capped at 3 CTAs per SM so that one all-reduce CTA fits on every SM beside it)
                    comm = _geom_stream(dev, 2)
                    step_g = ((P + overlap_chunks - 1) // overlap_chunks + 3) // 4 * 4
                    for g0 in range(0, P, step_g):
                        g1 = min(P, g0 + step_g)
                        bv(first=g0, count=g1 - g0, leave_room=overlap_leave_room)
                        comm.wait_stream(main)
                        with torch.cuda.stream(comm):
                            bucket.collective.allreduce_ranges(bucket.chunk_ranges(g0, g1))
                    main.wait_stream(comm)
                    bucket.reduced = True
                else:
                    bv()
                bucket.filled = True
                group = []
        else:
            grads = backend.rasterize_gaussians_backward(
                bg, params["means3D"], radii, e, params["scales"], params["rotations"], scale_modifier, e, cam.viewmatrix, cam.projmatrix,
                cam.tanfovx, cam.tanfovy, dpix, params["shs"], sh_degree, cam.campos, geom, N, binning, img, debug, low_pass,
                grads_out=bucket.grads_out(), accumulate=bucket.filled,       # the first view overwrites the bucket, later views add
                **({"omit": ("colors", "cov3D")} if backend is _native else {}))   # intermediates nobody reads (no zero fill, no read-modify-write)
            bucket.filled = True
            d2 = grads[0]
            if stats is not None:
                stats.add_view(radii, d2)
        results.append(ViewResult(v, color if keep_outputs else None, depth if keep_outputs else None, radii if keep_outputs else None,
                                  int(N), d2 if keep_outputs else None))
    return results


_geom_streams = {}


def _geom_stream(dev, which=0):
    """The high-priority side streams of the two-stream view pipeline (render_views_fwd_bwd)."""
    key = (torch.device(dev).index, which)
    if key not in _geom_streams:
        _geom_streams[key] = torch.cuda.Stream(device=dev, priority=-1)
    return _geom_streams[key]


class HostFeed:
    """Per-view host inputs (target images in pinned memory) streamed to the device on a side stream so that the copy of
    view i+1 overlaps the render of view i.  ``stage(step_tensors)`` enqueues the copies of one step after everything
    already queued on the compute stream (the previous step may still read the staging buffers) and returns, per slot,
    the device buffer; ``wait(slot)`` makes the compute stream wait for that slot's copy only."""

    def __init__(self, shapes, device, dtype=torch.float32):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self.buffers = [torch.empty(s, dtype=dtype, device=device) for s in shapes]
        self.events = [torch.cuda.Event() for _ in shapes]

    def stage(self, host_tensors):
        main = torch.cuda.current_stream(self.device)
        self.stream.wait_stream(main)
        with torch.cuda.stream(self.stream):
            for buf, ev, h in zip(self.buffers, self.events, host_tensors):
                buf.copy_(h, non_blocking=True)
                ev.record(self.stream)
        return self.buffers

    def stage_slot(self, slot, host_tensor):
        """Enqueue the copy of ONE slot (after what is queued on the compute stream so far).  Staging view i+1 when view i
        starts keeps a single short transfer in flight instead of a burst of all the step's images."""
        main = torch.cuda.current_stream(self.device)
        self.stream.wait_stream(main)
        with torch.cuda.stream(self.stream):
            self.buffers[slot].copy_(host_tensor, non_blocking=True)
            self.events[slot].record(self.stream)
        return self.buffers[slot]

    def wait(self, slot):
        torch.cuda.current_stream(self.device).wait_event(self.events[slot])
        return self.buffers[slot]


def allreduce_bucket(bucket: GradientBucket, group=None, average: bool = False):
    """The single collective of a step: sum (or mean) of the flat bucket over all ranks, in place."""
    import torch.distributed as dist
    if group == "skip":          # a single-rank step inside an initialised process group (tools/verify_multigpu.py)
        return bucket
    if bucket.reduced:           # already summed, chunk by chunk beside the step's last backward pass
        bucket.reduced = False
        if average:
            bucket.flat.div_(dist.get_world_size(group))
        return bucket
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        if bucket.collective is not None:
            bucket.collective.allreduce()        # one kernel over NVLink peer / multicast memory (csrc/allreduce.cu)
        else:
            dist.all_reduce(bucket.flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            bucket.flat.div_(dist.get_world_size(group))
    return bucket


def train_step_views(params, cameras, num_views, bucket, grad_fn, *, rank=0, world_size=1, group=None, **kw):
    """One data-parallel step: zero-copy accumulate the local views' gradients, then one all-reduce."""
    bucket.filled = False
    local = shard_views(num_views, rank, world_size)
    res = render_views_fwd_bwd(params, cameras, local, bucket, grad_fn, **kw)
    if not local:
        bucket.zero_()
    allreduce_bucket(bucket, group)
    return res
