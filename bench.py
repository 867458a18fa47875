"""bench.py — fwd+bwd renders/sec of the splat rasterizer on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C3|C2|C1|C4]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step": every rank renders its views of the step (different cameras of the same synthetic scene) forward + backward,
accumulating the parameter gradients of its views into one flat bucket (59 floats per Gaussian), followed — for N > 1 — by
ONE NCCL all-reduce of that bucket (view-sharded data parallelism).
  * default workload C3 = BASELINE.json configs[2], the one the metric is quoted on (1M Gaussians, 1920x1080, SH degree 3):
    8 views per rank and step; scaling is WEAK (per-GPU work fixed);
  * workload C4 = BASELINE.json configs[3] (6M Gaussians, a 64-view batch): the 64 views of a step are SHARDED over the
    ranks (64 / N each), one 1.42 GB all-reduce per step; scaling is STRONG (total work fixed).
`value`  = views rendered by all ranks / max-over-ranks device time, inputs resident in HBM.
`e2e`    = the same step driven from HOST buffers through the public API: per view the camera matrices and the target
           image are copied from pinned host memory, the loss gradient is formed on the device, and the loss scalar and
           num_rendered are read back.
`dropin` = (N = 1) the metric exactly as SURVEY.md §8(d) defines it: ONE view at a time through the unchanged reference
           API — GaussianRasterizationSettings + GaussianRasterizer(...)(...) + autograd backward, as
           gaussian_renderer/__init__.py:20-64 drives it — no view pipeline, no gradient bucket.
Prints ONE JSON line on rank 0.
"""
import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from rain_b200.synthetic import CONFIGS, make_camera, make_grad, make_scene  # noqa: E402

# diagnostics of the end-to-end arm only (never set for a reported number): skip the target-image H2D feed / the per-step host sync
E2E_DIAG_NO_FEED = os.environ.get("RAIN_E2E_DIAG", "").find("nofeed") >= 0
E2E_DIAG_NO_SYNC = os.environ.get("RAIN_E2E_DIAG", "").find("nosync") >= 0
NUM_CAMERAS = 64            # the view pool (BASELINE.json config 4 uses a 64-view batch)
NUM_TARGETS = 8             # distinct target images (8-bit, pinned on the host); view slot i of a step uses image i % 8
METRIC = "fwd+bwd renders/sec (1M Gaussians, 1920x1080) at 1/2/4/8 B200"
# flops per evaluated (pixel, list entry) pair, SURVEY.md §8(d): the reference's forward spends ~15 flop on a pair its alpha
# test rejects and ~26 on one it blends (forward.cu:325-351); its backward ~65 on a blended pair (backward.cu:481-544) and
# the same ~15 on a rejected one
FLOP_FWD_REJECT, FLOP_FWD_BLEND, FLOP_BWD_REJECT, FLOP_BWD_BLEND = 15, 26, 15, 65


def views_per_rank(workload, world):
    """views each rank renders per step: 8 (weak scaling) except for C4, whose 64-view batch is sharded over the ranks"""
    if workload != "C4":
        return 8
    if 64 % world:
        raise SystemExit("workload C4 shards a 64-view batch: --gpus must divide 64")
    return 64 // world


def views_of(step, rank, world, vps):
    """camera ids of rank `rank` in step `step`: consecutive blocks of the 64-camera pool; successive steps start one block
    further (and one more when the step's views are a multiple of the pool, so that a rank does not see the same cameras again)"""
    per_step = world * vps
    stride = per_step + (vps if per_step % NUM_CAMERAS == 0 else 0)
    return [(step * stride + rank * vps + i) % NUM_CAMERAS for i in range(vps)]


def load_json(name):
    p = os.path.join(ROOT, "profiles", name)
    return json.load(open(p)) if os.path.exists(p) else {}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons of one GPU during the timed region.  The polled query is kept light (SM clock
    and the four throttle reasons, every SMI_MS ms): polling the full recipe line every 50 ms stalled the timed steps in about one
    run out of five (profiles/r02_smi_polling.txt); the maximum clock and the power draw are read once, right after the region."""
    Q = "clocks.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    Q_FULL = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    SMI_MS = os.environ.get("RAIN_BENCH_SMI_MS", "100")                 # diagnostics: polling period; "0" = no polling process at all
    FULL = os.environ.get("RAIN_BENCH_SMI_FULL", "0") == "1"            # diagnostics: poll the full line (incl. clocks.max.sm, power.draw)

    def __init__(self, index):
        self.index, self.rows, self.proc, self.ready = index, [], None, threading.Event()

    def start(self):
        if self.SMI_MS == "0":
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q_FULL if self.FULL else self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", self.SMI_MS, "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            if self.FULL and len(r) >= 7:
                r = [r[0]] + r[3:7]
            self.rows.append((time.time(), r))
            self.ready.set()

    def wait_ready(self, timeout=5.0):
        """returns once the first sample has arrived (nvidia-smi is up and polling)"""
        if self.proc is not None:
            self.ready.wait(timeout)

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable" if self.SMI_MS != "0" else "polling disabled (diagnostic run)"]}
        time.sleep(0.12)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.1 and len(r) >= 5] or [r for _, r in self.rows if len(r) >= 5]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        reasons = [n for i, n in ((1, "hw_slowdown"), (2, "hw_thermal_slowdown"), (3, "sw_thermal_slowdown"), (4, "sw_power_cap"))
                   if any(r[i].lower().startswith("active") for r in rows)]
        out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": None, "samples": len(rows), "period_ms": int(self.SMI_MS), "reasons": reasons}
        try:       # one query outside the timed region
            one = subprocess.run(["nvidia-smi", "--query-gpu=clocks.max.sm,power.draw", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                 capture_output=True, text=True, timeout=10).stdout.strip().split(",")
            out["sm_max_mhz"] = float(one[0]); out["power_w_after"] = float(one[1])
        except Exception:
            pass
        return out


# stages whose roofline is HBM bandwidth; the two tile renderers are bound by instruction issue (FP32 / SFU / LSU slots)
HBM_STAGES = ("preprocess_fwd", "tile_plan", "depth_sort", "scan", "duplicate", "tile_sort", "ranges", "preprocess_bwd", "backward_views")


def stage_bytes(name, P, V, N, X, M=16, views=1):
    """ALGORITHMIC bytes of one launch group (SURVEY.md §8(d) per-unit figures adapted to this library's
    data layout).  P Gaussians, V visible, N (tile, Gaussian) pairs, X pixels; per render unless stated."""
    if views > 1 and name == "preprocess_fwd":
        # view-batched geometry pass: parameters (44 B + the SH row) in once, per view radii / tile count / depth key out always and
        # rect + clamp flags + the 48-byte record out if visible
        return (44 + 12 * M) * P + views * (12 * P + 67 * V)
    if views > 1 and name in ("depth_sort", "scan", "tile_plan"):
        return views * stage_bytes(name, P, V, N, X, M, 1)
    return {
        "preprocess_fwd": 52 * P + (12 * M + 67) * V,                 # 44 in + 8 out always; SH row in + projected record out if visible
        "tile_plan": 12 * P,                                           # tiles_touched + rect in (ranges and digit bases out: a few KB)
        "depth_sort": (4 + 4 * 16) * P,                                # histogram read + 4 passes x (key+value read, key+value write)
        "scan": 12 * P,                                                # order + tiles_touched in, offsets out
        "duplicate": 16 * P + 8 * N,                                   # order/offsets/rect in, tile key + Gaussian id out
        "tile_sort": (4 + 2 * 16) * N,                                 # (unfused fallback) histogram read + 2 passes x 16 B
        # 16-bit tile sort = duplicate + the two radix passes in one stage: order / offsets / rect in, 6-byte pairs out (duplicate),
        # in and out (pass 0), in (pass 1), 4-byte ids out (pass 1)
        "tile_sort_fused": 16 * P + 28 * N,
        "ranges": 4 * N,                                               # sorted tile keys in (ranges out: 8 B per tile)
        "render_fwd": 52 * N + N + 24 * X,                             # id + gathered record per list entry, 8 flag bits out; colour 12 + depth 4 + final_T 4 + n_contrib 4 out
        "render_bwd": 52 * N + N + 20 * X + 96 * P,                    # id + record + flag bits in, dL_dpix/T/n in, accumulator zero fill + read-modify-write
        "preprocess_bwd": 52 * P + (12 * M + 111) * V + (12 * M + 56) * P,   # accumulators + radii + inputs in, every gradient element out
        # view-batched backward of a whole step: parameters in once, per view radii + accumulator (+ record, clamp flags of a
        # visible Gaussian) in and dL_dmean2D out, the 59-float bucket row out once
        "backward_views": (44 + 12 * M) * P + views * (52 * P + 37 * V + 12 * P) + (11 + 3 * M) * 4 * P,
    }.get(name)


def dist_setup(n_gpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    elif n_gpus > 1:
        raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one process per GPU)")
    return rank, world, local


def barrier_sync(world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world, dev):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def make_target(W, H, slot):
    """target image of view slot `slot`: seeded, 8-bit (the native form of training photographs; the loss kernel converts)"""
    return make_grad(W, H, seed=100 + slot % NUM_TARGETS).mul_(0.1).add_(0.5).clamp_(0, 1).mul_(255).round_().to(torch.uint8)


def run_ours(args, rank, world, local):
    from rain_b200 import _lib, multiview as mv
    from rain_b200.loss import l2_loss_grad
    dev = torch.device("cuda", local)
    lib = _lib.load()
    cfg = CONFIGS[args.workload]
    P, W, H, D = cfg["P"], cfg["W"], cfg["H"], cfg["D"]
    VPS = views_per_rank(args.workload, world)
    sc = make_scene(P, D, seed=0).to(dev)
    cams_h = [make_camera(v, NUM_CAMERAS, W, H) for v in range(NUM_CAMERAS)]
    targets_h = [make_target(W, H, v).pin_memory() for v in range(NUM_TARGETS)]
    params = dict(means3D=sc.means3D, opacities=sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    bg = torch.zeros(3, device=dev)
    # N > 1: the bucket lives in symmetric memory and is reduced by the library's one-kernel NVLink all-reduce (csrc/allreduce.cu);
    # `--collective nccl` keeps it a plain tensor reduced by ncclAllReduce
    bucket = mv.GradientBucket(P, 16, dev, symmetric=(world > 1 and args.collective == "rr"))
    kw = dict(bg=bg, sh_degree=D, keep_outputs=False, overlap_allreduce=args.overlap_chunks, overlap_leave_room=not args.overlap_no_room)

    # ---------------- device-resident arm (`value`) ----------------
    cams_d = [c.to(dev) for c in cams_h]
    targets_d = [t.to(dev) for t in targets_h]
    slot = {}
    ar_events = []

    def grad_resident(v, color, depth):
        return l2_loss_grad(color, targets_d[slot[v] % NUM_TARGETS])            # dL/dcolor = colour - target, one fused pass (csrc/loss.cu)

    def step_resident(s, pipeline="two_stream", time_allreduce=False):
        ids = views_of(s, rank, world, VPS)
        slot.clear(); slot.update({v: i for i, v in enumerate(ids)})
        bucket.filled = False
        res = mv.render_views_fwd_bwd(params, cams_d, ids, bucket, grad_resident, pipeline=pipeline, **kw)
        if time_allreduce and world > 1:
            if os.environ.get("RAIN_BENCH_DIAG", "") == "arsync":      # diagnostic only: remove the rank skew from the all-reduce's time
                torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(); mv.allreduce_bucket(bucket); a1.record()
            ar_events.append((a0, a1))
        else:
            mv.allreduce_bucket(bucket)
        return res

    # the clock sampler (an nvidia-smi process polling every 50 ms) is started before the warm-up: its start-up (process launch,
    # NVML initialisation) would otherwise fall into the first timed steps and stall them; only the samples taken inside the
    # timed region [t0, t1] are used
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        sampler.wait_ready()
    # Python's cyclic collector: a full collection over everything torch has imported takes tens of ms of host time and, landing
    # in a step, drains the GPU (every view waits for its num_rendered on the host) — seen as one slow step in one run out of
    # five (profiles/r02_steptimes.txt).  Collect now and move the survivors out of the collector's reach; it stays enabled.
    gc.collect()
    gc.freeze()
    for s in range(args.warmup):
        step_resident(s)
    barrier_sync(world)
    launches0 = lib.rr_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    N_last = 0
    step_marks = []                                  # diagnostic only (RAIN_BENCH_DIAG=steptimes): an event after every timed step
    for s in range(args.steps):
        res = step_resident(args.warmup + s, time_allreduce=True)
        N_last = res[-1].num_rendered
        if os.environ.get("RAIN_BENCH_DIAG", "") == "steptimes":
            step_marks.append(torch.cuda.Event(enable_timing=True)); step_marks[-1].record()
    e1.record()
    barrier_sync(world)
    t1 = time.time()
    launches = lib.rr_launch_count() - launches0
    ms = max_over_ranks(e0.elapsed_time(e1), world, dev)
    if step_marks:
        marks = [e0] + step_marks
        print("steptimes_ms " + " ".join(f"{marks[i].elapsed_time(marks[i + 1]):.2f}" for i in range(len(step_marks))), file=sys.stderr)
    if os.environ.get("RAIN_BENCH_DIAG", "") == "artrace" and world > 1:      # diagnostic only: per-step all-reduce time of every rank
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"artrace_{args.collective}_r{rank}.txt"), "w") as f:
            f.write(" ".join(f"{a.elapsed_time(b):.3f}" for a, b in ar_events) + "\n")
    allreduce_ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in ar_events) / max(1, len(ar_events)), world, dev) if world > 1 else 0.0
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    renders = world * VPS * args.steps
    value = renders / (ms * 1e-3)

    # ---------------- per-stage device times (CUDA events on the launch stream, same steps) ----------------
    # one stream for this pass: in the timed region above the HBM-bound front of view i+1 runs beside the render kernels of
    # view i on a second stream, which stretches both; a kernel's own duration (the roofline numerator) is the serial one
    lib.rr_prof_enable(1)
    _lib.stage_times(reset=True)
    torch.cuda.synchronize()
    for s in range(min(args.steps, max(1, 160 // VPS))):
        step_resident(args.warmup + s, pipeline="one_stream")
    torch.cuda.synchronize()
    st = _lib.stage_times(reset=True)
    lib.rr_prof_enable(0)
    # Gaussians one view sees (radii > 0): the V of the per-stage byte model
    v_probe = views_of(args.warmup + args.steps - 1, rank, world, VPS)[-1]
    slot[v_probe] = VPS - 1
    bucket.filled = False
    V_seen = int((mv.render_views_fwd_bwd(params, cams_d, [v_probe], bucket, grad_resident, bg=bg, sh_degree=D, keep_outputs=True, pipeline="one_stream")[0].radii > 0).sum())

    # ---------------- end-to-end arm (host buffers -> public API -> host scalar) ----------------
    # cameras: one 48-float row per camera (view 16 | proj 16 | campos 3 | pad) in pinned host memory; the rows of a step's
    # views are gathered into one pinned staging block and cross PCIe as ONE copy per step
    cam_rows = torch.zeros(len(cams_h), 48)
    for ci, c in enumerate(cams_h):
        cam_rows[ci, :16] = c.viewmatrix.reshape(-1); cam_rows[ci, 16:32] = c.projmatrix.reshape(-1); cam_rows[ci, 32:35] = c.campos
    cam_rows = cam_rows.pin_memory()
    cam_stage = torch.zeros(VPS, 48).pin_memory()
    cam_dev = torch.zeros(VPS, 48, device=dev)
    dev_cam = [dict(view=cam_dev[i, :16].view(4, 4), proj=cam_dev[i, 16:32].view(4, 4), campos=cam_dev[i, 32:35]) for i in range(VPS)]
    feed = mv.HostFeed([(3, H, W)] * min(VPS, NUM_TARGETS), dev, dtype=torch.uint8)     # target images: pinned host -> device on a side stream, overlapped with the renders
    nbuf = len(feed.buffers)
    loss_dev = torch.zeros(1, device=dev)
    loss_host = torch.zeros(1).pin_memory()
    h2d = VPS * (48 * 4 + 3 * H * W)      # one padded camera row + one 8-bit target image per view
    d2h = 4 + VPS * 8                     # loss scalar + num_rendered (8 B, read inside the forward) per view
    from rain_b200.synthetic import Camera

    def step_e2e(s):
        ids = views_of(s, rank, world, VPS)
        cams_step = {}
        if not E2E_DIAG_NO_FEED:
            feed.stage_slot(0, targets_h[0])       # view i+1's image is staged when view i reaches its loss: one short transfer in flight
        torch.index_select(cam_rows, 0, torch.as_tensor(ids), out=cam_stage)      # the previous step's copy has completed (per-step sync)
        cam_dev.copy_(cam_stage, non_blocking=True)
        for i, v in enumerate(ids):
            cams_step[v] = Camera(W, H, cams_h[v].tanfovx, cams_h[v].tanfovy, dev_cam[i]["view"], dev_cam[i]["proj"], dev_cam[i]["campos"])
        slot.clear(); slot.update({v: i for i, v in enumerate(ids)})
        loss_dev.zero_()

        def grad_e2e(v, color, depth):
            i = slot[v]
            if not E2E_DIAG_NO_FEED and i + 1 < VPS:
                feed.stage_slot((i + 1) % nbuf, targets_h[(i + 1) % NUM_TARGETS])
            tgt = targets_d[i % NUM_TARGETS] if E2E_DIAG_NO_FEED else feed.wait(i % nbuf)
            return l2_loss_grad(color, tgt, loss_dev)     # gradient + 0.5 * sum d^2 into the step's loss scalar
        bucket.filled = False
        mv.render_views_fwd_bwd(params, cams_step, ids, bucket, grad_e2e, **kw)
        mv.allreduce_bucket(bucket)
        loss_host.copy_(loss_dev, non_blocking=True)
        if E2E_DIAG_NO_SYNC:
            return 0.0
        torch.cuda.current_stream().synchronize()
        return float(loss_host.item())

    for s in range(args.warmup):
        step_e2e(s)
    barrier_sync(world)
    e0.record()
    for s in range(args.steps):
        loss = step_e2e(args.warmup + s)
    e1.record()
    barrier_sync(world)
    ms_e2e = max_over_ranks(e0.elapsed_time(e1), world, dev)
    e2e_value = renders / (ms_e2e * 1e-3)

    # ---------------- drop-in arm: one view at a time through the reference's own Python API ----------------
    dropin = None
    if world == 1 and rank == 0 and args.dropin_iters > 0:
        dropin = run_dropin(cfg, sc, cams_d, targets_d, bg, dev, iters=args.dropin_iters, warm=10)

    if bucket.collective is not None:
        bucket.collective.check()                 # no cross-rank wait of the one-kernel all-reduce timed out
    if rank != 0:
        return None
    return dict(cfg=cfg, value=value, ms=ms, e2e_value=e2e_value, ms_e2e=ms_e2e, h2d=h2d, d2h=d2h, launches=launches, stage=st, clocks=clocks,
                N=N_last, V=V_seen, loss=loss, bucket_bytes=bucket.nbytes(), vps=VPS, allreduce_ms=allreduce_ms, dropin=dropin,
                collective=("rr_allreduce_sum_f32 " + json.dumps(bucket.collective.calibration)) if bucket.collective is not None else "ncclAllReduce")


def run_dropin(cfg, sc, cams, targets, bg, dev, iters, warm):
    """SURVEY.md §8(d)'s metric literally: 1 / (time of one _RasterizeGaussians.forward + .backward for one view), CUDA-event
    timed, steady state, >= 50 iterations after 10 warm-ups, through the drop-in module path the unchanged
    gaussian_renderer/__init__.py imports.  Per iteration, as render() + train.py do: a fresh zero screenspace tensor, the
    settings tuple, the module call, the backward from dL/dcolor (= colour - target, the bench's L2 loss), grads set to None."""
    from submodules.diff_gaussian_rasterization.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    W, H, D = cfg["W"], cfg["H"], cfg["D"]
    leaves = [t.detach().clone().requires_grad_(True) for t in (sc.means3D, sc.opacities, sc.shs, sc.scales, sc.rotations)]
    tg = [t.float().div_(255.0) for t in targets]

    def one(i):
        cam = cams[i % NUM_CAMERAS]
        screenspace = torch.zeros_like(leaves[0], requires_grad=True) + 0                 # gaussian_renderer/__init__.py:10
        screenspace.retain_grad()
        rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, bg, 1.0, cam.viewmatrix, cam.projmatrix, D, cam.campos, False, False, 0.3)
        color, radii, depth = GaussianRasterizer(raster_settings=rs)(means3D=leaves[0], means2D=screenspace, opacities=leaves[1], shs=leaves[2],
                                                                     scales=leaves[3], rotations=leaves[4])
        color.backward(gradient=torch.sub(color.detach(), tg[i % NUM_TARGETS]))
        for t in leaves:
            t.grad = None                                                                  # optimizer.zero_grad(set_to_none=True), train.py:147
    for i in range(warm):
        one(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        one(warm + i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return {"value": round(1e3 / ms, 2), "unit": "renders/s", "ms_per_render": round(ms, 4), "iterations": iters, "warmup": warm,
            "api": "GaussianRasterizationSettings + GaussianRasterizer.forward + autograd backward, one view at a time "
                   "(submodules.diff_gaussian_rasterization.diff_gaussian_rasterization, as gaussian_renderer/__init__.py:20-64 calls it)"}


def cpu_oracle_sample(workload, renders):
    """The CPU baseline: the oracle port (scalar C++/OpenMP restatement of forward.cu / backward.cu, all host threads) timed
    on `renders` fwd+bwd renders of the same workload.  Test infrastructure used here only as the measured baseline.  Its
    evaluated / blended (pixel, list entry) pair counts are the ALGORITHMIC work of one render."""
    from oracle.oracle_py import Oracle
    orc = Oracle()
    cfg = CONFIGS[workload]
    sc = make_scene(cfg["P"], cfg["D"], seed=0)
    bg = np.zeros(3, np.float32)
    arrs = [sc.means3D.numpy(), sc.opacities.numpy(), sc.scales.numpy(), sc.rotations.numpy()]
    cnt = dict(fwd_evals=0, fwd_contribs=0, bwd_evals=0, bwd_contribs=0)
    t = time.time()
    for v in range(renders):
        cam = make_camera(v, NUM_CAMERAS, cfg["W"], cfg["H"])
        N, color, radii, depth = orc.forward(bg, *arrs, cam.viewmatrix.numpy(), cam.projmatrix.numpy(), cam.campos.numpy(), cam.tanfovx, cam.tanfovy,
                                             cam.height, cam.width, shs=sc.shs.numpy(), sh_degree=cfg["D"])
        orc.backward(color - (make_target(cfg["W"], cfg["H"], v).float().numpy() / 255.0))
        for k, x in orc.counters().items():
            cnt[k] += x
    dt = time.time() - t
    return dict(value=renders / dt, unit="renders/s", cores=orc.num_threads(), kind="port",
                sample=f"{renders} fwd+bwd renders of {workload} ({cfg['P']} Gaussians, {cfg['W']}x{cfg['H']}, SH degree {cfg['D']}) in {dt:.1f} s, "
                       f"oracle/liboracle.so (g++ -O2 -fopenmp), host threads = {orc.num_threads()} of {os.cpu_count()} cores",
                pairs_per_render={k: x // renders for k, x in cnt.items()})


def run_reference(args, rank, world, local):
    """Reference arm: the UNMODIFIED reference CUDA rasterizer (oracle/_ref/libref_rast.so, compiled from /root/reference for
    sm_100a by oracle/Makefile) on the same GPU, same workload, same step — the reference has no CPU implementation of this
    path, so its own CUDA build is 'the reference' here; the scalar host port is reported beside it as cpu_baseline."""
    if rank != 0:
        return None
    from oracle.oracle_py import REF_SO, RefCuda
    cfg = CONFIGS[args.workload]
    VPS = views_per_rank(args.workload, 1)
    base = {"metric": METRIC, "unit": "renders/s", "impl": "reference", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong" if args.workload == "C4" else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    if not os.path.exists(REF_SO) or not torch.cuda.is_available():
        # fall back to timing the host port, as the tier contract allows
        cb = cpu_oracle_sample(args.workload, max(1, min(3, args.steps)))
        base.update(value=cb["value"], ms_per_step=1e3 * VPS / cb["value"], cpu_baseline=cb,
                    config={"workload": f"{args.workload}: {cfg['P']} Gaussians, {cfg['W']}x{cfg['H']}, SH degree {cfg['D']}", "arm": "host port (oracle)"},
                    e2e={"value": cb["value"], "unit": "renders/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        return base
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    ref = RefCuda()
    sc = make_scene(cfg["P"], cfg["D"], seed=0).to(dev)
    cams = [make_camera(v, NUM_CAMERAS, cfg["W"], cfg["H"]).to(dev) for v in range(NUM_CAMERAS)]
    targets = [make_target(cfg["W"], cfg["H"], v).float().div_(255.0).to(dev) for v in range(NUM_TARGETS)]
    bg = torch.zeros(3, device=dev)

    def step(s):
        for i, v in enumerate(views_of(s, 0, 1, VPS)):
            cam = cams[v]
            N, color, radii, depth = ref.forward(bg, sc.means3D, sc.opacities, sc.scales, sc.rotations, cam.viewmatrix, cam.projmatrix, cam.campos,
                                                 cam.tanfovx, cam.tanfovy, cam.height, cam.width, shs=sc.shs, sh_degree=cfg["D"], sync=False)
            ref.backward(torch.sub(color, targets[i % NUM_TARGETS]), sync=False)
    gc.collect()
    gc.freeze()          # same treatment as our arm: no full cyclic collection inside the timed steps
    for s in range(args.warmup):
        step(s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(args.steps):
        step(args.warmup + s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    value = VPS * args.steps / (ms * 1e-3)
    base.update(value=value, ms_per_step=ms / args.steps, n_gpus=1,
                config={"workload": f"{args.workload}: {cfg['P']} Gaussians, {cfg['W']}x{cfg['H']}, SH degree {cfg['D']}, {VPS} views per step",
                        "arm": "reference CUDA (unmodified sources, rebuilt for sm_100a) on the same B200, one GPU; no gradient accumulation or all-reduce"},
                cpu_baseline={"value": value, "unit": "renders/s", "cores": 0, "kind": "reference",
                              "sample": f"{args.steps} steps x {VPS} views, oracle/_ref/libref_rast.so on cuda:{local} (the reference has no CPU path)"},
                e2e={"value": value, "unit": "renders/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4 * VPS})
    return base


def fp32_peak_tflops(dev, clocks):
    """FP32 peak of this GPU: SMs x 128 lanes x 2 flop x SM clock, read at run time"""
    props = torch.cuda.get_device_properties(dev)
    mhz = (clocks or {}).get("sm_max_mhz") or 1965.0
    return props.multi_processor_count * 128 * 2 * mhz * 1e6 / 1e12, props.multi_processor_count, mhz


def main():
    # stdout carries exactly ONE line, the JSON result: anything a library prints there meanwhile (e.g. the NCCL version
    # banner) is sent to stderr instead
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    def emit(obj):
        real_stdout.write(json.dumps(obj) + "\n")
        real_stdout.flush()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", choices=["C1", "C2", "C3", "C4"])
    ap.add_argument("--cpu-renders", type=int, default=2, help="renders of the workload timed on the host for cpu_baseline (0 = skip)")
    ap.add_argument("--collective", default="rr", choices=["rr", "nccl"], help="N > 1: rr = the library's one-kernel NVLink all-reduce on a symmetric bucket, nccl = ncclAllReduce")
    ap.add_argument("--overlap-chunks", type=int, default=0, help="experiment: all-reduce in K chunks beside the last batched backward (0 = off, the default)")
    ap.add_argument("--overlap-no-room", action="store_true", help="experiment: do not cap the backward's occupancy while overlapping")
    ap.add_argument("--dropin-iters", type=int, default=50, help="timed single-view iterations of the drop-in arm at N = 1 (0 = skip, e.g. under a profiler)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        line = run_reference(args, rank, int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0")))
        if line is not None:
            emit(line)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: rain_b200 has no CPU path")
    rank, world, local = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    r = run_ours(args, rank, world, local)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    cfg = r["cfg"]
    P, X, N, VPS = cfg["P"], cfg["W"] * cfg["H"], r["N"], r["vps"]
    peaks, peak_kind = load_peaks()
    peak = float(peaks["hbm_gbs"])
    fp32_peak, sms, mhz = fp32_peak_tflops(torch.device("cuda", local), r["clocks"])
    cpu = cpu_oracle_sample(args.workload, args.cpu_renders) if (world == 1 and args.cpu_renders > 0) else None
    pairs = (cpu or {}).get("pairs_per_render") or load_json("pairs_per_render.json").get(args.workload)
    ncu_render = load_json("ncu_render.json").get(args.workload, {})
    live = load_json("live_lanes.json").get(args.workload, {})
    traffic = load_json("ncu_traffic.json").get(args.workload, {})
    stages = {}
    n_render = max(1, r["stage"]["render_fwd"][1])
    views_per_launch = {}
    for name, (ms, n) in r["stage"].items():
        if n == 0:
            continue
        per = ms / n
        vpl = views_per_launch[name] = max(1, round(n_render / n))      # the view-batched stages run once per batch of views
        e = {"ms": round(per, 4), "views_per_launch": vpl, "share": None, "bound": "hbm" if name in HBM_STAGES else "issue"}
        b = stage_bytes("tile_sort_fused" if (name == "tile_sort" and r["stage"].get("duplicate", (0, 0))[1] == 0) else name, P, r["V"], N, X, views=vpl)
        if b:
            e.update(alg_bytes=b, gbs=round(b / per / 1e6, 1), frac_of_hbm=round(b / per / 1e6 / peak, 4))
        if name in ("render_fwd", "render_bwd") and pairs:
            ev, co = (pairs["fwd_evals"], pairs["fwd_contribs"]) if name == "render_fwd" else (pairs["bwd_evals"], pairs["bwd_contribs"])
            fl = (FLOP_FWD_REJECT * (ev - co) + FLOP_FWD_BLEND * co) if name == "render_fwd" else (FLOP_BWD_REJECT * (ev - co) + FLOP_BWD_BLEND * co)
            e.update(alg_flops=fl, evaluated_pairs=ev, blended_pairs=co, tflops=round(fl / per / 1e9, 2), frac_of_fp32=round(fl / per / 1e9 / fp32_peak, 4))
            e.update(ncu_render.get(name, {}))          # issue_active, pipe_fma, smem_wavefronts_per_clk_per_sm ... (ncu capture, committed under profiles/)
            e.update(live.get(name, {}))                # live_lane_frac ... (cross-check build counters, committed under profiles/)
        tkey = "preprocess_views" if (name == "preprocess_fwd" and vpl > 1) else name      # the view-batched geometry pass is its own kernel
        if tkey in traffic:
            e["dram_traffic"] = traffic[tkey]
        stages[name] = e
    per_view = {k: 1.0 / views_per_launch[k] for k in stages}                           # launches per rendered view
    tot = sum(v["ms"] * per_view[k] for k, v in stages.items())
    for k, v in stages.items():
        v["share"] = round(v["ms"] * per_view[k] / tot, 4)
    top = max(stages, key=lambda k: stages[k]["ms"] * per_view[k])
    t = stages[top]
    if t["bound"] == "issue" and "tflops" in t:
        roof = {"kernel": top, "bound": "issue", "achieved": t["tflops"], "peak": round(fp32_peak, 2), "unit": "TFLOP/s", "frac": t["frac_of_fp32"],
                "peak_source": f"FP32 pipe: {sms} SMs x 128 lanes x 2 flop x {mhz:.0f} MHz (read at run time)",
                "achieved_def": "oracle-counted evaluated / blended (pixel, list entry) pairs of one render x SURVEY 8(d) flop figures / kernel time",
                "issue_active": t.get("issue_active"), "pipe_fma": t.get("pipe_fma"), "smem_wavefronts_per_clk_per_sm": t.get("smem_wavefronts_per_clk_per_sm"),
                "live_lane_frac": t.get("live_lane_frac")}
    else:
        roof = {"kernel": top, "bound": "hbm", "achieved": t.get("gbs"), "peak": peak, "unit": "GB/s", "frac": t.get("frac_of_hbm"),
                "peak_source": f"{peak_kind} (MEASURED_PEAKS.json hbm_gbs)" if peak_kind == "measured" else "fallback"}
    roof.update(traffic=traffic.get(top), ms_per_launch=t["ms"], share_of_step=t["share"])
    strong = args.workload == "C4"
    line = {
        "metric": METRIC, "value": round(r["value"], 2), "unit": "renders/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(r["ms"] / args.steps, 4), "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: {P} random Gaussians, {cfg['W']}x{cfg['H']}, SH degree {cfg['D']} (BASELINE.json configs[{int(args.workload[1]) - 1}])",
                   "views_per_step_per_gpu": VPS, "global_views_per_step": VPS * world, "num_rendered_per_view": N, "visible_per_view": r["V"],
                   "parallelism": (f"view-sharded dp{world}, one f32 all-reduce of {r['bucket_bytes']} B per step ({r['allreduce_ms']:.3f} ms, "
                                   f"{r['allreduce_ms'] / (r['ms'] / args.steps) * 100:.1f} % of the step)") if world > 1 else "single GPU",
                   "pipeline": ("two streams: preprocess/sort/binning of the next view on a high-priority stream beside the render kernels of the current view; "
                                "stages[] and roofline come from a one-stream pass over the same steps (a kernel's own duration), so their sum exceeds ms_per_step / views"),
                   "l2": f"inputs larger than L2: {P * 59 * 4 // 1000000} MB of Gaussian parameters + {P * 59 * 4 // 1000000} MB gradient bucket per render; "
                         f"every step and rank takes the next block of the {NUM_CAMERAS}-camera pool"},
        "e2e": {"value": round(r["e2e_value"], 2), "unit": "renders/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                "ms_per_step": round(r["ms_e2e"] / args.steps, 4), "loss": r["loss"]},
        "gpu_launches": int(r["launches"]),
        "clocks": r["clocks"],
        "roofline": roof,
        "stages": stages,
    }
    if world > 1:
        line["allreduce"] = {"impl": r["collective"], "ms_per_step": round(r["allreduce_ms"], 4), "bytes": r["bucket_bytes"],
                             "note": (f"--overlap-chunks {args.overlap_chunks}: the reduction runs chunk by chunk beside the step's last backward pass; ms_per_step is "
                                      "only what remains after it") if args.overlap_chunks else "exposed: issued after the step's last backward pass",
                             "bus_gbs": round(2 * (world - 1) / world * r["bucket_bytes"] / max(r["allreduce_ms"], 1e-9) / 1e6, 1)}
    if r["dropin"]:
        line["dropin"] = r["dropin"]
    if cpu:
        line["cpu_baseline"] = cpu
    emit(line)


if __name__ == "__main__":
    main()
