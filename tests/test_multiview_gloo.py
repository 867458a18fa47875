"""CPU tests of the view-sharded data-parallel host logic (rain_b200.multiview) with the gloo backend, world_size 2.
The CUDA kernels cannot run here, so a test double stands in for rain_b200._C and produces a deterministic gradient
per (view, Gaussian); what is under test is the sharding, the bucket layout/accumulation protocol and the single
all-reduce — not rendering."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rain_b200 import multiview as mv
from rain_b200.synthetic import make_camera


def test_shard_views_partitions_exactly():
    for n in (0, 1, 7, 8, 64, 65):
        for w in (1, 2, 3, 8):
            parts = [mv.shard_views(n, r, w) for r in range(w)]
            flat = [v for p in parts for v in p]
            assert flat == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        mv.shard_views(4, 2, 2)


def test_bucket_layout_is_one_flat_buffer_of_59_floats_per_gaussian():
    b = mv.GradientBucket(10, 16, "cpu")
    assert b.floats_per_gaussian == 59 and b.flat.numel() == 590 and b.nbytes() == 2360
    base = b.flat.data_ptr()
    offs = [b.views[k].data_ptr() - base for k in mv.SEGMENTS]
    assert offs == [0, 120, 120 + 1920, 120 + 1920 + 40, 120 + 1920 + 40 + 120]
    assert b.views["sh"].shape == (10, 16, 3) and b.views["rotations"].shape == (10, 4)
    b.views["opacity"].fill_(2.0)
    assert b.flat.sum().item() == 20.0
    b.zero_()
    assert not b.flat.any()


class FakeBackend:
    """Stands in for rain_b200._C: gradient of view v for Gaussian i, slot s is (v+1) * (i+1) * scale(s) on the Gaussians
    the view 'sees' (i % (v+2) != 0), honouring the overwrite / accumulate protocol of rr_backward."""
    scale = dict(means3D=1.0, sh=0.5, opacity=2.0, scales=3.0, rotations=-1.0)

    def __init__(self):
        self.calls = []

    def rasterize_gaussians(self, bg, means3D, colors, opac, scales, rots, smod, cov, view, proj, tx, ty, H, W, sh, D, campos, pre, dbg, lp):
        P = means3D.shape[0]
        v = int(round(float(campos[0]) * 1000)) % 1000          # the test encodes the view id in campos
        radii = torch.tensor([(i % (v + 2) != 0) for i in range(P)], dtype=torch.int32)
        return P, torch.full((3, H, W), float(v)), radii, torch.zeros(1, H, W), torch.tensor([v]), torch.empty(0), torch.empty(0)

    def rasterize_gaussians_backward(self, bg, means3D, radii, colors, scales, rots, smod, cov, view, proj, tx, ty, dpix, sh, D, campos,
                                     geom, R, binning, img, dbg, lp, grads_out=None, accumulate=False):
        v = int(geom[0])
        P = means3D.shape[0]
        self.calls.append((v, accumulate))
        seen = (radii > 0).float()
        idx = torch.arange(1, P + 1, dtype=torch.float32)
        for k, t in grads_out.items():
            val = (v + 1) * idx * self.scale[k] * seen * float(dpix.flatten()[0])
            val = val.view(P, *([1] * (t.dim() - 1))).expand_as(t)
            if accumulate:
                t += val
            else:
                t.copy_(val)
        return (torch.zeros(P, 3),) + (None,) * 7


def expected_sum(P, views, slot):
    idx = torch.arange(1, P + 1, dtype=torch.float32)
    tot = torch.zeros(P)
    for v in views:
        seen = torch.tensor([(i % (v + 2) != 0) for i in range(P)], dtype=torch.float32)
        tot += (v + 1) * idx * FakeBackend.scale[slot] * seen
    return tot


def make_inputs(P, num_views):
    params = dict(means3D=torch.zeros(P, 3), opacities=torch.ones(P, 1), shs=torch.zeros(P, 16, 3), scales=torch.ones(P, 3), rotations=torch.ones(P, 4))
    cams = []
    for v in range(num_views):
        c = make_camera(v, num_views, 32, 16)
        c.campos = torch.tensor([v / 1000.0, 0.0, 0.0])
        cams.append(c)
    return params, cams


def test_local_accumulation_protocol_first_view_overwrites_then_accumulates():
    P, V = 13, 5
    params, cams = make_inputs(P, V)
    bucket = mv.GradientBucket(P, 16, "cpu")
    bucket.flat.fill_(123.0)                                 # stale contents must not leak into the result
    be = FakeBackend()
    res = mv.train_step_views(params, cams, V, bucket, lambda v, color, depth: torch.ones_like(color), bg=torch.zeros(3), sh_degree=3, backend=be)
    assert [c[1] for c in be.calls] == [False, True, True, True, True]
    assert [r.view for r in res] == list(range(V)) and float(res[3].color[0, 0, 0]) == 3.0
    for k in mv.SEGMENTS:
        got = bucket.views[k].reshape(P, -1)
        assert torch.equal(got, expected_sum(P, range(V), k)[:, None].expand_as(got)), k


def _worker(rank, world, port, P, V, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        params, cams = make_inputs(P, V)
        bucket = mv.GradientBucket(P, 16, "cpu")
        be = FakeBackend()
        res = mv.train_step_views(params, cams, V, bucket, lambda v, color, depth: torch.ones_like(color), rank=rank, world_size=world,
                                  bg=torch.zeros(3), sh_degree=3, backend=be)
        torch.save(dict(flat=bucket.flat.clone(), views=[r.view for r in res]), os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("V", [5, 1])
def test_two_ranks_allreduce_equals_single_rank_sum(tmp_path, V):
    P, world = 11, 2
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(world, port, P, V, str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(os.path.join(str(tmp_path), f"r{r}.pt")) for r in range(world)]
    assert sorted(outs[0]["views"] + outs[1]["views"]) == list(range(V))          # every view rendered exactly once
    assert torch.equal(outs[0]["flat"], outs[1]["flat"])                           # replicas agree after the all-reduce
    ref = mv.GradientBucket(P, 16, "cpu")
    for k in mv.SEGMENTS:
        ref.views[k].copy_(expected_sum(P, range(V), k).view(P, *([1] * (ref.views[k].dim() - 1))).expand_as(ref.views[k]))
    assert torch.equal(outs[0]["flat"], ref.flat)                                  # == the sum over ALL views (a rank with no view adds zeros)


def _stats_worker(rank, world, port, P, out_dir):
    from rain_b200.optim import DensificationStats
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        st = DensificationStats(P, "cpu")
        st.max_radii2D.fill_(3.0); st.xyz_gradient_accum.fill_(1.0); st.denom.fill_(2.0)      # replicated totals of earlier steps
        idx = torch.arange(P, dtype=torch.float32)
        # this rank's delta of the step (what add_view's kernel leaves behind; the kernel itself needs a GPU)
        st._sum[:P] = (rank + 1) * idx; st._sum[P:] = float(rank + 1); st._max.copy_(idx * (1 + rank) % 7)
        st.sync()
        assert not st._sum.any() and not st._max.any()                                        # deltas are consumed
        torch.save(dict(m=st.max_radii2D, a=st.xyz_gradient_accum, d=st.denom), os.path.join(out_dir, f"s{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_densification_stats_are_summed_and_maxed_across_ranks(tmp_path):
    """scene/gaussian_model.py:419-421 over a view-sharded step: xyz_gradient_accum / denom are sums over ALL ranks' views,
    max_radii2D the max, and every replica ends up with the same totals (SURVEY.md 8(e) caveat)."""
    P, world = 9, 2
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_stats_worker, args=(world, port, P, str(tmp_path)), nprocs=world, join=True)
    a, b = (torch.load(os.path.join(str(tmp_path), f"s{r}.pt")) for r in range(world))
    for k in "mad":
        assert torch.equal(a[k], b[k])
    idx = torch.arange(P, dtype=torch.float32)
    assert torch.equal(a["a"].view(-1), 1.0 + 3 * idx) and torch.equal(a["d"].view(-1), torch.full((P,), 5.0))
    assert torch.equal(a["m"], torch.maximum(torch.full((P,), 3.0), torch.maximum(idx % 7, idx * 2 % 7)))
