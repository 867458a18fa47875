"""Multi-GPU probe: which algorithm NCCL picks for the gradient-bucket all-reduce, whether the box
offers multicast (NVLS / multimem) through torch's symmetric memory, and what torch's own multimem / two-shot all-reduce
reach on the bucket — the headroom over ncclAllReduce.  One JSON line on rank 0.
    NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING python -m torch.distributed.run --nproc-per-node N ... tools/probe_symm.py [P]"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = P * 59
out = {"world": world, "floats": n, "bytes": n * 4}


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


x = torch.randn(n, device=dev)
ms = timeit(lambda: dist.all_reduce(x))
out["nccl_allreduce_ms"] = ms
out["nccl_busbw_gbs"] = 2 * (world - 1) / world * n * 4 / ms / 1e6
try:
    import torch.distributed._symmetric_memory as symm_mem
    t = symm_mem.empty(n, dtype=torch.float32, device=dev)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    out["has_multicast"] = bool(hdl.multicast_ptr)
    out["multicast_ptr_nonzero"] = hdl.multicast_ptr != 0
    out["signal_pad_size"] = hdl.signal_pad_size
    out["buffer_size"] = hdl.buffer_size
    gname = dist.group.WORLD.group_name
    t.copy_(torch.full((n,), float(rank + 1), device=dev))
    dist.barrier()
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_"):
        try:
            op = getattr(torch.ops.symm_mem, name)
            t.fill_(float(rank + 1)); torch.cuda.synchronize(); dist.barrier()
            op(t, "sum", gname)
            torch.cuda.synchronize()
            ok = bool((t == world * (world + 1) / 2).all())
            out[name + "ok"] = ok
            out[name + "ms"] = timeit(lambda: op(t, "sum", gname))
            out[name + "busbw_gbs"] = 2 * (world - 1) / world * n * 4 / out[name + "ms"] / 1e6
        except Exception as ex:
            out[name + "error"] = repr(ex)[:300]
except Exception as ex:
    out["symm_mem_error"] = repr(ex)[:400]
# this library's one-kernel all-reduce (csrc/allreduce.cu) on a symmetric bucket: correctness against NCCL's result, then time
try:
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from rain_b200.collective import SymmetricAllReduce
    for mc in (True, False):
        tag = "rr_multimem_" if mc else "rr_peer_"
        ar = SymmetricAllReduce(n, dev, use_multicast=mc)
        if mc and not ar.multicast:
            out[tag + "skipped"] = "no multicast address"
            continue
        g = torch.Generator(device=dev).manual_seed(1234 + rank)
        src = torch.randn(n, device=dev, generator=g)
        ar.flat.copy_(src)
        ref = src.clone(); dist.all_reduce(ref)
        torch.cuda.synchronize(); dist.barrier()
        ar.allreduce()
        torch.cuda.synchronize()
        ar.check()
        err = float((ar.flat - ref).abs().max() / ref.abs().max())
        lo, hi = ar.flat.clone(), ar.flat.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        out[tag + "rel_err_vs_nccl"] = err
        out[tag + "replicas_bit_identical"] = bool(torch.equal(lo, hi))
        for ctas in (148, 256, 64):
            ms = timeit(lambda: ar.allreduce(ctas))
            out[tag + f"ms_ctas{ctas}"] = ms
        ar.check()
        best = min(out[tag + f"ms_ctas{c}"] for c in (148, 256, 64))
        out[tag + "busbw_gbs"] = 2 * (world - 1) / world * n * 4 / best / 1e6
except Exception as ex:
    import traceback
    out["rr_allreduce_error"] = traceback.format_exc()[-600:]
if rank == 0:
    print(json.dumps(out), flush=True)
dist.barrier()
dist.destroy_process_group()
