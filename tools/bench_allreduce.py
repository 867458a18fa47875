"""All-reduce of the gradient bucket (59 floats x 1M Gaussians = 236 MB) over the GPUs of one box: a plain torch allocation
against a buffer from NCCL's own allocator registered with the communicator (ncclMemAlloc + ncclCommRegister, which makes
the NVLS / zero-copy paths eligible).  torchrun --nproc-per-node N tools/bench_allreduce.py"""
import json, os, sys, time
import torch, torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = 59 * 1_000_000


def timed(buf, reps=20):
    for _ in range(5):
        dist.all_reduce(buf)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        dist.all_reduce(buf)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


out = {"world": world, "bytes": n * 4}
plain = torch.zeros(n, device=dev)
out["plain_ms"] = round(timed(plain), 4)
try:
    backend = dist.distributed_c10d._get_default_group()._get_backend(dev)
    pool = torch.cuda.MemPool(backend.mem_allocator)
    with torch.cuda.use_mem_pool(pool):
        reg = torch.zeros(n, device=dev)
    backend.register_mem_pool(pool)
    out["registered_ms"] = round(timed(reg), 4)
except Exception as e:
    out["registered_error"] = repr(e)[:200]
for k in ("plain_ms", "registered_ms"):
    if k in out:
        out[k.replace("_ms", "_busbw_gbs")] = round(2 * (world - 1) / world * n * 4 / out[k] / 1e6, 1)
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
