"""Probe: does the one-kernel all-reduce slow down when the bucket has just been overwritten by a local kernel (as in a step)?"""
import json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rain_b200.collective import SymmetricAllReduce

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = 59_000_000
out = {}


def timed(pre, fn, iters=10):
    tot = 0.0
    for it in range(iters + 3):
        pre()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if it >= 3:
            tot += e0.elapsed_time(e1)
    t = torch.tensor([tot / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


plain = torch.zeros(n, device=dev)
other = torch.zeros(n, device=dev)
for mc in ("peer", "multimem"):
    ar = SymmetricAllReduce(n, dev, use_multicast=(mc == "multimem"))
    if mc == "multimem" and not ar.multicast:
        continue
    out[mc + "_clean"] = timed(lambda: None, ar.allreduce)
    out[mc + "_after_local_overwrite"] = timed(lambda: ar.flat.fill_(1.0), ar.allreduce)
    out[mc + "_after_other_buffer_write"] = timed(lambda: other.fill_(1.0), ar.allreduce)
    ar.check()
out["nccl_clean"] = timed(lambda: None, lambda: dist.all_reduce(plain))
out["nccl_after_local_overwrite"] = timed(lambda: plain.fill_(1.0), lambda: dist.all_reduce(plain))
if rank == 0:
    print(json.dumps(out), flush=True)
dist.barrier(); dist.destroy_process_group()
