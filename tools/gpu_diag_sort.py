"""GPU diagnostic for the stage-level entry points: onesweep sort vs a stable torch sort, distCUDA2 vs oracle/reference."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rain_b200 import _lib, _C
lib = _lib.load()
dev = torch.device("cuda:0")
st = lambda: torch.cuda.current_stream().cuda_stream

def sort_case(n, key_bytes, b0, b1, seed, skew=False):
    g = torch.Generator(device="cpu").manual_seed(seed)
    if key_bytes == 4:
        keys = torch.randint(0, 2**31 - 1, (n,), generator=g, dtype=torch.int64)
        if skew: keys = keys % 7
        keys_d = keys.to(torch.int32).to(dev)   # values < 2^31: int32 view is fine
        kdt = torch.int32
    else:
        hi = torch.randint(0, 8160, (n,), generator=g, dtype=torch.int64)
        lo = torch.randint(0, 2**31 - 1, (n,), generator=g, dtype=torch.int64)
        if skew: lo = lo % 3
        keys = (hi << 32) | lo
        keys_d = keys.to(dev); kdt = torch.int64
    vals = torch.arange(n, dtype=torch.int32)
    mask = ((1 << b1) - 1) ^ ((1 << b0) - 1)
    exp_perm = torch.sort(keys & mask, stable=True).indices
    ws = torch.empty(lib.rr_sort_ws_bytes(n, key_bytes, b0, b1), dtype=torch.uint8, device=dev)
    kin = keys_d.clone(); kout = torch.empty_like(kin); vin = vals.to(dev); vout = torch.empty_like(vin)
    fn = lib.rr_sort_pairs_u32 if key_bytes == 4 else lib.rr_sort_pairs_u64
    _lib.check(fn(st(), n, kin.data_ptr(), kout.data_ptr(), vin.data_ptr(), vout.data_ptr(), b0, b1, ws.data_ptr()), "sort")
    torch.cuda.synchronize()
    ok_v = bool((vout.cpu().long() == exp_perm).all()); ok_k = bool((kout.cpu().long() == keys[exp_perm]).all())
    # timing
    reps = 5
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    kin2 = keys_d.clone()
    s.record()
    for _ in range(reps):
        fn(st(), n, kin2.data_ptr(), kout.data_ptr(), vin.data_ptr(), vout.data_ptr(), b0, b1, ws.data_ptr())
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / reps
    npass = (b1 - b0 + 7) // 8
    gb = n * (key_bytes + (key_bytes + 4) * 2 * npass) / 1e9
    print(f"sort n={n} key={key_bytes}B bits[{b0},{b1}) skew={skew}: vals_ok={ok_v} keys_ok={ok_k}  {ms:.3f} ms  ~{gb/ms*1e3:.0f} GB/s algorithmic")
    return ok_v and ok_k

ok = True
for (n, kb, b0, b1, skew) in [(1, 4, 0, 32, False), (1000, 4, 0, 32, False), (4096, 4, 0, 13, False), (4097, 4, 0, 13, True), (100_003, 4, 0, 32, False),
                               (1_000_000, 4, 0, 32, False), (8_400_000, 4, 0, 13, False), (8_400_000, 4, 0, 13, True), (3073, 8, 0, 45, False), (2_000_000, 8, 0, 45, False), (2_000_000, 8, 0, 45, True)]:
    ok &= sort_case(n, kb, b0, b1, 1234 + n, skew)
print("SORT_ALL_OK", ok)

# distCUDA2
from oracle.oracle_py import Oracle, RefCuda
orc = Oracle(); ref = RefCuda()
for P in (7, 1000, 100_000, 1_000_000):
    g = torch.Generator().manual_seed(P)
    pts = (torch.rand(P, 3, generator=g) * 2 - 1).float()
    if P == 1000: pts[10] = pts[11]   # coincident points
    d = _C.distCUDA2(pts.to(dev)); torch.cuda.synchronize()
    dr = ref.dist2(pts.to(dev))
    do = torch.from_numpy(orc.dist2(pts.numpy(), brute=(P <= 1000)))
    print(f"dist2 P={P}: mine==ref {int((d.cpu() != dr.cpu()).sum())} mismatches, mine==oracle {int((d.cpu() != do).sum())}, ref==oracle {int((dr.cpu() != do).sum())}")
for P in (1_000_000, 5_000_000):
    pts = (torch.rand(P, 3, generator=torch.Generator().manual_seed(0)) * 2 - 1).float().to(dev)
    for nm, fn in (("mine", lambda: _C.distCUDA2(pts)), ("ref", lambda: ref.dist2(pts))):
        fn(); torch.cuda.synchronize(); t = time.time(); fn(); torch.cuda.synchronize()
        print(f"dist2 P={P} {nm}: {(time.time()-t)*1e3:.2f} ms")
