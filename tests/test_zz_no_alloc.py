"""rr_sort_pairs_u32_noalloc (include/rain_rast.h): the u32 pair sort with its workspace in a device array owned by the
library, for callers that cannot allocate.  Checked through the C ABI against rr_sort_pairs_u32 (bit-exact: it is integer
work) and against torch's stable sort, on two calls in a row that share the built-in array.  Named test_zz_* so that it is
collected last."""
import pytest
import torch

from rain_b200 import _lib

pytestmark = pytest.mark.gpu


def _inputs(n, b1, seed, device):
    g = torch.Generator().manual_seed(seed)
    keys = torch.randint(0, 2 ** 31 - 1, (n,), generator=g, dtype=torch.int64)
    if seed % 2:
        keys = keys % 1000                                        # heavy ties: stability matters
    return keys, keys.to(torch.int32).to(device), torch.arange(n, dtype=torch.int32, device=device)


def test_noalloc_sort_matches_the_workspace_entry_point_twice_in_a_row(device):
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    nmax = lib.rr_sort_noalloc_max_pairs()
    assert nmax == 1 << 22
    # two different sizes and bit ranges back to back: the second call finds the first one's histogram / look-back words in the array
    cases = [(1_000_003, 0, 32, 7), (300_001, 0, 13, 8), (nmax, 0, 30, 9)]
    # every buffer of the test is allocated before the first call under test
    bufs = []
    for n, b0, b1, seed in cases:
        keys, kd, vin = _inputs(n, b1, seed, device)
        ws = torch.empty(lib.rr_sort_ws_bytes(n, 4, b0, b1), dtype=torch.uint8, device=device)
        bufs.append(dict(keys=keys, k_ref=kd.clone(), v_ref=vin.clone(), k_new=kd.clone(), v_new=vin.clone(), ws=ws,
                         ko_ref=torch.empty_like(kd), vo_ref=torch.empty_like(vin), ko_new=torch.full_like(kd, -1), vo_new=torch.full_like(vin, -1)))
    for (n, b0, b1, seed), b in zip(cases, bufs):
        _lib.check(lib.rr_sort_pairs_u32(stream, n, b["k_ref"].data_ptr(), b["ko_ref"].data_ptr(), b["v_ref"].data_ptr(), b["vo_ref"].data_ptr(),
                                         b0, b1, b["ws"].data_ptr()), "sort")
    torch.cuda.synchronize()
    for i, ((n, b0, b1, seed), b) in enumerate(zip(cases, bufs)):
        held, free = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        _lib.check(lib.rr_sort_pairs_u32_noalloc(stream, n, b["k_new"].data_ptr(), b["ko_new"].data_ptr(), b["v_new"].data_ptr(), b["vo_new"].data_ptr(), b0, b1),
                   "sort_noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == held
        if i > 0:                                                   # (the first call may load the kernels' module)
            assert torch.cuda.mem_get_info()[0] >= free - (1 << 20)
        assert torch.equal(b["ko_new"], b["ko_ref"]) and torch.equal(b["vo_new"], b["vo_ref"]), (n, b0, b1)
        mask = ((1 << b1) - 1) ^ ((1 << b0) - 1)
        perm = torch.sort(b["keys"] & mask, stable=True).indices
        assert torch.equal(b["vo_new"].long().cpu(), perm) and torch.equal(b["ko_new"].long().cpu(), b["keys"][perm])


def test_noalloc_sort_rejects_sizes_beyond_its_workspace(device):
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    t = torch.zeros(16, dtype=torch.int32, device=device)
    n = lib.rr_sort_noalloc_max_pairs() + 1
    rc = lib.rr_sort_pairs_u32_noalloc(stream, n, t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), 0, 32)
    assert rc != 0
    assert "limit" in lib.rr_last_error().decode()
    assert lib.rr_sort_pairs_u32_noalloc(stream, 16, t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), 5, 5) != 0      # empty bit range
    assert lib.rr_sort_pairs_u32_noalloc(stream, 0, None, None, None, None, 0, 32) == 0                                        # nothing to do
    torch.cuda.synchronize()
