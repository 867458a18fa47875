"""The all-reduce of the view-sharded step as ONE hand-written kernel over NVLink peer / multicast memory
(csrc/allreduce.cu, rr_allreduce_sum_f32) instead of ncclAllReduce.  No reference counterpart: sharonal10/rain is single-GPU.

The gradient bucket is allocated in SYMMETRIC memory (torch.distributed._symmetric_memory: every rank's buffer is mapped into
every rank's address space and — where the NVSwitch fabric offers it — aliased by one multicast address).  PyTorch is plumbing
here: it allocates and exchanges the mappings; the reduction itself is the library's kernel (in-switch multimem.ld_reduce +
multimem.st when multicast is available, peer loads / stores otherwise), launched on the caller's stream by every rank."""
import ctypes

import torch
import torch.distributed as dist

from . import _lib


class SymmetricAllReduce:
    """An f32 buffer of ``numel`` floats in symmetric memory plus what the one-kernel all-reduce needs (peer mappings, flag
    buffer, epoch).  Collective: every rank of ``group`` constructs it with the same arguments."""

    def __init__(self, numel, device, group=None, use_multicast="auto", ctas=0):
        """use_multicast: True (in-switch multimem reduction; needs a multicast address), False (peer loads / stores), or "auto":
        both variants are timed once on this buffer at construction and the faster one (max over ranks) is kept — at 2 GPUs
        the peer variant wins (a GPU moves the same bytes either way and plain loads stream faster), with more ranks the
        in-switch reduction moves ~1 bucket per GPU instead of 2 (world - 1) / world."""
        import torch.distributed._symmetric_memory as symm_mem
        lib = _lib.load()
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > 8:
            raise RuntimeError("the one-kernel all-reduce covers one NVSwitch domain of at most 8 GPUs")
        self.numel = int(numel)
        self.flat = symm_mem.empty(self.numel, dtype=torch.float32, device=device)
        self._h = symm_mem.rendezvous(self.flat, self.group)
        nsig = lib.rr_allreduce_signal_bytes() // 4
        self._sig = symm_mem.empty(nsig, dtype=torch.int32, device=device)
        self._sig.zero_()
        self._hs = symm_mem.rendezvous(self._sig, self.group)
        self.error = torch.zeros(1, dtype=torch.int32, device=device)
        off, soff = int(getattr(self._h, "offset", 0)), int(getattr(self._hs, "offset", 0))
        vp = ctypes.c_void_p
        self._bufs = (vp * self.world)(*[int(p) + off for p in self._h.buffer_ptrs])
        self._sigs = (vp * self.world)(*[int(p) + soff for p in self._hs.buffer_ptrs])
        mc = int(self._h.multicast_ptr)
        self._multicast_addr = (mc + off) if mc else 0
        self.multicast = self._multicast_addr if use_multicast else 0
        self.ctas = int(ctas)
        assert int(self._bufs[self.rank]) == self.flat.data_ptr(), "symmetric memory: local mapping differs from the tensor's address"
        self.epoch = 1
        self.calibration = None
        torch.cuda.synchronize(device)          # the zeroed flags are in memory before any peer can signal
        dist.barrier(self.group)
        if use_multicast == "auto":
            self._calibrate()

    def _calibrate(self, iters=4):
        """time both variants on the (zero-filled) buffer and keep the faster; every rank takes the same decision"""
        self.flat.zero_()
        times = {}
        for name, addr in (("peer", 0), ("multimem", self._multicast_addr)):
            if name == "multimem" and not addr:
                continue
            self.multicast = addr
            self.allreduce()                                            # warm-up
            torch.cuda.synchronize(self.flat.device); dist.barrier(self.group)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                self.allreduce()
            e1.record(); torch.cuda.synchronize(self.flat.device)
            t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=self.flat.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            times[name] = float(t.item())
        best = min(times, key=times.get)
        self.multicast = self._multicast_addr if best == "multimem" else 0
        self.calibration = dict(times_ms=times, chosen=best)
        self.check()

    def allreduce(self, ctas=None):
        """in-place sum over the ranks, on the current stream; every rank calls it once per step"""
        lib = _lib.load()
        with torch.cuda.device(self.flat.device):
            _lib.check(lib.rr_allreduce_sum_f32(torch.cuda.current_stream().cuda_stream, self.rank, self.world, self.numel, self._bufs,
                                                self.multicast or None, self._sigs, self.error.data_ptr(), self.epoch,
                                                self.ctas if ctas is None else int(ctas)), "rr_allreduce_sum_f32")
        self.epoch = (self.epoch + 2) & 0x7FFFFFFF
        return self.flat

    def allreduce_ranges(self, ranges, ctas=None, threads=256):
        """in-place sum of the sub-ranges [(offset, length), ...] (floats, multiples of 4, at most 8) of the buffer in one kernel,
        on the current stream; every rank calls it with the same ranges.  Smaller CTAs (256 threads) so that one fits on every
        SM beside a co-running kernel that left room for it."""
        lib = _lib.load()
        n = len(ranges)
        offs = (ctypes.c_int64 * n)(*[int(o) for o, _ in ranges])
        lens = (ctypes.c_int64 * n)(*[int(l) for _, l in ranges])
        with torch.cuda.device(self.flat.device):
            _lib.check(lib.rr_allreduce_sum_f32_ranges(torch.cuda.current_stream().cuda_stream, self.rank, self.world, n, offs, lens, self._bufs,
                                                       self.multicast or None, self._sigs, self.error.data_ptr(), self.epoch,
                                                       self.ctas if ctas is None else int(ctas), int(threads)), "rr_allreduce_sum_f32_ranges")
        self.epoch = (self.epoch + 2) & 0x7FFFFFFF

    def check(self):
        """raise if a cross-rank wait of an earlier all-reduce timed out (synchronises)"""
        if int(self.error.item()) != 0:
            raise RuntimeError("rr_allreduce_sum_f32: a cross-rank wait timed out (did every rank launch the all-reduce?)")
