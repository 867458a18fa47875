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

    def __init__(self, numel, device, group=None, use_multicast=True):
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
        mc = int(self._h.multicast_ptr) if use_multicast else 0
        self.multicast = (mc + off) if mc else 0
        assert int(self._bufs[self.rank]) == self.flat.data_ptr(), "symmetric memory: local mapping differs from the tensor's address"
        self.epoch = 1
        torch.cuda.synchronize(device)          # the zeroed flags are in memory before any peer can signal
        dist.barrier(self.group)

    def allreduce(self, ctas=0):
        """in-place sum over the ranks, on the current stream; every rank calls it once per step"""
        lib = _lib.load()
        with torch.cuda.device(self.flat.device):
            _lib.check(lib.rr_allreduce_sum_f32(torch.cuda.current_stream().cuda_stream, self.rank, self.world, self.numel, self._bufs,
                                                self.multicast or None, self._sigs, self.error.data_ptr(), self.epoch, int(ctas)), "rr_allreduce_sum_f32")
        self.epoch = (self.epoch + 2) & 0x7FFFFFFF
        return self.flat

    def check(self):
        """raise if a cross-rank wait of an earlier all-reduce timed out (synchronises)"""
        if int(self.error.item()) != 0:
            raise RuntimeError("rr_allreduce_sum_f32: a cross-rank wait timed out (did every rank launch the all-reduce?)")
