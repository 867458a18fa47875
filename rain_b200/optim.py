"""Host side of the fused parameter step (csrc/param_step.cu, rr_param_step):
activation backward + torch.optim.Adam semantics + re-activation for the next render in one pass per Gaussian, on the flat
gradient bucket that rain_b200.multiview accumulates and all-reduces.  Mirrors GaussianModel.training_setup /
optimizer.step() of the reference (scene/gaussian_model.py:139-157, train.py:146).  CUDA tensors only."""
import ctypes

import torch

from . import _lib

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


class GaussianParamStep:
    """leaves: dict(xyz[P,3], f_dc[P,1,3], f_rest[P,M-1,3], opacity[P,1], scaling[P,3], rotation[P,4]) raw CUDA tensors
    (updated in place); lrs: dict over GROUPS (scene/gaussian_model.py:144-151)."""

    def __init__(self, leaves, lrs, betas=(0.9, 0.999), eps=1e-15):
        self.leaves = {k: leaves[k] for k in GROUPS}
        for k, t in self.leaves.items():
            if not t.is_cuda:
                raise RuntimeError(f"rain_b200 has no CPU path: leaf '{k}' must be a CUDA tensor")
            if t.dtype != torch.float32 or not t.is_contiguous():
                raise RuntimeError(f"leaf '{k}' must be contiguous float32")
        self.P = int(self.leaves["xyz"].shape[0])
        self.M = 1 + int(self.leaves["f_rest"].shape[1])
        dev = self.leaves["xyz"].device
        n = self.P * (11 + 3 * self.M)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.lrs = dict(lrs)
        self.betas, self.eps, self.steps = betas, eps, 0
        P, M = self.P, self.M
        self.activated = dict(means3D=self.leaves["xyz"], shs=torch.empty(P, M, 3, device=dev), opacities=torch.empty(P, 1, device=dev),
                              scales=torch.empty(P, 3, device=dev), rotations=torch.empty(P, 4, device=dev))

    def step_leaf(self, grads_flat):
        """grads_flat: flat f32 [P*(11+3M)] in the LEAF bucket layout (GradientBucket(layout="leaf"), gradients w.r.t. the raw leaves
        as the leaf fast path of rain_b200.multiview produces them).  Plain Adam over the six leaves; returns the leaves dict, which
        is what the next leaf-mode render reads (no activated copies)."""
        lib = _lib.load()
        if not grads_flat.is_cuda or grads_flat.numel() != self.exp_avg.numel():
            raise RuntimeError("grads_flat must be the CUDA gradient bucket of this parameter set")
        self.steps += 1
        hyper = (ctypes.c_float * 9)(*[float(self.lrs[k]) for k in GROUPS], float(self.betas[0]), float(self.betas[1]), float(self.eps))
        L = self.leaves
        with torch.cuda.device(grads_flat.device):
            _lib.check(lib.rr_param_step_leaf(torch.cuda.current_stream().cuda_stream, self.P, self.M, L["xyz"].data_ptr(), L["f_dc"].data_ptr(),
                                              L["f_rest"].data_ptr() if self.M > 1 else None, L["opacity"].data_ptr(), L["scaling"].data_ptr(), L["rotation"].data_ptr(),
                                              self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), grads_flat.data_ptr(), hyper, self.steps), "rr_param_step_leaf")
        return L

    def step(self, grads_flat):
        """grads_flat: flat f32 [P*(11+3M)] in the bucket layout (w.r.t. the activated values).  Returns the activated
        inputs of the next render: dict(means3D, shs, opacities, scales, rotations)."""
        lib = _lib.load()
        if not grads_flat.is_cuda or grads_flat.numel() != self.exp_avg.numel():
            raise RuntimeError("grads_flat must be the CUDA gradient bucket of this parameter set")
        self.steps += 1
        hyper = (ctypes.c_float * 9)(*[float(self.lrs[k]) for k in GROUPS], float(self.betas[0]), float(self.betas[1]), float(self.eps))
        L, A = self.leaves, self.activated
        with torch.cuda.device(grads_flat.device):
            _lib.check(lib.rr_param_step(torch.cuda.current_stream().cuda_stream, self.P, self.M, L["xyz"].data_ptr(), L["f_dc"].data_ptr(),
                                         L["f_rest"].data_ptr() if self.M > 1 else None, L["opacity"].data_ptr(), L["scaling"].data_ptr(), L["rotation"].data_ptr(),
                                         self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), grads_flat.data_ptr(), hyper, self.steps,
                                         A["shs"].data_ptr(), A["opacities"].data_ptr(), A["scales"].data_ptr(), A["rotations"].data_ptr()), "rr_param_step")
        return A


def add_densification_stats(radii, dL_dmean2D, max_radii2D, xyz_gradient_accum, denom):
    """train.py:132-134 + GaussianModel.add_densification_stats (scene/gaussian_model.py:419-421) for one rendered view, in one
    kernel: radii i32[P], dL_dmean2D f32[P,3] (viewspace_points.grad); max_radii2D[P], xyz_gradient_accum[P,1], denom[P,1] f32
    are updated in place for the Gaussians with radii > 0."""
    lib = _lib.load()
    for t in (radii, dL_dmean2D, max_radii2D, xyz_gradient_accum, denom):
        if not t.is_cuda or not t.is_contiguous():
            raise RuntimeError("rain_b200 has no CPU path: contiguous CUDA tensors required")
    if radii.dtype != torch.int32 or any(t.dtype != torch.float32 for t in (dL_dmean2D, max_radii2D, xyz_gradient_accum, denom)):
        raise RuntimeError("radii must be int32 and the statistics float32")
    P = radii.numel()
    if dL_dmean2D.numel() != 3 * P or max_radii2D.numel() != P or xyz_gradient_accum.numel() != P or denom.numel() != P:
        raise RuntimeError("shape mismatch between radii and the statistics")
    with torch.cuda.device(radii.device):
        _lib.check(lib.rr_densify_stats(torch.cuda.current_stream().cuda_stream, P, radii.data_ptr(), dL_dmean2D.data_ptr(), max_radii2D.data_ptr(),
                                        xyz_gradient_accum.data_ptr(), denom.data_ptr()), "rr_densify_stats")


class DensificationStats:
    """Multi-GPU-consistent densification statistics (SURVEY.md 8(e) "caveat for callers", 8(f) rank 3): the three per-Gaussian
    statistics of GaussianModel (scene/gaussian_model.py:419-421, train.py:132-134) kept identical on every replica of a
    view-sharded step.  They are per view and non-linear (a norm, a max), so they cannot be derived from the all-reduced
    gradients: each rank folds its views into a per-step DELTA (``add_view``), and ``sync`` combines the deltas across ranks —
    ONE sum all-reduce over [xyz_gradient_accum | denom] (2P floats) and ONE max all-reduce over max_radii2D (P floats) —
    and folds them into the replicated totals.  With one rank ``sync`` only folds."""

    def __init__(self, P, device):
        f = dict(dtype=torch.float32, device=device)
        self.P = int(P)
        self.max_radii2D = torch.zeros(self.P, **f)
        self.xyz_gradient_accum = torch.zeros(self.P, 1, **f)
        self.denom = torch.zeros(self.P, 1, **f)
        self._sum = torch.zeros(2 * self.P, **f)          # [delta of xyz_gradient_accum | delta of denom]
        self._max = torch.zeros(self.P, **f)              # max of radii over this rank's views of the step

    def add_view(self, radii, dL_dmean2D):
        """fold one rendered view (its radii and viewspace_points.grad) into this rank's delta of the step"""
        add_densification_stats(radii, dL_dmean2D, self._max, self._sum[:self.P], self._sum[self.P:])

    def delta_tensors(self):
        """(max_radii2D, xyz_gradient_accum, denom) delta tensors of this rank's step, for kernels that fold views in directly
        (rr_backward_views)"""
        return self._max, self._sum[:self.P], self._sum[self.P:]

    def sync(self, group=None):
        """combine the step's deltas across the ranks of ``group`` and fold them into the replicated statistics"""
        import torch.distributed as dist
        if group != "skip" and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self._sum, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(self._max, op=dist.ReduceOp.MAX, group=group)
        self.xyz_gradient_accum.view(-1).add_(self._sum[:self.P])
        self.denom.view(-1).add_(self._sum[self.P:])
        torch.maximum(self.max_radii2D, self._max, out=self.max_radii2D)
        self._sum.zero_()
        self._max.zero_()
