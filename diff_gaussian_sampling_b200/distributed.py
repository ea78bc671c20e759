"""Point-sharded multi-GPU use of the sampler (new functionality — the reference is
single-device, `grep nccl|mpi|torch.distributed` over /root/reference finds nothing).

One process per GPU (`torchrun`), `torch.distributed` for the plumbing.  Query points are sharded,
Gaussian parameters are replicated, every rank bins its slice with the GLOBAL sample bounds (the
reference derives the cell grid from `samples.min/max`, sample_points.cu:70-73, and cell membership
is part of the result, so local bounds would make results depend on the rank count).  Forward needs
no communication; the backward ends with ONE all_reduce(SUM) over the flat
[dL_dmeans | dL_dvalues | dL_dconics] buffer (the three gradients are views of one allocation).

Nothing here touches the device itself: it works on CPU tensors over gloo (tests) and on CUDA
tensors over NCCL / NVLink (bench.py --gpus N).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def _world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_slice(n_points: int, rank: int, world: int):
    """Contiguous, balanced slice [start, end) of the points owned by `rank`."""
    base, extra = divmod(int(n_points), int(world))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def global_bounds(samples_local: torch.Tensor, group=None):
    """(lo, hi) over the points of ALL ranks: local min/max + one MIN and one MAX all_reduce of D floats.
    Ranks without points contribute +/-inf."""
    D = samples_local.shape[-1]
    if samples_local.shape[0]:
        lo = samples_local.min(0).values.clone()
        hi = samples_local.max(0).values.clone()
    else:
        lo = torch.full((D,), float("inf"), dtype=samples_local.dtype, device=samples_local.device)
        hi = -lo
    if _world(group)[1] > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    return lo, hi


def _common_flat_base(tensors):
    """The single contiguous allocation the tensors are views of, if they tile it exactly."""
    base = tensors[0]._base
    if base is None or not base.is_contiguous() or any(t._base is not base for t in tensors):
        return None
    if sum(t.numel() for t in tensors) != base.numel():
        return None
    return base


def allreduce_gradients(grad_means, grad_values, grad_conics, group=None):
    """ONE all_reduce(SUM) for the three gradients.  When they are views of one flat buffer (what
    `_C.sample_gaussians*_backward` returns) the buffer is reduced in place; otherwise they are packed,
    reduced and copied back.  Returns the three (summed) tensors."""
    grads = (grad_means, grad_values, grad_conics)
    if _world(group)[1] == 1:
        return grads
    flat = _common_flat_base(grads)
    nvtx = os.environ.get("DGS_NVTX", "") in ("1", "true", "on") and flat is not None and flat.is_cuda
    if nvtx:
        torch.cuda.nvtx.range_push("dgs/allreduce")
    try:
        if flat is not None:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
            return grads
    finally:
        if nvtx:
            torch.cuda.nvtx.range_pop()
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    out, o = [], 0
    for g in grads:
        out.append(flat[o:o + g.numel()].view_as(g))
        o += g.numel()
    return tuple(out)


class _SyncGrads(torch.autograd.Function):
    """Identity on (means, values, conics); its backward sums their gradients over all ranks with one
    all_reduce.  Placed between the (replicated) parameters and the rank-local sampling ops."""

    @staticmethod
    def forward(ctx, group, means, values, conics):
        ctx.group = group
        return means.view_as(means), values.view_as(values), conics.view_as(conics)

    @staticmethod
    def backward(ctx, gm, gv, gc):
        if gm is None or gv is None or gc is None:
            raise RuntimeError("ShardedGaussianSampler: all three of means/values/conics need gradients")
        gm, gv, gc = allreduce_gradients(gm, gv, gc, ctx.group)
        return None, gm, gv, gc


def sync_grads(means, values, conics, group=None):
    return _SyncGrads.apply(group, means, values, conics)


class ShardedGaussianSampler:
    """GaussianSampler for one rank of a point-sharded job: same methods; `preprocess` takes this rank's slice
    of the samples and bins it with the global bounds; gradients of means/values/conics arriving through the
    sampled values are summed over ranks (one all_reduce per backward of each sampled function)."""

    def __init__(self, debug=False, group=None):
        from . import GaussianSampler
        self._s = GaussianSampler(debug)
        self.group = group

    def preprocess(self, means, values, covariances, conics, samples_local, bounds=None):
        if bounds is None:
            lo, hi = global_bounds(samples_local, self.group)
            bounds = (lo.tolist(), hi.tolist())
        m, v, c = sync_grads(means, values, conics, self.group)
        self._s.preprocess(m, v, covariances, c, samples_local, bounds=bounds)
        self.bounds = bounds

    def __getattr__(self, name):
        return getattr(self._s, name)
