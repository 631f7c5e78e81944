"""View-parallel data parallelism for the rasterizer (the reference is single-GPU, so this axis is new).

One process per GPU. Every rank holds a full replica of the Gaussian parameters, renders a disjoint subset of the
batch's cameras, accumulates dL/dtheta over its views locally, and then takes part in ONE exchange step: a sum
all-reduce of the per-Gaussian parameter gradients (59 floats = 236 B per Gaussian at SH degree 3; 708 MB at 3M)
before the (identical) optimiser step. There is no data-path collective inside the rasterizer itself.

The gradients live in one flat fp32 bucket (`GradBucket`): the parameters' .grad tensors are views into it, so the
exchange is a single collective over one contiguous buffer with no packing copies, and zeroing is one memset.
Works with any torch.distributed backend (NCCL over NVLink on the GPU box; gloo in the CPU tests).
"""
from typing import Dict, List, Sequence

import torch
import torch.distributed as dist


def shard_views(num_views: int, rank: int, world_size: int) -> List[int]:
    """Views r, r+n, r+2n, ... of a batch go to rank r. Ragged batches are allowed."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    return list(range(rank, num_views, world_size))


class GradBucket:
    """Flat gradient storage for a dict of leaf parameters."""

    def __init__(self, params: Dict[str, torch.Tensor]):
        if not params:
            raise ValueError("GradBucket needs at least one parameter")
        dev = {p.device for p in params.values()}
        if len(dev) != 1:
            raise ValueError("all parameters must live on one device")
        self.params = dict(params)
        self.device = dev.pop()
        sizes = [p.numel() for p in self.params.values()]
        # 128-byte aligned slices so vectorised kernels may write them
        self.offsets, off = {}, 0
        for (name, p), n in zip(self.params.items(), sizes):
            if p.dtype != torch.float32:
                raise ValueError(f"{name}: gradients are exchanged in fp32")
            self.offsets[name] = off
            off += (n + 31) // 32 * 32
        self.flat = torch.zeros(off, dtype=torch.float32, device=self.device)
        self.views = {name: self.flat[o:o + p.numel()].view_as(p) for (name, p), o in
                      zip(self.params.items(), self.offsets.values())}

    def attach(self):
        """Points every parameter's .grad at its slice of the bucket (autograd then accumulates in place)."""
        for name, p in self.params.items():
            p.grad = self.views[name]

    def zero(self):
        self.flat.zero_()

    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def all_reduce(self, group=None, async_op=False, average=False):
        """The exchange step: sum (optionally mean) of the bucket over all ranks."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if average and not async_op:
            self.flat.div_(dist.get_world_size(group))
        return work


def reduce_densification_stats(grad_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: torch.Tensor, group=None):
    """Every densification interval: xyz_gradient_accum / denom are summed, max_radii2D is max-reduced
    (scene/gaussian_model.py:431-433, train.py:159 on one GPU)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    dist.all_reduce(grad_accum, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)


def render_views_accumulate(rasterize_one, views: Sequence, bucket: GradBucket):
    """Renders `views` one after the other with `rasterize_one(view) -> (outputs, grad_outputs)` and backpropagates
    each into the bucket. Returns the number of views rendered."""
    n = 0
    for v in views:
        outs, gouts = rasterize_one(v)
        torch.autograd.backward(list(outs), list(gouts))
        n += 1
    return n
