"""View-parallel training support (no counterpart in the
single-GPU reference).

Every rank holds a full replica of the Gaussians and the DeformNetwork and
renders its own camera views.  Gradients of all parameters live in ONE flat
fp32 arena (each `.grad` is a view into it), so a step needs one memset and
one `all_reduce` (NCCL over NVLink on GPUs, gloo in the CPU tests).  The
arena after the all-reduce equals the sum over views of the single-GPU
gradients, which is the parity definition for this path.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence

import torch
import torch.distributed as dist


class GradArena:
    """Flat gradient buffer shared by a list of parameters."""

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("GradArena needs at least one parameter that requires grad")
        dev, dt = self.params[0].device, self.params[0].dtype
        if any(p.device != dev or p.dtype != dt for p in self.params):
            raise ValueError("all parameters of an arena must share device and dtype")
        self.offsets = []
        n = 0
        for p in self.params:
            self.offsets.append(n)
            n += p.numel()
        self.flat = torch.zeros(n, device=dev, dtype=dt)
        for p, off in zip(self.params, self.offsets):
            p.grad = self.flat[off:off + p.numel()].view_as(p)

    def zero(self) -> None:
        self.flat.zero_()

    def all_reduce(self, group=None) -> None:
        """Sum over ranks (sum, not mean: parity is defined against the sum of per-view gradients)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)

    def view_of(self, p: torch.nn.Parameter) -> torch.Tensor:
        for q, off in zip(self.params, self.offsets):
            if q is p:
                return self.flat[off:off + p.numel()].view_as(p)
        raise KeyError("parameter is not part of this arena")


def shard_views(num_views: int, rank: int, world_size: int) -> List[int]:
    """Round-robin view assignment: rank g gets views {g, g+G, ...} (training and inference)."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    return list(range(rank, num_views, world_size))


def reduce_densification_stats(grad_norm_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: torch.Tensor,
                               group=None) -> None:
    """The reference accumulates, per view, ||d mean2D[:, :2]|| and a visibility count, and keeps a running
    max of the screen radius (scene/gaussian_model.py:637-639, train.py:364).  Across ranks these are a
    sum, a sum and a max."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(grad_norm_accum, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)
