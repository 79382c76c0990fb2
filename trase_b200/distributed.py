"""View-parallel training support (no counterpart in the
single-GPU reference).

Every rank holds a full replica of the Gaussians and the DeformNetwork and
renders its own camera views.  Gradients of all parameters live in ONE flat
fp32 arena (each `.grad` is a view into it), so a step needs one memset and
one `all_reduce` (NCCL over NVLink on GPUs, gloo in the CPU tests).  The
arena after the all-reduce equals the sum over views of the single-GPU
gradients, which is the parity definition for this path.

Overlap: the Gaussians' own gradients (the bulk of the arena: ~90 floats per
Gaussian) are final as soon as the rasterizer backward and the activation
glue have run, while the DeformNetwork backward (about 40 % of a step) is
still ahead.  Parameters passed as `early` are laid out first in the arena;
once `arm()` has been called, the hook of the last early parameter to
accumulate starts an asynchronous all-reduce of that segment, which then runs
under the MLP backward, and `all_reduce()` only has the small tail left.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence

import torch
import torch.distributed as dist


class GradArena:
    """Flat gradient buffer shared by a list of parameters.

    Every `.grad` becomes a view into `flat`.  Parameters of an arena are marked so that the native DeformNetwork
    backward and the fused parameter-activation backward (`_assemble.py`) may add gradients straight into their slots
    instead of returning them to autograd; drive such models with `.backward()` (not torch.autograd.grad), and do not
    hang tensor (pre-accumulation) gradient hooks on those parameters if the direct path is wanted (such parameters
    are handled the ordinary way; post-accumulate hooks are fine)."""

    def __init__(self, params: Iterable[torch.nn.Parameter], early: Iterable[torch.nn.Parameter] = ()):
        params = [p for p in params if p.requires_grad]
        early_ids = {id(p) for p in early}
        first = [p for p in params if id(p) in early_ids]
        self.params: List[torch.nn.Parameter] = first + [p for p in params if id(p) not in early_ids]
        if not self.params:
            raise ValueError("GradArena needs at least one parameter that requires grad")
        self.n_early_params = len(first)
        self.n_early = sum(p.numel() for p in first)
        self._armed, self._seen, self._work, self._group = False, 0, None, None
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
            # opt-in mark for kernels that can add a gradient straight into its arena slot (the native DeformNetwork
            # backward, trase_b200/_deform_native.py DIRECT_GRAD_ACCUM): records WHICH buffer was agreed on
            p._trase_arena_grad_ptr = p.grad.data_ptr()
        for p in first:
            # fires when autograd reaches the parameter's AccumulateGrad node — also when a kernel has already added
            # the gradient into the slot and handed autograd None (trase_b200/_assemble.py)
            p.register_post_accumulate_grad_hook(self._early_hook)

    def zero(self) -> None:
        self.flat.zero_()

    @staticmethod
    def _multi_rank(group) -> bool:
        return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1

    def arm(self, group=None) -> None:
        """Call before the LAST backward of a step (earlier views of the same step still accumulate
        into the early segment, so it is not final before that)."""
        self._armed, self._seen, self._group = self.n_early > 0 and self._multi_rank(group), 0, group

    def _early_hook(self, _param) -> None:
        if not self._armed:
            return
        self._seen += 1
        if self._seen == self.n_early_params:
            self._armed = False
            self._work = dist.all_reduce(self.flat[:self.n_early], op=dist.ReduceOp.SUM, group=self._group,
                                         async_op=True)

    def all_reduce(self, group=None) -> None:
        """Sum over ranks (sum, not mean: parity is defined against the sum of per-view gradients).
        Finishes an overlapped early-segment reduction if one was started by `arm()`."""
        if not self._multi_rank(group):
            return
        if self._work is not None:
            self._work.wait()
            self._work = None
            if self.n_early < self.flat.numel():
                dist.all_reduce(self.flat[self.n_early:], op=dist.ReduceOp.SUM, group=group)
        else:
            self._armed = False        # armed but some early parameter got no gradient: reduce everything
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
