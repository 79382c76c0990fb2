"""Fused Adam over the flat arenas + densification statistics.

`ArenaAdam` has the semantics of the optimizers TRASE builds —
`torch.optim.Adam(param_groups, lr=0.0, eps=1e-15)` with one group per Gaussian
attribute (scene/gaussian_model.py:259-289) or one group for the DeformNetwork
(scene/deform_model.py:38-43), learning rates rewritten per iteration through
`optimizer.param_groups[i]["lr"]` (scene/gaussian_model.py:303-309) — but the
parameters, their gradients (the all-reduced `GradArena`) and both moments live in
flat fp32 buffers, and a step is ONE kernel instead of a few elementwise kernels per
tensor.  CUDA only.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, List, Sequence

import torch

from . import _cuda, _lib
from .distributed import GradArena


class ArenaAdam:
    def __init__(self, arena: GradArena, param_groups: Sequence[Dict], betas=(0.9, 0.999), eps: float = 1e-15):
        self.arena = arena
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.param_groups: List[Dict] = [dict(g) for g in param_groups]
        flat_g = arena.flat
        if not flat_g.is_cuda:
            raise RuntimeError("trase_b200: ArenaAdam runs on CUDA tensors only")
        owner = {}
        for gi, g in enumerate(self.param_groups):
            g.setdefault("lr", 0.0)
            g["params"] = list(g["params"])
            for p in g["params"]:
                owner[id(p)] = gi
        if len(arena.params) > 64:
            raise ValueError("ArenaAdam handles at most 64 parameter tensors per arena")
        # parameters move into one flat buffer laid out like the gradient arena
        self.flat_p = torch.empty_like(flat_g)
        self.exp_avg = torch.zeros_like(flat_g)
        self.exp_avg_sq = torch.zeros_like(flat_g)
        self._seg_group, self._seg_end = [], []
        for p, off in zip(arena.params, arena.offsets):
            n = p.numel()
            self.flat_p[off:off + n].copy_(p.detach().reshape(-1))
            p.data = self.flat_p[off:off + n].view_as(p)
            self._seg_group.append(owner.get(id(p)))       # None: in the arena but not optimised (frozen)
            self._seg_end.append(off + n)
        self._steps = [0] * len(self.param_groups)

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.arena.zero()

    @torch.no_grad()
    def step(self, groups: Sequence[str] = None) -> None:
        """One Adam update of every group (or of the groups named in `groups`: TRASE alternates its GAUSSIAN and
        FEATURE optimizers, train.py:362-372)."""
        lib = _lib.load()
        b1, b2 = self.betas
        active = [groups is None or g.get("name") in groups for g in self.param_groups]
        for gi, a in enumerate(active):
            if a:
                self._steps[gi] += 1
        nseg = len(self._seg_end)
        ends = (C.c_int64 * nseg)(*self._seg_end)
        step_size, bc2 = (C.c_float * nseg)(), (C.c_float * nseg)()
        for i, gi in enumerate(self._seg_group):
            if gi is None or not active[gi]:
                step_size[i], bc2[i] = 0.0, 0.0
            else:
                t = self._steps[gi]
                step_size[i] = float(self.param_groups[gi]["lr"]) / (1.0 - b1 ** t)
                bc2[i] = math.sqrt(1.0 - b2 ** t)
        dev = self.flat_p.device
        with _cuda.on_device(dev):
            _lib.check(lib.trase_b200_adam_step(self.flat_p.numel(), self.flat_p.data_ptr(), self.arena.flat.data_ptr(),
                                                self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), nseg, ends,
                                                step_size, bc2, b1, b2, self.eps,
                                                _cuda.raw_stream(dev)))


def add_densification_stats(xyz_gradient_accum: torch.Tensor, denom: torch.Tensor, viewspace_point_tensor: torch.Tensor,
                            update_filter: torch.Tensor) -> None:
    """GaussianModel.add_densification_stats (scene/gaussian_model.py:637-639) in one kernel:
    accum[filter] += ||viewspace.grad[filter, :2]||, denom[filter] += 1."""
    g = viewspace_point_tensor.grad if viewspace_point_tensor.grad is not None else viewspace_point_tensor
    if not g.is_cuda:
        raise RuntimeError("trase_b200: add_densification_stats runs on CUDA tensors only")
    lib = _lib.load()
    g = g.detach().contiguous().float()
    f = update_filter.contiguous()
    if f.dtype != torch.bool:
        f = f != 0
    P = g.shape[0]
    if not (xyz_gradient_accum.is_contiguous() and denom.is_contiguous() and xyz_gradient_accum.numel() == P
            and denom.numel() == P and g.shape[1] >= 2 and g.shape[1] == 3 and f.numel() == P):
        raise RuntimeError("trase_b200: add_densification_stats expects [P,3] gradients, [P] filter, [P,1] accumulators")
    with _cuda.on_device(g.device):
        _lib.check(lib.trase_b200_densification_stats(P, g.data_ptr(), f.data_ptr(), xyz_gradient_accum.data_ptr(),
                                                      denom.data_ptr(), _cuda.raw_stream(g.device)))
