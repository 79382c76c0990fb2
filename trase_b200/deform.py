"""DeformNetwork — host-side mirror of the reference's time-conditioned
deformation MLP (utils/time_utils.py:60-131, wrapped by scene/deform_model.py:26-35).

Same constructor arguments, parameter names (`linear.{i}`, `gaussian_warp`,
`gaussian_rotation`, `gaussian_scaling`, `timenet.{0,2}`) and forward
signature, so a reference `deform.pth` state_dict loads unchanged.  Only the
default non-6DoF variant is on the hot path (arguments/__init__.py:70,124).

Backends:
  "torch"  — eager fp32 (what the reference runs; cuBLAS SGEMM on CUDA cores)
  "native" — fused sm_100a kernel chain (positional encoding + 8 layers + heads);
             raises if the kernels are not built.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def positional_encoding(x: torch.Tensor, num_freqs: int) -> torch.Tensor:
    """[x, sin(2^0 x), cos(2^0 x), ..., sin(2^(L-1) x), cos(2^(L-1) x)] along the last dim
    (utils/time_utils.py:27-57 with log-sampled bands and include_input)."""
    outs = [x]
    for k in range(num_freqs):
        f = 2.0 ** k
        outs.append(torch.sin(x * f))
        outs.append(torch.cos(x * f))
    return torch.cat(outs, -1)


class DeformNetwork(nn.Module):
    def __init__(self, D=8, W=256, input_ch=3, output_ch=59, multires=10, is_blender=False, is_6dof=False,
                 backend: str = "torch"):
        super().__init__()
        if is_6dof:
            raise NotImplementedError("the 6-DoF head (exp_se3) is outside the accelerated path")
        self.D, self.W = D, W
        self.output_ch = output_ch
        self.multires = multires
        self.t_multires = 6 if is_blender else 10
        self.skips = [D // 2]
        time_input_ch = 1 + 2 * self.t_multires
        xyz_input_ch = 3 + 6 * multires
        self.input_ch = xyz_input_ch + time_input_ch
        self.is_blender, self.is_6dof = is_blender, is_6dof
        if is_blender:
            self.time_out = 30
            self.timenet = nn.Sequential(nn.Linear(time_input_ch, 256), nn.ReLU(inplace=True),
                                         nn.Linear(256, self.time_out))
            first = xyz_input_ch + self.time_out
        else:
            first = self.input_ch
        self.skip_in = first
        self.linear = nn.ModuleList([nn.Linear(first, W)] + [
            nn.Linear(W, W) if i not in self.skips else nn.Linear(W + first, W) for i in range(D - 1)])
        self.gaussian_warp = nn.Linear(W, 3)
        self.gaussian_rotation = nn.Linear(W, 4)
        self.gaussian_scaling = nn.Linear(W, 3)
        self.backend = backend

    def forward(self, x, t):
        if self.backend == "native":
            from . import _deform_native
            return _deform_native.deform_forward(self, x, t)
        t_emb = positional_encoding(t, self.t_multires)
        if self.is_blender:
            t_emb = self.timenet(t_emb)
        x_emb = positional_encoding(x, self.multires)
        h = torch.cat([x_emb, t_emb], dim=-1)
        for i, _ in enumerate(self.linear):
            h = F.relu(self.linear[i](h))
            if i in self.skips:
                h = torch.cat([x_emb, t_emb, h], -1)
        return self.gaussian_warp(h), self.gaussian_rotation(h), self.gaussian_scaling(h)


class DeformModel:
    """scene/deform_model.py:26-35 (`step` only; optimiser and checkpoint
    plumbing stay with the caller)."""

    def __init__(self, is_blender=False, is_6dof=False, model_type="DeformNetwork", backend="torch", device="cuda"):
        if model_type != "DeformNetwork":
            raise NotImplementedError("only the default DeformNetwork is on the accelerated path")
        self.deform = DeformNetwork(is_blender=is_blender, is_6dof=is_6dof, backend=backend).to(device)
        self.model_type = model_type

    def step(self, xyz, time_emb, f=None):
        return self.deform(xyz, time_emb)
