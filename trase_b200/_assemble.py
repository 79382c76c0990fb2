"""Fused parameter activation + rasterizer-input assembly (autograd wrapper of
trase_b200_assemble_forward/backward).  Equivalent to the torch glue of the
reference's render() (gaussian_renderer/__init__.py:82-121) with
scene/gaussian_model.py's activations, in one forward and one backward launch group."""
from __future__ import annotations

import torch

from . import _lib


def _c(t):
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    return t if t.data_ptr() % 16 == 0 else t.clone()


class _Assemble(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat):
        lib = _lib.load()
        dev = xyz.device
        P = xyz.shape[0]
        Mrest = f_rest.shape[1]
        with torch.cuda.device(dev):
            ins = [_c(t) for t in (xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat)]
            f32 = dict(dtype=torch.float32, device=dev)
            means3D = torch.empty((P, 3), **f32)
            scales = torch.empty((P, 3), **f32)
            rotations = torch.empty((P, 4), **f32)
            opac = torch.empty((P, 1), **f32)
            shs = torch.empty((P, Mrest + 1, 3), **f32)
            sh_objs = torch.empty((P, 1, 32), **f32)
            _lib.check(lib.trase_b200_assemble_forward(
                P, Mrest, *[t.data_ptr() for t in ins], means3D.data_ptr(), scales.data_ptr(), rotations.data_ptr(),
                opac.data_ptr(), shs.data_ptr(), sh_objs.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
        ctx.save_for_backward(ins[2], ins[4], ins[6], ins[9])
        ctx.Mrest = Mrest
        return means3D, scales, rotations, opac, shs, sh_objs

    @staticmethod
    def backward(ctx, g_means3D, g_scales, g_rotations, g_opac, g_shs, g_objs):
        lib = _lib.load()
        scaling, rotation, opacity, feat = ctx.saved_tensors
        dev = scaling.device
        P, Mrest = scaling.shape[0], ctx.Mrest
        with torch.cuda.device(dev):
            ups = [None if g is None else _c(g) for g in (g_means3D, g_scales, g_rotations, g_opac, g_shs, g_objs)]
            f32 = dict(dtype=torch.float32, device=dev)
            outs = [torch.empty((P, 3), **f32), torch.empty((P, 3), **f32), torch.empty((P, 3), **f32),
                    torch.empty((P, 3), **f32), torch.empty((P, 4), **f32), torch.empty((P, 4), **f32),
                    torch.empty((P, 1), **f32), torch.empty((P, 1, 3), **f32), torch.empty((P, Mrest, 3), **f32),
                    torch.empty((P, 1, 32), **f32)]
            _lib.check(lib.trase_b200_assemble_backward(
                P, Mrest, scaling.data_ptr(), rotation.data_ptr(), opacity.data_ptr(), feat.data_ptr(),
                *[None if u is None else u.data_ptr() for u in ups], *[o.data_ptr() for o in outs],
                torch.cuda.current_stream(dev).cuda_stream))
        return tuple(outs)


def assemble_inputs(xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat):
    """-> (means3D, scales, rotations, opacity, shs, sh_objs); all arguments CUDA fp32 tensors."""
    return _Assemble.apply(xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat)
