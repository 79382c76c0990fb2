"""Fused parameter activation + rasterizer-input assembly (autograd wrapper of
trase_b200_assemble_forward/backward).  Equivalent to the torch glue of the
reference's render() (gaussian_renderer/__init__.py:82-121) with
scene/gaussian_model.py's activations, in one forward and one backward launch group."""
from __future__ import annotations

import ctypes as C

import torch

from . import _cuda, _lib


def _c(t):
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    return t if t.data_ptr() % 16 == 0 else t.clone()


def sh_block_of(f_dc, f_rest):
    """If f_dc [P,1,3] and f_rest [P,Mrest,3] are adjacent views of ONE contiguous [P,Mrest+1,3] block (as
    GaussianParams lays them out), return that block as a tensor (no copy); else None."""
    if f_dc.dim() != 3 or f_rest.dim() != 3 or f_dc.dtype != torch.float32 or f_rest.dtype != torch.float32:
        return None
    P, Mrest = f_rest.shape[0], f_rest.shape[1]
    row = 3 * (Mrest + 1)
    if P == 0 or tuple(f_dc.shape) != (P, 1, 3) or f_rest.shape[2] != 3 or f_dc.device != f_rest.device:
        return None
    if f_dc.stride(0) != row or f_dc.stride(2) != 1 or tuple(f_rest.stride()) != (row, 3, 1):
        return None
    if f_rest.data_ptr() != f_dc.data_ptr() + 12 or \
            f_dc.untyped_storage().data_ptr() != f_rest.untyped_storage().data_ptr():
        return None
    return torch.as_strided(f_dc.detach(), (P, Mrest + 1, 3), (row, 3, 1))


# Parameters whose `.grad` is a slot of a GradArena (trase_b200/distributed.py marks them) get their gradient ADDED
# straight into that slot by the backward kernels instead of being returned to autograd (which would launch one
# accumulation kernel per parameter, two of them on strided SH slices).  Autograd still runs the parameter's
# AccumulateGrad node with an undefined gradient, and post-accumulate hooks (the arena's own among them) still fire
# there — after the kernels that did the adding were queued — so they see the accumulated `.grad` as before.
# Opt-in through the arena; parameters with tensor (pre-accumulation) hooks, or models driven through
# torch.autograd.grad, take the ordinary path.
DIRECT_GRAD_ACCUM = True


def _arena_slot(p, dev, align):
    """The gradient tensor to add into, or None when the parameter is not an arena parameter in good standing."""
    if not (DIRECT_GRAD_ACCUM and isinstance(p, torch.nn.Parameter)):
        return None
    g = p.grad
    if (getattr(p, "_trase_arena_grad_ptr", None) is None or g is None or g.data_ptr() != p._trase_arena_grad_ptr or
            not g.is_contiguous() or g.dtype != torch.float32 or g.device != dev or g.data_ptr() % align or
            g.shape != p.shape or getattr(p, "_backward_hooks", None)):
        return None
    return g


class _Assemble(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat):
        lib = _lib.load()
        dev = xyz.device
        P = xyz.shape[0]
        Mrest = f_rest.shape[1]
        with _cuda.on_device(dev):
            block = sh_block_of(f_dc, f_rest)        # parameters already laid out as the rasterizer's SH input?
            ins = [_c(t) for t in (xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity)] + \
                ([None, None] if block is not None else [_c(f_dc), _c(f_rest)]) + [_c(feat)]
            f32 = dict(dtype=torch.float32, device=dev)
            means3D = torch.empty((P, 3), **f32)
            scales = torch.empty((P, 3), **f32)
            rotations = torch.empty((P, 4), **f32)
            opac = torch.empty((P, 1), **f32)
            shs = block if block is not None else torch.empty((P, Mrest + 1, 3), **f32)
            sh_objs = torch.empty((P, 1, 32), **f32)
            _lib.check(lib.trase_b200_assemble_forward(
                P, Mrest, *[None if t is None else t.data_ptr() for t in ins], means3D.data_ptr(), scales.data_ptr(),
                rotations.data_ptr(), opac.data_ptr(), None if block is not None else shs.data_ptr(), sh_objs.data_ptr(),
                _cuda.raw_stream(dev)))
        ctx.save_for_backward(ins[2], ins[4], ins[6], ins[9])
        ctx.leaves = (xyz, scaling, rotation, opacity, f_dc, f_rest)      # the objects themselves: backward looks at .grad
        ctx.Mrest = Mrest
        ctx.sh_aliased = block is not None
        # an output whose sources are all frozen is not differentiable: the rasterizer then knows (needs_input_grad)
        # that nobody wants that gradient and skips the work behind it (FEATURE state: everything but sh_objs)
        n = ctx.needs_input_grad
        frozen = [t for t, live in ((means3D, n[0] or n[1]), (scales, n[2] or n[3]), (rotations, n[4] or n[5]),
                                    (opac, n[6]), (shs, n[7] or n[8]), (sh_objs, n[9])) if not live]
        if frozen:
            ctx.mark_non_differentiable(*frozen)
        return means3D, scales, rotations, opac, shs, sh_objs

    @staticmethod
    def backward(ctx, g_means3D, g_scales, g_rotations, g_opac, g_shs, g_objs):
        lib = _lib.load()
        scaling, rotation, opacity, feat = ctx.saved_tensors
        dev = scaling.device
        P, Mrest = scaling.shape[0], ctx.Mrest
        with _cuda.on_device(dev):
            ups = [None if g is None else _c(g) for g in (g_means3D, g_scales, g_rotations, g_opac, g_shs, g_objs)]
            f32 = dict(dtype=torch.float32, device=dev)
            alias = ctx.sh_aliased
            need = ctx.needs_input_grad
            geom = any(need[0:7])            # FEATURE state: every geometry parameter and the deformation are frozen
            mk = (lambda *shape: torch.empty(shape, **f32)) if geom else (lambda *shape: None)
            outs = [mk(P, 3), mk(P, 3), mk(P, 3), mk(P, 3), mk(P, 4), mk(P, 4), mk(P, 1),
                    None if alias else torch.empty((P, 1, 3), **f32),
                    None if alias else torch.empty((P, Mrest, 3), **f32),
                    torch.empty((P, 1, 32), **f32) if need[9] else None]     # frozen feature: skipped
            # arena parameters: the kernels add into the parameter's own gradient slot (accumulate_mask)
            xyz_p, scaling_p, rotation_p, opacity_p, fdc_p, frest_p = ctx.leaves
            acc, direct = 0, []
            if geom:
                for bit, idx, leaf, align in ((1, 0, xyz_p, 4), (2, 2, scaling_p, 4), (4, 4, rotation_p, 16),
                                              (8, 6, opacity_p, 4)):
                    slot = _arena_slot(leaf, dev, align) if need[idx] else None
                    if slot is not None:
                        acc |= bit
                        outs[idx] = slot
                        direct.append((idx, leaf))
            sh_slots = None
            if need[7] and need[8] and ups[4] is not None and Mrest > 0:
                a, b = _arena_slot(fdc_p, dev, 4), _arena_slot(frest_p, dev, 4)
                if a is not None and b is not None:
                    sh_slots = (a, b)
                    acc |= 16
                    outs[7], outs[8] = a, b
                    direct += [(7, fdc_p), (8, frest_p)]
            _lib.check(lib.trase_b200_assemble_backward(
                P, Mrest, scaling.data_ptr(), rotation.data_ptr(), opacity.data_ptr(), feat.data_ptr(),
                *[None if u is None else u.data_ptr() for u in ups], *[None if o is None else o.data_ptr() for o in outs],
                acc, _cuda.raw_stream(dev)))
            if alias and sh_slots is None:
                # the SH gradient block is already laid out like the parameters: hand out its two slices
                g = ups[4] if ups[4] is not None else torch.zeros((P, Mrest + 1, 3), **f32)
                outs[7], outs[8] = g[:, :1], g[:, 1:]
            for idx, leaf in direct:
                outs[idx] = None                       # nothing left for autograd to add (its hooks still fire)
        return tuple(outs)


def assemble_inputs(xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat):
    """-> (means3D, scales, rotations, opacity, shs, sh_objs); all arguments CUDA fp32 tensors."""
    return _Assemble.apply(xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat)


def build_smooth_reverse(knn_idx):
    """Reverse of a [P,K] neighbour map for the gather form of the smoothed-feature backward: (rev_ptr [P+1] int32,
    rev_ent [P*K] int32) with the entries (p << 8 | column) of all (p, column) that name Gaussian j, row by row.
    Built once per map (a sort of P*K integers); None when the map does not fit the kernel's limits."""
    P, K = knn_idx.shape
    if K > 64 or P >= (1 << 23) or P == 0:
        return None
    flat = knn_idx.reshape(-1)
    order = torch.argsort(flat, stable=True)
    ent = ((torch.div(order, K, rounding_mode="floor") << 8) | (order % K)).to(torch.int32).contiguous()
    ptr = torch.zeros(P + 1, dtype=torch.int32, device=knn_idx.device)
    ptr[1:] = torch.cumsum(torch.bincount(flat, minlength=P), 0).to(torch.int32)
    return ptr, ent


class _SmoothFeatures(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat, knn_idx, sel_cols, rev):
        lib = _lib.load()
        if not feat.is_cuda:
            raise RuntimeError("trase_b200: smoothed features run on CUDA tensors only")
        dev = feat.device
        P, K = knn_idx.shape
        f = _c(feat).view(P, 32)
        idx = knn_idx.to(device=dev, dtype=torch.int64).contiguous()
        sel = (C.c_int * len(sel_cols))(*[int(c) for c in sel_cols])
        with _cuda.on_device(dev):
            out = torch.empty((P, 1, 32), dtype=torch.float32, device=dev)
            unit = torch.empty((P, 32), dtype=torch.float32, device=dev)     # every row normalised once, then gathered
            _lib.check(lib.trase_b200_smooth_features_forward(P, K, sel, len(sel_cols), f.data_ptr(), idx.data_ptr(),
                                                              out.data_ptr(), unit.data_ptr(),
                                                              _cuda.raw_stream(dev)))
        ctx.save_for_backward(f, idx)
        ctx.sel = sel
        ctx.shape = feat.shape
        # the gather form of the backward needs the reverse map and distinct columns (the reference's randperm draw)
        ctx.rev = rev if (rev is not None and len(set(int(c) for c in sel_cols)) == len(sel_cols)) else None
        return out

    @staticmethod
    def backward(ctx, g_out):
        lib = _lib.load()
        f, idx = ctx.saved_tensors
        P, K = idx.shape
        dev = f.device
        with _cuda.on_device(dev):
            g = _c(g_out)
            g_feat = torch.empty((P, 32), dtype=torch.float32, device=dev)
            s = _cuda.raw_stream(dev)
            if ctx.rev is not None:
                _lib.check(lib.trase_b200_smooth_features_backward_csr(
                    P, K, ctx.sel, len(ctx.sel), f.data_ptr(), ctx.rev[0].data_ptr(), ctx.rev[1].data_ptr(),
                    g.data_ptr(), g_feat.data_ptr(), s))
            else:
                _lib.check(lib.trase_b200_smooth_features_backward(P, K, ctx.sel, len(ctx.sel), f.data_ptr(),
                                                                   idx.data_ptr(), g.data_ptr(), g_feat.data_ptr(), s))
        return g_feat.view(ctx.shape), None, None, None


def smooth_features(feat, knn_idx, sel_cols, rev=None):
    """mean over the selected columns of knn_idx of normalize(feat[neighbour]) -> [P,1,32] (scene/gaussian_model.py:94-103
    runs normalize + index + mean + their autograd).  `rev` = build_smooth_reverse(knn_idx) turns the backward from an
    atomic scatter into a gather."""
    return _SmoothFeatures.apply(feat, knn_idx, sel_cols, rev)
