"""Drop-in replacement for the reference's `diff_gaussian_rasterization` package.

Same public surface, argument order, return values and error behaviour as
DGR/diff_gaussian_rasterization/__init__.py (reference lines cited inline), on
top of the sm_100a kernels in libtrase_b200.so.  TRASE's
`gaussian_renderer/__init__.py:21` import works unchanged when this package is
importable as `diff_gaussian_rasterization`.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from .. import _raster


def cpu_deep_copy_tuple(input_tuple):
    # debug snapshots, as in the reference (…/__init__.py:20-22)
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


def rasterize_gaussians(means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    # …/__init__.py:24-47
    return _RasterizeGaussians.apply(means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    # …/__init__.py:49-164
    @staticmethod
    def forward(ctx, means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        args = (raster_settings.bg, means3D, colors_precomp, opacities, scales, rotations,
                raster_settings.scale_modifier, cov3Ds_precomp, raster_settings.viewmatrix,
                raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy,
                raster_settings.image_height, raster_settings.image_width, sh, sh_objs, raster_settings.sh_degree,
                raster_settings.campos, raster_settings.prefiltered, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # copy them before they can be corrupted
            try:
                res = _raster.rasterize_forward(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            res = _raster.rasterize_forward(*args)

        # outputs the loss never read arrive in backward as None instead of materialised zero images
        ctx.set_materialize_grads(False)
        ctx.raster_settings = raster_settings
        ctx.num_rendered = res.num_rendered
        ctx.capacity = res.capacity
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, res.radii, sh, sh_objs,
                              res.geom, res.binning, res.img)
        return res.color, res.radii, res.objects, res.depth

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_out_objects, grad_depth):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, sh_objs, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        H, W = raster_settings.image_height, raster_settings.image_width
        # Only what autograd will actually use is computed (the reference always computes all nine gradients): the
        # feature gradient is skipped while the feature parameter is frozen (GAUSSIAN state), and when ONLY the
        # feature parameter trains (FEATURE state) the dL/dalpha chain and the preprocess backward are skipped.
        need = ctx.needs_input_grad            # means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales, ...
        need_features = bool(need[3])
        need_geom = any(need[i] for i in (0, 1, 2, 4, 5, 6, 7, 8))
        if grad_out_color is None and grad_out_objects is None:
            return (None,) * 10

        args = (raster_settings.bg, means3D, radii, colors_precomp, scales, rotations,
                raster_settings.scale_modifier, cov3Ds_precomp, raster_settings.viewmatrix,
                raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy, grad_out_color, sh,
                grad_out_objects, sh_objs, raster_settings.sh_degree, raster_settings.campos, geomBuffer,
                num_rendered, binningBuffer, imgBuffer, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                grads9 = _raster.rasterize_backward(*args, capacity=ctx.capacity, need_geom=need_geom,
                                                    need_features=need_features)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            grads9 = _raster.rasterize_backward(*args, capacity=ctx.capacity, need_geom=need_geom,
                                                need_features=need_features)
        (grad_means2D, grad_colors_precomp, grad_objects, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = grads9

        # input order of forward (…/__init__.py:151-162)
        return (grad_means3D, grad_means2D, grad_sh, None if grad_objects is None else grad_objects.unsqueeze(1),
                grad_colors_precomp, grad_opacities, grad_scales, grad_rotations, grad_cov3Ds_precomp, None)


class GaussianRasterizationSettings(NamedTuple):
    # …/__init__.py:166-178 — field order is part of the API
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    # …/__init__.py:180-232
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            raster_settings = self.raster_settings
            visible = _raster.mark_visible(positions, raster_settings.viewmatrix, raster_settings.projmatrix)
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, sh_objs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        if sh_objs is None:
            sh_objs = torch.Tensor([])

        return rasterize_gaussians(means3D, means2D, shs, sh_objs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, raster_settings)
