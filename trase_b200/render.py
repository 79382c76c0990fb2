"""Render glue: mirror of the reference's `render()`
(gaussian_renderer/__init__.py:37-155) on top of the B200 rasterizer, plus the
minimal parameter container it reads (the `get_*` properties of
scene/gaussian_model.py:181-205).  Same arguments, same output dict keys.
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import torch

from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer


class GaussianParams(torch.nn.Module):
    """Raw (pre-activation) per-Gaussian parameters with the reference's
    activations: exp for scaling, sigmoid for opacity, normalize for rotation
    (scene/gaussian_model.py:45-60, 181-205)."""

    def __init__(self, xyz, features_dc, features_rest, scaling, rotation, opacity, gaussian_features,
                 active_sh_degree=3, max_sh_degree=3):
        super().__init__()
        P = torch.nn.Parameter
        self._xyz = P(xyz)
        # The two SH parameters are adjacent views of ONE [P,16,3] block, i.e. already the rasterizer's `shs`
        # input: render() then needs no torch.cat copy per view and no split of dL/dSH on the way back
        # (trase_b200/_assemble.py sh_block_of).  Shapes, names and state_dict entries are the reference's;
        # if anything re-allocates them separately (Module.to, an optimizer that re-homes .data) the copying
        # path is used again automatically.
        block = torch.cat((features_dc, features_rest), dim=1).contiguous()
        self._features_dc = P(block[:, :1])         # [P,1,3]
        self._features_rest = P(block[:, 1:])       # [P,15,3]
        self._scaling = P(scaling)                  # log-scale
        self._rotation = P(rotation)
        self._opacity = P(opacity)                  # logit
        self._gaussian_features = P(gaussian_features)  # [P,1,32]
        self.active_sh_degree = active_sh_degree
        self.max_sh_degree = max_sh_degree
        self.feature_smooth_map = None              # {"K": K, "m": [P,K] int64 neighbour map}, as in the reference

    @classmethod
    def from_activated(cls, scene: dict, device) -> "GaussianParams":
        """From a synthetic scene holding activated values (trase_b200.synth)."""
        d = lambda t: t.to(device)
        op = scene["opacities"].clamp(1e-6, 1 - 1e-6)
        return cls(d(scene["means3D"]), d(scene["shs"][:, :1].contiguous()), d(scene["shs"][:, 1:].contiguous()),
                   d(torch.log(scene["scales"])), d(scene["rotations"]), d(torch.log(op / (1 - op))),
                   d(scene["sh_objs"]), active_sh_degree=scene["sh_degree"])

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_scaling(self):
        return torch.exp(self._scaling)

    @property
    def get_rotation(self):
        return torch.nn.functional.normalize(self._rotation)

    @property
    def get_opacity(self):
        return torch.sigmoid(self._opacity)

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    @property
    def get_gaussian_features(self):
        return self._gaussian_features


    def set_feature_smooth_map(self, knn_idx):
        """Install a precomputed [P,K] nearest-neighbour index map (the reference fills the same cache from
        pytorch3d.ops.knn_points on first use, scene/gaussian_model.py:85-92)."""
        self.feature_smooth_map = {"K": int(knn_idx.shape[1]), "m": knn_idx.to(self._xyz.device, torch.int64)}

    def get_smoothed_gaussian_features(self, K=16, dropout=0.5, select_point=None):
        """scene/gaussian_model.py:79-104.  Same arguments and result; `select_point` (optional) fixes the neighbour
        subset the reference draws with torch.randperm(K)[:int(K*dropout)], for reproducible tests.  The neighbour
        map is taken from `feature_smooth_map` (set_feature_smooth_map) or built once with a host k-d tree; the
        gather-normalise-mean runs as one CUDA kernel per direction on CUDA tensors."""
        if K <= 1:
            return self._gaussian_features
        assert dropout < 0 or int(K * dropout) >= 1
        if self.feature_smooth_map is None or self.feature_smooth_map["K"] != K:
            from .feature_utils import build_feature_smooth_map
            with torch.no_grad():
                self.set_feature_smooth_map(build_feature_smooth_map(self.get_xyz, K))
        m = self.feature_smooth_map["m"]
        if select_point is None:
            select_point = torch.randperm(K)[:int(K * dropout)] if 0 < dropout < 1 else torch.arange(K)
        cols = [int(c) for c in select_point]
        if self._gaussian_features.is_cuda and FUSED_GLUE:
            from ._assemble import build_smooth_reverse, smooth_features
            rev = None
            if self._gaussian_features.requires_grad and torch.is_grad_enabled():
                if "rev" not in self.feature_smooth_map:       # once per map: the reverse map of the backward gather
                    self.feature_smooth_map["rev"] = build_smooth_reverse(m)
                rev = self.feature_smooth_map["rev"]
            return smooth_features(self._gaussian_features, m, cols, rev)
        normed = torch.nn.functional.normalize(self._gaussian_features, dim=-1, p=2)
        return normed[m[:, cols], 0, :].mean(dim=1).unsqueeze(1)


def make_view(cam: dict, device) -> SimpleNamespace:
    """Camera object with the attributes `render()` reads (scene/cameras.py:27-79)."""
    return SimpleNamespace(FoVx=2 * math.atan(cam["tanfovx"]), FoVy=2 * math.atan(cam["tanfovy"]),
                           image_height=cam["H"], image_width=cam["W"],
                           world_view_transform=cam["viewmatrix"].to(device),
                           full_proj_transform=cam["projmatrix"].to(device), camera_center=cam["campos"].to(device))


FUSED_GLUE = True   # module switch: set False to run the reference's elementwise torch glue instead

PIPE_DEFAULT = SimpleNamespace(convert_SHs_python=False, compute_cov3D_python=False, debug=False)


def render(viewpoint_camera, pc, pipe, bg_color, d_xyz, d_rotation, d_scaling, is_6dof=False, scaling_modifier=1.0,
           override_color=None, mask=None, norm_gaussian_features=True, is_smooth_gaussian_features=False,
           smooth_K=16, want_viewspace_grad=True):
    """gaussian_renderer/__init__.py:37-155.  Background tensor must be on the GPU.

    `want_viewspace_grad` (extension, default = the reference's behaviour): the reference always makes
    `viewspace_points` a gradient leaf so that `.grad` can feed the densification statistics
    (train.py:362-366).  A caller that will not read it (FEATURE state past `densify_until_iter`, style transfer past
    it, evaluation) passes False; if then nothing but the feature parameter requires grad, the rasterizer's backward
    computes the feature gradient only."""
    if is_6dof or getattr(pipe, "convert_SHs_python", False) or getattr(pipe, "compute_cov3D_python", False):
        raise NotImplementedError("6-DoF deformation and the *_python pipeline switches are outside the "
                                  "accelerated path")
    if want_viewspace_grad:
        # a LEAF that requires grad: its .grad is kept like the reference's `zeros(..., requires_grad=True) + 0` followed
        # by retain_grad() (gaussian_renderer/__init__.py:40-44), without the extra add kernel
        screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True)
    else:
        screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center, prefiltered=False, debug=pipe.debug)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    means2D = screenspace_points
    fused = (FUSED_GLUE and override_color is None and norm_gaussian_features and isinstance(pc, GaussianParams)
             and all(torch.is_tensor(d) and d.shape[0] == pc._xyz.shape[0] for d in (d_xyz, d_rotation, d_scaling)))
    if fused:
        # one fused kernel group instead of ~20 elementwise torch kernels (and as many in autograd)
        from ._assemble import assemble_inputs
        # FEATURE state (train.py:274-275): the kNN-smoothed features take the place of the raw ones
        feats = pc.get_smoothed_gaussian_features(K=smooth_K, dropout=0.5) if is_smooth_gaussian_features \
            else pc._gaussian_features
        means3D, scales, rotations, opacity, shs, sh_objs = assemble_inputs(
            pc._xyz, d_xyz, pc._scaling, d_scaling, pc._rotation, d_rotation, pc._opacity, pc._features_dc,
            pc._features_rest, feats)
        colors_precomp = None
    else:
        means3D = pc.get_xyz + d_xyz
        opacity = pc.get_opacity
        scales = pc.get_scaling + d_scaling
        rotations = pc.get_rotation + d_rotation
        shs, colors_precomp = (pc.get_features, None) if override_color is None else (None, override_color)
        sh_objs = pc.get_smoothed_gaussian_features(K=smooth_K, dropout=0.5) if is_smooth_gaussian_features \
            else pc.get_gaussian_features
        if norm_gaussian_features:
            sh_objs = sh_objs / (sh_objs.norm(dim=2, keepdim=True) + 1e-9)
    if mask is not None:
        means3D, means2D, opacity, sh_objs = means3D[mask], means2D[mask], opacity[mask], sh_objs[mask]
        scales, rotations = scales[mask], rotations[mask]
        if colors_precomp is not None:
            colors_precomp = colors_precomp[mask]
        else:
            shs = shs[mask]
    rendered_image, radii, rendered_feats, depth = rasterizer(
        means3D=means3D, means2D=means2D, shs=shs, sh_objs=sh_objs, colors_precomp=colors_precomp,
        opacities=opacity, scales=scales, rotations=rotations, cov3D_precomp=None)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii, "render_gaussian_features": rendered_feats, "depth": depth}
