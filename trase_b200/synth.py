"""Synthetic scenes and cameras for tests and benchmarks.

Everything is generated on the CPU with a seeded torch.Generator so that the
same bits reach the reference, the oracle and the CUDA path; move the result
with `scene_to(scene, device)`.
"""
from __future__ import annotations

import math
from typing import Dict

import torch

CONFIGS = {
    # name: (P, W, H)
    "tiny": (20_000, 128, 128),          # BASELINE.json configs[0]
    "neu3d": (300_000, 1352, 1014),      # configs[1]
    "hypernerf": (200_000, 536, 960),    # configs[2]
    "inference2m": (2_000_000, 1920, 1080),  # configs[3]
}


def projection_matrix(znear: float, zfar: float, fovx: float, fovy: float) -> torch.Tensor:
    """Perspective matrix with the conventions of utils/graphics_utils.py:51-71
    (z_sign = +1, depth mapped to [0, 1])."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def make_camera(W: int, H: int, world_view: torch.Tensor | None = None) -> Dict[str, object]:
    """Camera dict in the row-vector (transposed) convention of scene/cameras.py:76-79."""
    f = 1.1 * W
    tanfovx, tanfovy = W / (2 * f), H / (2 * f)
    fovx, fovy = 2 * math.atan(tanfovx), 2 * math.atan(tanfovy)
    wv = torch.eye(4) if world_view is None else world_view.clone().float()  # already transposed
    proj = projection_matrix(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = wv @ proj
    campos = torch.linalg.inv(wv)[3, :3].contiguous()
    return dict(W=W, H=H, tanfovx=tanfovx, tanfovy=tanfovy, viewmatrix=wv.contiguous(),
                projmatrix=full.contiguous(), campos=campos, focal=f)


def orbit_world_view(yaw, pitch, centre=(0.0, 0.0, 4.75)):
    """World->view in the row-vector convention: the camera orbits the scene centre by (yaw, pitch) radians."""
    cy, sy, cp, sp = math.cos(yaw), math.sin(yaw), math.cos(pitch), math.sin(pitch)
    Ry = torch.tensor([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rx = torch.tensor([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    Rm = (Rx @ Ry).float()
    c = torch.tensor(centre)
    wv = torch.eye(4)
    wv[:3, :3] = Rm.T
    wv[3, :3] = c - c @ Rm.T
    return wv


def make_scene(P: int, W: int, H: int, seed: int = 1, sh_degree: int = 3, target_radius_px: float = 10.0
               ) -> Dict[str, object]:
    g = torch.Generator().manual_seed(seed)
    cam = make_camera(W, H)
    tanx, tany, f = cam["tanfovx"], cam["tanfovy"], cam["focal"]

    def U(n, lo, hi):
        return torch.rand(n, generator=g) * (hi - lo) + lo

    def N(*shape):
        return torch.randn(*shape, generator=g)

    n_far = int(round(0.98 * P))
    z = torch.cat([U(n_far, 1.5, 8.0), U(P - n_far, -1.0, 0.2)])
    z = z[torch.randperm(P, generator=g)]
    x = U(P, -1.15, 1.15) * tanx * z
    y = U(P, -1.15, 1.15) * tany * z
    means3D = torch.stack([x, y, z], 1).contiguous()

    s0 = (target_radius_px / 3.0) * 4.75 / f
    # transcendental / reduction steps run in float64 and are narrowed once, so that the fp32 scene does not
    # depend on how the CPU vectorises (thread count, SIMD tails)
    scales = torch.exp(math.log(s0) + 0.6 * N(P, 3).double()).float().contiguous()
    q = N(P, 4)
    qd = q.double()
    rotations = ((qd / qd.norm(dim=1, keepdim=True)).float() + 0.01 * N(P, 4)).contiguous()  # deliberately non-unit
    opacities = torch.sigmoid(2.0 * N(P, 1).double()).float().contiguous()
    shs = torch.cat([N(P, 1, 3), 0.1 * N(P, 15, 3)], 1).contiguous()
    feats = N(P, 1, 32)
    fd = feats.double()
    sh_objs = (fd / (fd.norm(dim=2, keepdim=True) + 1e-9)).float().contiguous()
    return dict(P=P, sh_degree=sh_degree, means3D=means3D, scales=scales, rotations=rotations,
                opacities=opacities, shs=shs, sh_objs=sh_objs, cam=cam, seed=seed)


def make_upstream_grads(W: int, H: int, seed: int = 1) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(10_000 + seed)
    n = W * H
    return dict(dL_dcolor=(torch.randn(3, H, W, generator=g) / n).contiguous(),
                dL_dobjects=(torch.randn(32, H, W, generator=g) / n).contiguous(),
                dL_ddepth=torch.randn(1, H, W, generator=g).contiguous())


def scene_to(scene: Dict[str, object], device) -> Dict[str, object]:
    out = {}
    for k, v in scene.items():
        if isinstance(v, torch.Tensor):
            out[k] = v.to(device)
        elif isinstance(v, dict):
            out[k] = scene_to(v, device)
        else:
            out[k] = v
    return out


def settings_from(scene: Dict[str, object], bg: torch.Tensor, settings_cls, debug: bool = False,
                  scale_modifier: float = 1.0):
    cam = scene["cam"]
    return settings_cls(image_height=cam["H"], image_width=cam["W"], tanfovx=cam["tanfovx"],
                        tanfovy=cam["tanfovy"], bg=bg, scale_modifier=scale_modifier,
                        viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"],
                        sh_degree=scene["sh_degree"], campos=cam["campos"], prefiltered=False, debug=debug)
