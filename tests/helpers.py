"""Shared helpers for the parity tests (test infrastructure)."""
import numpy as np
import torch

GRAD_NAMES = ["means2D", "colors_precomp", "objects", "opacities", "means3D", "cov3D", "sh", "scales", "rotations"]


def oracle_forward_from_golden(g):
    from oracle import oracle
    pre = "in_colors_precomp" in g.files
    return oracle.forward(
        g["in_means3D"], g["in_opacities"], g["in_sh_objs"], g["in_viewmatrix"], g["in_projmatrix"], g["in_campos"],
        int(g["W"]), int(g["H"]), float(g["in_tanfovx"]), float(g["in_tanfovy"]),
        np.full(3, float(g["bg"]), np.float32), shs=None if pre else g["in_shs"], sh_degree=int(g["sh_degree"]),
        colors_precomp=g["in_colors_precomp"] if pre else None, scales=None if pre else g["in_scales"],
        rotations=None if pre else g["in_rotations"], cov3D_precomp=g["in_cov3D_precomp"] if pre else None)


def golden_inputs_cuda(g, dev):
    """(raw-args dict) for trase_b200._raster from a golden file."""
    pre = "in_colors_precomp" in g.files
    t = lambda k: torch.from_numpy(np.ascontiguousarray(g[k])).to(dev)
    e = torch.Tensor([])
    return dict(bg=torch.full((3,), float(g["bg"]), device=dev), means3D=t("in_means3D"),
                colors_precomp=t("in_colors_precomp") if pre else e, opacities=t("in_opacities"),
                scales=e if pre else t("in_scales"), rotations=e if pre else t("in_rotations"),
                cov3D_precomp=t("in_cov3D_precomp") if pre else e, viewmatrix=t("in_viewmatrix"),
                projmatrix=t("in_projmatrix"), tanfovx=float(g["in_tanfovx"]), tanfovy=float(g["in_tanfovy"]),
                H=int(g["H"]), W=int(g["W"]), sh=e if pre else t("in_shs"), sh_objs=t("in_sh_objs"),
                degree=int(g["sh_degree"]), campos=t("in_campos"))


def cuda_forward(a, debug=False, scale_modifier=1.0):
    from trase_b200 import _raster
    return _raster.rasterize_forward(a["bg"], a["means3D"], a["colors_precomp"], a["opacities"], a["scales"],
                                     a["rotations"], scale_modifier, a["cov3D_precomp"], a["viewmatrix"],
                                     a["projmatrix"], a["tanfovx"], a["tanfovy"], a["H"], a["W"], a["sh"],
                                     a["sh_objs"], a["degree"], a["campos"], False, debug)


def cuda_backward(a, res, dL_dcolor, dL_dobjects, scale_modifier=1.0):
    from trase_b200 import _raster
    out = _raster.rasterize_backward(a["bg"], a["means3D"], res.radii, a["colors_precomp"], a["scales"],
                                     a["rotations"], scale_modifier, a["cov3D_precomp"], a["viewmatrix"],
                                     a["projmatrix"], a["tanfovx"], a["tanfovy"], dL_dcolor, a["sh"], dL_dobjects,
                                     a["sh_objs"], a["degree"], a["campos"], res.geom, res.num_rendered, res.binning,
                                     res.img, False, capacity=res.capacity)
    return dict(zip(GRAD_NAMES, out))


def scene_args(sc, bgval, dev, colors_precomp=None, cov3D_precomp=None):
    cam = sc["cam"]
    e = torch.Tensor([])
    d = lambda x: x.to(dev)
    return dict(bg=torch.full((3,), float(bgval), device=dev), means3D=d(sc["means3D"]),
                colors_precomp=e if colors_precomp is None else d(colors_precomp), opacities=d(sc["opacities"]),
                scales=e if cov3D_precomp is not None else d(sc["scales"]),
                rotations=e if cov3D_precomp is not None else d(sc["rotations"]),
                cov3D_precomp=e if cov3D_precomp is None else d(cov3D_precomp), viewmatrix=d(cam["viewmatrix"]),
                projmatrix=d(cam["projmatrix"]), tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], H=cam["H"],
                W=cam["W"], sh=e if colors_precomp is not None else d(sc["shs"]), sh_objs=d(sc["sh_objs"]),
                degree=sc["sh_degree"], campos=d(cam["campos"]))


def views(res, a):
    from trase_b200 import _raster
    return _raster.workspace_views(res.geom, res.img, res.binning, a["means3D"].shape[0], a["W"], a["H"],
                                   res.capacity, res.num_rendered)


def rel_err(ref, got):
    ref = np.asarray(ref, np.float64)
    got = np.asarray(got, np.float64).reshape(ref.shape)
    if ref.size == 0:
        return 0.0
    return float(np.abs(ref - got).max() / max(np.abs(ref).max(), 1e-30))


def elementwise_grad_check(ref_runs, got, rtol=1e-3, k=16.0):
    """Per-element gradient check against a reference whose own result varies from run to run (it accumulates with
    fp32 atomics in scheduling order).  `ref_runs`: the same backward run several times.  An element passes when
        |got - mean_ref| <= rtol * |mean_ref| + k * spread_e + atol,
    spread_e = that element's own run-to-run half-range and atol = k * (99.9th percentile of spread_e over the
    tensor) — i.e. the absolute floor is the reference's measured atomics noise, not a fraction of the tensor's
    largest entry.  Returns (fraction of elements failing, atol / max|ref|, worst excess in units of the bound)."""
    runs = np.stack([np.asarray(r, np.float64).reshape(-1) for r in ref_runs])
    ref = runs.mean(0)
    got = np.asarray(got, np.float64).reshape(-1)
    if ref.size == 0:
        return 0.0, 0.0, 0.0
    spread = np.abs(runs - ref).max(0)
    atol = k * float(np.quantile(spread, 0.999))
    bound = rtol * np.abs(ref) + k * spread + atol
    excess = np.abs(got - ref) / np.maximum(bound, 1e-300)
    bad = excess > 1.0
    return float(bad.mean()), atol / max(float(np.abs(ref).max()), 1e-300), float(excess.max())
