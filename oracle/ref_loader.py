"""TEST INFRASTRUCTURE — not part of the product.  Only tests/, bench.py's
reference/cpu_baseline legs and __graft_entry__.smoke() may import this.

Loads the UNMODIFIED reference rasterizer built by oracle/build_ref.sh into
oracle/_ref/ (git-ignored, travels to the GPU box) and exposes helpers to read
the reference's three byte workspaces with the `fromChunk` layouts
(DGR/cuda_rasterizer/rasterizer_impl.cu:158-197, 128-byte aligned arrays).
"""
from __future__ import annotations

import importlib.util
import os
import sys

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
_mod = None


def available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "diff_gaussian_rasterization", "_C.so"))


def load():
    """Returns the reference's `diff_gaussian_rasterization` module (its own
    __init__.py + compiled _C) under the private name `_trase_ref_dgr`."""
    global _mod
    if _mod is not None:
        return _mod
    if not available():
        raise RuntimeError("oracle/_ref is not built; run oracle/build_ref.sh where /root/reference exists")
    pkg_dir = os.path.join(REF_DIR, "diff_gaussian_rasterization")
    name = "_trase_ref_dgr"
    spec = importlib.util.spec_from_file_location(name, os.path.join(pkg_dir, "__init__.py"),
                                                  submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    _mod = mod
    return mod


_utils = None


def utils_available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "trase_ref", "utils", "time_utils.py"))


def load_utils():
    """The reference's own `utils` package (time_utils.DeformNetwork, loss_utils, feature_utils), installed
    unmodified by oracle/build_ref.sh.  loss_utils.py imports pytorch3d and torchvision at module level for helpers
    this path never calls; neither ships in this image, so empty stand-in modules are registered for exactly those
    imports.  Returns a namespace with .time_utils, .loss_utils, .feature_utils."""
    global _utils
    if _utils is not None:
        return _utils
    if not utils_available():
        raise RuntimeError("oracle/_ref/trase_ref is not installed; run oracle/build_ref.sh where /root/reference exists")
    import importlib
    import types
    for name in ("pytorch3d", "pytorch3d.ops", "torchvision", "torchvision.utils"):
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                m = types.ModuleType(name)
                m.save_image = None
                sys.modules[name] = m
    if not hasattr(sys.modules["pytorch3d"], "ops"):
        sys.modules["pytorch3d"].ops = sys.modules["pytorch3d.ops"]
    root = os.path.join(REF_DIR, "trase_ref")
    if "utils" in sys.modules and not getattr(sys.modules["utils"], "__file__", "").startswith(root):
        raise RuntimeError("another top-level `utils` package is already imported")
    sys.path.insert(0, root)
    try:
        ns = types.SimpleNamespace(time_utils=importlib.import_module("utils.time_utils"),
                                   loss_utils=importlib.import_module("utils.loss_utils"),
                                   feature_utils=importlib.import_module("utils.feature_utils"))
    finally:
        sys.path.remove(root)
    _utils = ns
    return ns


def _align(x: int, a: int = 128) -> int:
    return (x + a - 1) // a * a


def parse_geom(buf: torch.Tensor, P: int) -> dict:
    """GeometryState::fromChunk up to tiles_touched (the CUB scan temp that
    follows has a library-defined size, so point_offsets is not located)."""
    base = buf.data_ptr()
    cur = base
    out = {}

    def take(name, count, dtype, shape=None):
        nonlocal cur
        cur = _align(cur)
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        v = buf[cur - base:cur - base + nbytes].view(dtype)
        out[name] = v.view(*shape) if shape else v
        cur += nbytes

    take("depths", P, torch.float32)
    take("clamped", 3 * P, torch.uint8, (P, 3))
    take("internal_radii", P, torch.int32)
    take("means2D", 2 * P, torch.float32, (P, 2))
    take("cov3D", 6 * P, torch.float32, (P, 6))
    take("conic_opacity", 4 * P, torch.float32, (P, 4))
    take("rgb", 3 * P, torch.float32, (P, 3))
    take("tiles_touched", P, torch.int32)
    return out


def parse_img(buf: torch.Tensor, N: int, T: int) -> dict:
    base = buf.data_ptr()
    cur = base
    out = {}

    def take(name, count, dtype, shape=None):
        nonlocal cur
        cur = _align(cur)
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        v = buf[cur - base:cur - base + nbytes].view(dtype)
        out[name] = v.view(*shape) if shape else v
        cur += nbytes

    take("accum_alpha", N, torch.float32)
    take("n_contrib", N, torch.int32)
    take("ranges_all", 2 * N, torch.int32, (N, 2))
    out["ranges"] = out["ranges_all"][:T]
    return out


def parse_binning(buf: torch.Tensor, R: int) -> dict:
    base = buf.data_ptr()
    cur = base
    out = {}

    def take(name, count, dtype):
        nonlocal cur
        cur = _align(cur)
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        out[name] = buf[cur - base:cur - base + nbytes].view(dtype)
        cur += nbytes

    take("point_list", R, torch.int32)
    take("point_list_unsorted", R, torch.int32)
    take("point_list_keys", R, torch.int64)
    take("point_list_keys_unsorted", R, torch.int64)
    return out


def run_forward(scene: dict, bg: torch.Tensor, colors_precomp=None, cov3D_precomp=None, scale_modifier=1.0):
    """Calls the reference's `_C.rasterize_gaussians` directly (same args the
    reference's autograd wrapper passes, …/__init__.py:68-89) and returns all
    eight results as a dict plus parsed workspaces."""
    ref = load()
    cam = scene["cam"]
    empty = torch.Tensor([])
    use_sh = colors_precomp is None
    args = (bg, scene["means3D"], empty if use_sh else colors_precomp, scene["opacities"],
            empty if cov3D_precomp is not None else scene["scales"],
            empty if cov3D_precomp is not None else scene["rotations"],
            scale_modifier, empty if cov3D_precomp is None else cov3D_precomp,
            cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cam["H"], cam["W"],
            scene["shs"] if use_sh else empty, scene["sh_objs"], scene["sh_degree"], cam["campos"], False, False)
    R, color, objects, depth, radii, geom, binning, img = ref._C.rasterize_gaussians(*args)
    P = scene["means3D"].shape[0]
    W, H = cam["W"], cam["H"]
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out = dict(num_rendered=R, color=color, objects=objects, depth=depth, radii=radii, geomBuffer=geom,
               binningBuffer=binning, imgBuffer=img, args=args)
    if P > 0:
        out.update(parse_geom(geom, P))
        out.update(parse_img(img, W * H, T))
        out.update(parse_binning(binning, R))
    return out


def run_backward(scene: dict, bg: torch.Tensor, fwd: dict, dL_dcolor, dL_dobjects, colors_precomp=None,
                 cov3D_precomp=None, scale_modifier=1.0):
    """Calls `_C.rasterize_gaussians_backward` (…/__init__.py:116-138) and
    returns the nine gradients by name."""
    ref = load()
    cam = scene["cam"]
    empty = torch.Tensor([])
    use_sh = colors_precomp is None
    args = (bg, scene["means3D"], fwd["radii"], empty if use_sh else colors_precomp,
            empty if cov3D_precomp is not None else scene["scales"],
            empty if cov3D_precomp is not None else scene["rotations"], scale_modifier,
            empty if cov3D_precomp is None else cov3D_precomp, cam["viewmatrix"], cam["projmatrix"],
            cam["tanfovx"], cam["tanfovy"], dL_dcolor, scene["shs"] if use_sh else empty, dL_dobjects,
            scene["sh_objs"], scene["sh_degree"], cam["campos"], fwd["geomBuffer"], fwd["num_rendered"],
            fwd["binningBuffer"], fwd["imgBuffer"], False)
    names = ["means2D", "colors_precomp", "objects", "opacities", "means3D", "cov3D", "sh", "scales", "rotations"]
    return dict(zip(names, ref._C.rasterize_gaussians_backward(*args)))
