"""CPU: the Python surface mirrors the reference's
(DGR/diff_gaussian_rasterization/__init__.py:166-232): names, field order,
argument checks and error behaviour; CPU tensors are rejected (no fallback)."""
import inspect

import pytest
import torch

from trase_b200.diff_gaussian_rasterization import (GaussianRasterizationSettings, GaussianRasterizer,
                                                    rasterize_gaussians)
from trase_b200 import synth


def _settings(W=32, H=32):
    sc = synth.make_scene(10, W, H, 1)
    return sc, synth.settings_from(sc, torch.zeros(3), GaussianRasterizationSettings)


def test_settings_field_order_is_the_reference_one():
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_forward_signature_is_the_reference_one():
    params = list(inspect.signature(GaussianRasterizer.forward).parameters)
    assert params == ["self", "means3D", "means2D", "opacities", "shs", "sh_objs", "colors_precomp", "scales",
                      "rotations", "cov3D_precomp"]
    assert list(inspect.signature(rasterize_gaussians).parameters) == [
        "means3D", "means2D", "sh", "sh_objs", "colors_precomp", "opacities", "scales", "rotations",
        "cov3Ds_precomp", "raster_settings"]
    assert hasattr(GaussianRasterizer, "markVisible")


def test_exactly_one_of_colour_inputs():
    sc, st = _settings()
    r = GaussianRasterizer(st)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(sc["means3D"], sc["means3D"], sc["opacities"], shs=None, colors_precomp=None, scales=sc["scales"],
          rotations=sc["rotations"], sh_objs=sc["sh_objs"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(sc["means3D"], sc["means3D"], sc["opacities"], shs=sc["shs"], colors_precomp=torch.zeros(10, 3),
          scales=sc["scales"], rotations=sc["rotations"], sh_objs=sc["sh_objs"])


def test_exactly_one_of_covariance_inputs():
    sc, st = _settings()
    r = GaussianRasterizer(st)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(sc["means3D"], sc["means3D"], sc["opacities"], shs=sc["shs"], sh_objs=sc["sh_objs"])
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(sc["means3D"], sc["means3D"], sc["opacities"], shs=sc["shs"], sh_objs=sc["sh_objs"], scales=sc["scales"],
          rotations=sc["rotations"], cov3D_precomp=torch.zeros(10, 6))


def test_cpu_tensors_are_rejected_not_emulated(native_lib):
    sc, st = _settings()
    r = GaussianRasterizer(st)
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        r(sc["means3D"], sc["means3D"], sc["opacities"], shs=sc["shs"], sh_objs=sc["sh_objs"], scales=sc["scales"],
          rotations=sc["rotations"])
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        r.markVisible(sc["means3D"])


def test_bad_means_shape_raises_like_the_reference(native_lib):
    sc, st = _settings()
    r = GaussianRasterizer(st)
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        r(torch.zeros(10, 4), sc["means3D"], sc["opacities"], shs=sc["shs"], sh_objs=sc["sh_objs"],
          scales=sc["scales"], rotations=sc["rotations"])


def test_synthetic_camera_matches_reference_conventions():
    # scene/cameras.py:76-79: full_proj = world_view^T-convention @ projection^T; campos = inverse(view)[3,:3]
    cam = synth.make_camera(1352, 1014)
    assert torch.allclose(cam["projmatrix"], cam["viewmatrix"] @ synth.projection_matrix(
        0.01, 100.0, 2 * torch.atan(torch.tensor(cam["tanfovx"])).item(),
        2 * torch.atan(torch.tensor(cam["tanfovy"])).item()).T)
    assert cam["projmatrix"][2, 3] == 1.0 and abs(cam["projmatrix"][0, 0].item() - 2.2) < 1e-5
    assert torch.equal(cam["campos"], torch.zeros(3))


def test_smooth_reverse_map_lists_every_entry_once():
    """build_smooth_reverse (the gather form of the smoothed-feature backward): row j of the reverse map holds exactly
    the (p, column) pairs with knn_idx[p][column] == j, encoded (p << 8 | column); every entry appears once."""
    import torch
    from trase_b200._assemble import build_smooth_reverse
    g = torch.Generator().manual_seed(3)
    P, K = 257, 16
    knn = torch.randint(0, P, (P, K), generator=g)
    knn[:, 0] = torch.arange(P)
    knn[10:40, 5] = 3                                     # a long reverse row
    ptr, ent = build_smooth_reverse(knn)
    assert ptr.dtype == torch.int32 and ent.dtype == torch.int32
    assert int(ptr[0]) == 0 and int(ptr[-1]) == P * K and bool((ptr[1:] >= ptr[:-1]).all())
    seen = set()
    for j in range(P):
        for e in ent[int(ptr[j]):int(ptr[j + 1])].tolist():
            p, c = e >> 8, e & 255
            assert int(knn[p, c]) == j
            seen.add((p, c))
    assert len(seen) == P * K
    assert build_smooth_reverse(torch.zeros((4, 65), dtype=torch.int64)) is None      # more columns than the kernel takes


def test_arena_slot_only_for_arena_parameters_in_good_standing():
    """_assemble._arena_slot: the direct-accumulation target exists only for a Parameter whose .grad is the arena's
    own slot; a replaced .grad, a tensor hook, a misaligned slot or the switch turned off all fall back to autograd."""
    import torch
    from trase_b200 import _assemble
    from trase_b200.distributed import GradArena
    a, b = torch.nn.Parameter(torch.zeros(5, 3)), torch.nn.Parameter(torch.zeros(5, 4))
    loose = torch.nn.Parameter(torch.zeros(5, 3))
    loose.grad = torch.zeros(5, 3)
    arena = GradArena([a, b], early=[a, b])
    dev = a.device
    assert _assemble._arena_slot(a, dev, 4) is a.grad and _assemble._arena_slot(b, dev, 4) is b.grad
    assert _assemble._arena_slot(loose, dev, 4) is None                     # not an arena parameter
    assert _assemble._arena_slot(torch.zeros(5, 3), dev, 4) is None         # not a Parameter at all
    assert _assemble._arena_slot(b, dev, 1 << 20) is None                   # alignment the kernel would need
    saved = _assemble.DIRECT_GRAD_ACCUM
    try:
        _assemble.DIRECT_GRAD_ACCUM = False
        assert _assemble._arena_slot(a, dev, 4) is None
    finally:
        _assemble.DIRECT_GRAD_ACCUM = saved
    a.register_hook(lambda g: g * 2)                                         # a tensor hook must see the gradient
    assert _assemble._arena_slot(a, dev, 4) is None
    b.grad = torch.zeros(5, 4)                                              # .grad replaced: no longer the arena's slot
    assert _assemble._arena_slot(b, dev, 4) is None
    assert arena.flat.numel() == 35
