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
