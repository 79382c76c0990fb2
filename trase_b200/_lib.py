"""ctypes binding of the C ABI declared in include/trase_b200.h.

This is the whole Python<->native boundary: raw device pointers, sizes and a
CUDA stream handle.  torch only supplies memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtrase_b200.so")

c_float_p = C.c_void_p  # device pointers are passed as integers


class FwdArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("sh_degree", C.c_int), ("M", C.c_int), ("W", C.c_int), ("H", C.c_int),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int), ("debug", C.c_int),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p), ("sh_objs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
        ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p),
        ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
        ("out_color", C.c_void_p), ("out_objects", C.c_void_p), ("out_depth", C.c_void_p), ("radii", C.c_void_p),
        ("geom", C.c_void_p), ("geom_bytes", C.c_size_t),
        ("img", C.c_void_p), ("img_bytes", C.c_size_t),
        ("binning", C.c_void_p), ("binning_bytes", C.c_size_t),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("capacity", C.c_int64),
        ("num_rendered_host", C.c_void_p),
        ("stream", C.c_void_p),
        ("readback_event", C.c_void_p),
    ]


class BwdArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("sh_degree", C.c_int), ("M", C.c_int), ("W", C.c_int), ("H", C.c_int),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("debug", C.c_int),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p), ("sh_objs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
        ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("campos", C.c_void_p), ("radii", C.c_void_p),
        ("geom", C.c_void_p), ("img", C.c_void_p), ("binning", C.c_void_p), ("capacity", C.c_int64),
        ("dL_dout_color", C.c_void_p), ("dL_dout_objects", C.c_void_p),
        ("dL_dmeans2D", C.c_void_p), ("dL_dconic", C.c_void_p), ("dL_dopacity", C.c_void_p),
        ("dL_dcolors", C.c_void_p), ("dL_dobjects", C.c_void_p), ("dL_dmeans3D", C.c_void_p),
        ("dL_dcov3D", C.c_void_p), ("dL_dsh", C.c_void_p), ("dL_dscales", C.c_void_p),
        ("dL_drotations", C.c_void_p),
        ("stream", C.c_void_p),
    ]


GEOM_ARRAYS = ["depths", "means2D", "conic_opacity", "rgbd", "cov3D", "clamped", "tiles_touched",
               "point_offsets", "block_sums", "counters"]
IMG_ARRAYS = ["final_T", "n_contrib", "ranges"]
BIN_ARRAYS = ["point_list", "point_list_keys"]

# every symbol include/trase_b200.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "trase_b200_version", "trase_b200_last_error", "trase_b200_geom_bytes", "trase_b200_img_bytes",
    "trase_b200_binning_bytes", "trase_b200_scratch_bytes", "trase_b200_geom_offsets",
    "trase_b200_img_offsets", "trase_b200_binning_offsets", "trase_b200_rasterize_forward",
    "trase_b200_rasterize_forward_rebin", "trase_b200_rasterize_backward", "trase_b200_mark_visible",
    "trase_b200_sort_scratch_bytes", "trase_b200_sort_pairs", "trase_b200_higher_msb",
    "trase_b200_set_tuning", "trase_b200_rasterize_backward_stages", "trase_b200_profile_enable",
    "trase_b200_profile_read", "trase_b200_sort_pairs_static", "trase_b200_gemm_tf32", "trase_b200_wgrad_tf32", "trase_b200_deform_embed", "trase_b200_assemble_forward", "trase_b200_assemble_backward", "trase_b200_mlp_round_transpose", "trase_b200_mlp_pack_heads",
    "trase_b200_mlp_pack_dout", "trase_b200_mlp_heads_grad_accumulate", "trase_b200_launch_count", "trase_b200_loss_scratch_bytes",
    "trase_b200_l1_ssim_forward", "trase_b200_l1_ssim_backward", "trase_b200_feature_norm_scratch_bytes",
    "trase_b200_feature_norm_reg_forward", "trase_b200_feature_norm_reg_backward",
    "trase_b200_mlp_fused_pack_floats", "trase_b200_mlp_pack_fused", "trase_b200_mlp_forward_fused",
    "trase_b200_mlp_fused_pack_bwd_floats", "trase_b200_mlp_pack_fused_bwd", "trase_b200_mlp_backward_fused",
    "trase_b200_adam_step", "trase_b200_densification_stats", "trase_b200_sample_features_forward",
    "trase_b200_sample_features_backward", "trase_b200_mlp_fused_set_prof",
    "trase_b200_smooth_features_forward", "trase_b200_smooth_features_backward",
    "trase_b200_smooth_features_backward_csr",
    "trase_b200_mlp_forward_fused_bf16", "trase_b200_mlp_backward_fused_bf16", "trase_b200_wgrad_bf16", "trase_b200_wgrad_bf16_multi",
    "trase_b200_f32_to_bf16", "trase_b200_arena_gather_rows", "trase_b200_contrastive_scratch_bytes",
    "trase_b200_pixel_mask_info", "trase_b200_contrastive_soft_forward", "trase_b200_contrastive_soft_backward",
    "trase_b200_deform_embed_bf16", "trase_b200_mlp_pack_dout_bf16",
]

_lib = None


def load() -> C.CDLL:
    """Loads libtrase_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "trase_b200: native library %s is missing. Build it with `python -m trase_b200.build` "
            "(there is no CPU or PyTorch fallback for the rasterizer)." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.trase_b200_version.restype = C.c_int
    lib.trase_b200_last_error.restype = C.c_char_p
    for name in ("geom_bytes",):
        getattr(lib, "trase_b200_" + name).restype = C.c_size_t
        getattr(lib, "trase_b200_" + name).argtypes = [C.c_int]
    lib.trase_b200_img_bytes.restype = C.c_size_t
    lib.trase_b200_img_bytes.argtypes = [C.c_int, C.c_int]
    lib.trase_b200_binning_bytes.restype = C.c_size_t
    lib.trase_b200_binning_bytes.argtypes = [C.c_int64]
    lib.trase_b200_scratch_bytes.restype = C.c_size_t
    lib.trase_b200_scratch_bytes.argtypes = [C.c_int, C.c_int64]
    lib.trase_b200_geom_offsets.restype = None
    lib.trase_b200_geom_offsets.argtypes = [C.c_int, C.POINTER(C.c_size_t)]
    lib.trase_b200_img_offsets.restype = None
    lib.trase_b200_img_offsets.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_size_t)]
    lib.trase_b200_binning_offsets.restype = None
    lib.trase_b200_binning_offsets.argtypes = [C.c_int64, C.POINTER(C.c_size_t)]
    lib.trase_b200_rasterize_forward.restype = C.c_int
    lib.trase_b200_rasterize_forward.argtypes = [C.POINTER(FwdArgs)]
    lib.trase_b200_rasterize_forward_rebin.restype = C.c_int
    lib.trase_b200_rasterize_forward_rebin.argtypes = [C.POINTER(FwdArgs)]
    lib.trase_b200_rasterize_backward.restype = C.c_int
    lib.trase_b200_rasterize_backward.argtypes = [C.POINTER(BwdArgs)]
    lib.trase_b200_mark_visible.restype = C.c_int
    lib.trase_b200_mark_visible.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.trase_b200_sort_scratch_bytes.restype = C.c_size_t
    lib.trase_b200_sort_scratch_bytes.argtypes = [C.c_int64]
    lib.trase_b200_sort_pairs.restype = C.c_int
    lib.trase_b200_sort_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                          C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.trase_b200_sort_pairs_static.restype = C.c_int
    lib.trase_b200_sort_pairs_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                 C.c_int64, C.c_int, C.c_void_p]
    lib.trase_b200_smooth_features_forward.restype = C.c_int
    lib.trase_b200_smooth_features_forward.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p,
                                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.trase_b200_smooth_features_backward_csr.restype = C.c_int
    lib.trase_b200_smooth_features_backward_csr.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p,
                                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.trase_b200_smooth_features_backward.restype = C.c_int
    lib.trase_b200_smooth_features_backward.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p,
                                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.trase_b200_gemm_tf32.restype = C.c_int
    lib.trase_b200_gemm_tf32.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.trase_b200_wgrad_tf32.restype = C.c_int
    lib.trase_b200_wgrad_tf32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p, C.c_int, C.c_void_p]
    lib.trase_b200_deform_embed.restype = C.c_int
    lib.trase_b200_deform_embed.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                            C.c_int, C.c_void_p]
    lib.trase_b200_assemble_forward.restype = C.c_int
    lib.trase_b200_assemble_forward.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 17
    lib.trase_b200_assemble_backward.restype = C.c_int
    lib.trase_b200_assemble_backward.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 20 + [C.c_int, C.c_void_p]
    lib.trase_b200_mlp_round_transpose.restype = C.c_int
    lib.trase_b200_mlp_round_transpose.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                   C.c_void_p]
    lib.trase_b200_mlp_pack_heads.restype = C.c_int
    lib.trase_b200_mlp_pack_heads.argtypes = [C.c_void_p] * 6 + [C.c_int] + [C.c_void_p] * 4
    lib.trase_b200_mlp_pack_dout.restype = C.c_int
    lib.trase_b200_mlp_pack_dout.argtypes = [C.c_int] + [C.c_void_p] * 6
    lib.trase_b200_loss_scratch_bytes.restype = C.c_size_t
    lib.trase_b200_loss_scratch_bytes.argtypes = [C.c_int] * 3
    lib.trase_b200_l1_ssim_forward.restype = C.c_int
    lib.trase_b200_l1_ssim_forward.argtypes = [C.c_int] * 3 + [C.c_void_p] * 2 + [C.c_float] * 3 + [C.c_void_p] * 4
    lib.trase_b200_feature_norm_scratch_bytes.restype = C.c_size_t
    lib.trase_b200_feature_norm_scratch_bytes.argtypes = [C.c_int64]
    lib.trase_b200_feature_norm_reg_forward.restype = C.c_int
    lib.trase_b200_feature_norm_reg_forward.argtypes = [C.c_int, C.c_int64] + [C.c_void_p] * 5
    lib.trase_b200_feature_norm_reg_backward.restype = C.c_int
    lib.trase_b200_feature_norm_reg_backward.argtypes = [C.c_int, C.c_int64] + [C.c_void_p] * 6
    lib.trase_b200_mlp_fused_pack_floats.restype = C.c_size_t
    lib.trase_b200_mlp_fused_pack_floats.argtypes = [C.c_int] * 3
    lib.trase_b200_mlp_pack_fused.restype = C.c_int
    lib.trase_b200_mlp_pack_fused.argtypes = [C.c_int] * 3 + [C.POINTER(C.c_void_p)] + [C.c_void_p] * 3
    lib.trase_b200_mlp_fused_pack_bwd_floats.restype = C.c_size_t
    lib.trase_b200_mlp_fused_pack_bwd_floats.argtypes = [C.c_int]
    lib.trase_b200_mlp_pack_fused_bwd.restype = C.c_int
    lib.trase_b200_mlp_pack_fused_bwd.argtypes = [C.c_int] * 3 + [C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p]
    lib.trase_b200_mlp_forward_fused_bf16.restype = C.c_int
    lib.trase_b200_mlp_forward_fused_bf16.argtypes = [C.c_int] * 4 + [C.c_void_p, C.c_int] + [C.c_void_p] * 9
    lib.trase_b200_mlp_backward_fused_bf16.restype = C.c_int
    lib.trase_b200_mlp_backward_fused_bf16.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.POINTER(C.c_void_p),
                                                                                             C.c_void_p]
    lib.trase_b200_arena_gather_rows.restype = C.c_int
    lib.trase_b200_arena_gather_rows.argtypes = [C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int),
                                                 C.c_int64] + [C.c_void_p] * 9
    lib.trase_b200_contrastive_scratch_bytes.restype = C.c_size_t
    lib.trase_b200_contrastive_scratch_bytes.argtypes = [C.c_int]
    lib.trase_b200_pixel_mask_info.restype = C.c_int
    lib.trase_b200_pixel_mask_info.argtypes = [C.c_int, C.c_int, C.c_int64] + [C.c_void_p] * 7
    lib.trase_b200_contrastive_soft_forward.restype = C.c_int
    lib.trase_b200_contrastive_soft_forward.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_float, C.c_float] + \
        [C.c_void_p] * 3
    lib.trase_b200_contrastive_soft_backward.restype = C.c_int
    lib.trase_b200_contrastive_soft_backward.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_float, C.c_float] + \
        [C.c_void_p] * 5
    lib.trase_b200_deform_embed_bf16.restype = C.c_int
    lib.trase_b200_deform_embed_bf16.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                                 C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    lib.trase_b200_mlp_pack_dout_bf16.restype = C.c_int
    lib.trase_b200_mlp_pack_dout_bf16.argtypes = [C.c_int] + [C.c_void_p] * 7
    lib.trase_b200_mlp_heads_grad_accumulate.restype = C.c_int
    lib.trase_b200_mlp_heads_grad_accumulate.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 7 + [C.c_int,
                                                                                                         C.c_void_p]
    lib.trase_b200_wgrad_bf16_multi.restype = C.c_int
    lib.trase_b200_wgrad_bf16_multi.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.trase_b200_wgrad_bf16.restype = C.c_int
    lib.trase_b200_wgrad_bf16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p, C.c_int, C.c_void_p]
    lib.trase_b200_f32_to_bf16.restype = C.c_int
    lib.trase_b200_f32_to_bf16.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    lib.trase_b200_mlp_backward_fused.restype = C.c_int
    lib.trase_b200_mlp_backward_fused.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.POINTER(C.c_void_p), C.c_void_p]
    lib.trase_b200_sample_features_forward.restype = C.c_int
    lib.trase_b200_sample_features_forward.argtypes = [C.c_int] * 6 + [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 3
    lib.trase_b200_sample_features_backward.restype = C.c_int
    lib.trase_b200_sample_features_backward.argtypes = [C.c_int] * 6 + [C.c_void_p, C.c_int] + [C.c_void_p] * 5
    lib.trase_b200_adam_step.restype = C.c_int
    lib.trase_b200_adam_step.argtypes = [C.c_int64] + [C.c_void_p] * 4 + [C.c_int, C.POINTER(C.c_int64),
                                                                          C.POINTER(C.c_float), C.POINTER(C.c_float),
                                                                          C.c_float, C.c_float, C.c_float, C.c_void_p]
    lib.trase_b200_densification_stats.restype = C.c_int
    lib.trase_b200_densification_stats.argtypes = [C.c_int] + [C.c_void_p] * 5
    lib.trase_b200_mlp_forward_fused.restype = C.c_int
    lib.trase_b200_mlp_forward_fused.argtypes = [C.c_int] * 4 + [C.c_void_p, C.c_int] + [C.c_void_p] * 9
    lib.trase_b200_l1_ssim_backward.restype = C.c_int
    lib.trase_b200_l1_ssim_backward.argtypes = [C.c_int] * 3 + [C.c_void_p] * 3 + [C.c_float] * 2 + [C.c_void_p] * 3
    lib.trase_b200_launch_count.restype = C.c_ulonglong
    lib.trase_b200_higher_msb.restype = C.c_uint32
    lib.trase_b200_higher_msb.argtypes = [C.c_uint32]
    lib.trase_b200_rasterize_backward_stages.restype = C.c_int
    lib.trase_b200_rasterize_backward_stages.argtypes = [C.POINTER(BwdArgs), C.c_int]
    lib.trase_b200_profile_enable.restype = None
    lib.trase_b200_profile_enable.argtypes = [C.c_int]
    lib.trase_b200_profile_read.restype = C.c_int
    lib.trase_b200_profile_read.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_int)]
    lib.trase_b200_set_tuning.restype = None
    lib.trase_b200_set_tuning.argtypes = [C.c_char_p, C.c_int]
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError(load().trase_b200_last_error().decode() or "trase_b200: native call failed (%d)" % rc)


def geom_offsets(P: int) -> dict:
    arr = (C.c_size_t * len(GEOM_ARRAYS))()
    load().trase_b200_geom_offsets(P, arr)
    return dict(zip(GEOM_ARRAYS, list(arr)))


def img_offsets(W: int, H: int) -> dict:
    arr = (C.c_size_t * len(IMG_ARRAYS))()
    load().trase_b200_img_offsets(W, H, arr)
    return dict(zip(IMG_ARRAYS, list(arr)))


def binning_offsets(capacity: int) -> dict:
    arr = (C.c_size_t * len(BIN_ARRAYS))()
    load().trase_b200_binning_offsets(capacity, arr)
    return dict(zip(BIN_ARRAYS, list(arr)))


STAGES = ["preprocess_fwd", "scan_duplicate", "sort", "ranges", "render_fwd", "render_bwd", "preprocess_bwd", "other"]


def profile_enable(on: bool) -> None:
    load().trase_b200_profile_enable(int(on))


def profile_read() -> dict:
    ms = (C.c_double * 8)()
    cnt = (C.c_int * 8)()
    load().trase_b200_profile_read(ms, cnt)
    return {STAGES[i]: (ms[i], cnt[i]) for i in range(8)}


def launch_count() -> int:
    return int(load().trase_b200_launch_count())


def set_tuning(key: str, value: int) -> None:
    load().trase_b200_set_tuning(key.encode(), int(value))
