"""Feature-loss front end of the FEATURE training state, fused.

`get_features_correspondence_matrix` has the reference's name and result
(utils/feature_utils.py:50-55) but takes the FULL-resolution rendered feature image
plus the mask resolution: the bilinear resize of train.py:283 is evaluated only at
the sampled pixels (csrc/loss.cu), and the backward touches only the 4 x C entries
each sample read instead of writing whole gradient planes several times.
CUDA tensors only.
"""
from __future__ import annotations

import torch

from . import _lib


class _SampleFeatures(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat, idx, Hm, Wm, normalize):
        if not feat.is_cuda or feat.dim() != 3:
            raise RuntimeError("trase_b200: sample_resized_features expects a CUDA [C,H,W] feature image")
        lib = _lib.load()
        f = feat.detach().contiguous().float()
        idx = idx.to(device=f.device, dtype=torch.int64).contiguous()
        C, H, W = f.shape
        N = idx.numel()
        dev = f.device
        with torch.cuda.device(dev):
            out = torch.empty((N, C), dtype=torch.float32, device=dev)
            inv = torch.empty((N,), dtype=torch.float32, device=dev)
            _lib.check(lib.trase_b200_sample_features_forward(C, H, W, Hm, Wm, N, f.data_ptr(), idx.data_ptr(),
                                                              int(normalize), out.data_ptr(), inv.data_ptr(),
                                                              torch.cuda.current_stream(dev).cuda_stream))
        ctx.save_for_backward(idx, out, inv)
        ctx.meta = (C, H, W, Hm, Wm, N, int(normalize))
        return out

    @staticmethod
    def backward(ctx, g_out):
        idx, out, inv = ctx.saved_tensors
        C, H, W, Hm, Wm, N, normalize = ctx.meta
        lib = _lib.load()
        dev = out.device
        with torch.cuda.device(dev):
            d = torch.zeros((C, H, W), dtype=torch.float32, device=dev)
            g = g_out.detach().contiguous().float()
            _lib.check(lib.trase_b200_sample_features_backward(C, H, W, Hm, Wm, N, idx.data_ptr(), normalize,
                                                               out.data_ptr(), inv.data_ptr(), g.data_ptr(), d.data_ptr(),
                                                               torch.cuda.current_stream(dev).cuda_stream))
        return d, None, None, None, None


def sample_resized_features(rendered_features, out_hw, sampled_pixel, normalize=True):
    """rows of normalize(F.interpolate(rendered_features[None], out_hw, mode='bilinear')[0][:, sampled_pixel].T):
    [N, C] for the N True entries of the boolean [Hm, Wm] mask `sampled_pixel` (row-major order, as boolean
    indexing yields them)."""
    Hm, Wm = int(out_hw[0]), int(out_hw[1])
    if tuple(sampled_pixel.shape) != (Hm, Wm):
        raise RuntimeError("trase_b200: sampled_pixel must have the mask resolution")
    idx = sampled_pixel.reshape(-1).nonzero().reshape(-1)       # one host sync, like the reference's boolean index
    return _SampleFeatures.apply(rendered_features, idx, Hm, Wm, bool(normalize))


def get_features_correspondence_matrix(rendered_features, sampled_pixel, out_hw=None):
    """utils/feature_utils.py:50-55 with train.py:283's resize folded in: cosine similarities of the sampled
    pixels' (resized) features, [N, N]."""
    out_hw = tuple(sampled_pixel.shape[-2:]) if out_hw is None else out_hw
    f = sample_resized_features(rendered_features, out_hw, sampled_pixel, normalize=True)
    return f @ f.t()
