"""Image losses of the GAUSSIAN training state, fused on the GPU
(SURVEY.md §8(f) rank 2).  Same names and arguments as the reference's
`utils/loss_utils.py`: `l1_loss` (:30-31) and `ssim` (:61-86), plus
`l1_ssim_loss`, the combination train.py:235-243 builds from them,
`(1 - lambda_dssim) * L1 + lambda_dssim * (1 - SSIM)`, in one pass.

One CUDA kernel per direction (trase_b200/csrc/loss.cu) replaces the five
depthwise 11x11 convolutions and the elementwise kernels around them; the
backward writes the image gradient directly.  CUDA tensors only, no fallback.
"""
from __future__ import annotations

import torch

from . import _cuda, _lib


def _stream(dev):
    return _cuda.raw_stream(dev)


class _L1SSIM(torch.autograd.Function):
    """value = w_l1 * mean|x - y| + w_ssim * mean(SSIM(x, y)) + bias."""

    @staticmethod
    def forward(ctx, img, gt, w_l1, w_ssim, bias):
        if not img.is_cuda or not gt.is_cuda:
            raise RuntimeError("trase_b200: the fused image losses run on CUDA tensors only")
        if img.shape != gt.shape or img.dim() not in (3, 4):
            raise RuntimeError("trase_b200: image and target must share a [C,H,W] or [B,C,H,W] shape")
        lib = _lib.load()
        x = img.detach().contiguous().float()
        y = gt.detach().contiguous().float()
        H, W = x.shape[-2:]
        C = x.numel() // (H * W)
        dev = x.device
        with _cuda.on_device(dev):
            need_maps = w_ssim != 0.0 and ctx.needs_input_grad[0]
            maps = torch.empty((3, C, H, W), dtype=torch.float32, device=dev) if need_maps else None
            scratch = torch.empty(lib.trase_b200_loss_scratch_bytes(C, H, W) // 4, dtype=torch.float32, device=dev)
            out3 = torch.empty(3, dtype=torch.float32, device=dev)
            _lib.check(lib.trase_b200_l1_ssim_forward(C, H, W, x.data_ptr(), y.data_ptr(), w_l1, w_ssim, bias,
                                                      maps.data_ptr() if maps is not None else None,
                                                      scratch.data_ptr(), out3.data_ptr(), _stream(dev)))
        ctx.save_for_backward(x, y, maps)
        ctx.meta = (C, H, W, w_l1, w_ssim, img.shape)
        return out3                                   # {L1, SSIM, w_l1 * L1 + w_ssim * SSIM + bias}

    @staticmethod
    def backward(ctx, g_out3):
        x, y, maps = ctx.saved_tensors
        C, H, W, w_l1, w_ssim, shape = ctx.meta
        lib = _lib.load()
        dev = x.device
        with _cuda.on_device(dev):
            d_img = torch.empty_like(x)
            go = g_out3.detach().contiguous().float()
            _lib.check(lib.trase_b200_l1_ssim_backward(C, H, W, x.data_ptr(), y.data_ptr(),
                                                       maps.data_ptr() if maps is not None else None, w_l1, w_ssim,
                                                       go.data_ptr(), d_img.data_ptr(), _stream(dev)))
        return d_img.view(shape), None, None, None, None


def l1_loss(network_output, gt):
    """utils/loss_utils.py:30-31."""
    return _L1SSIM.apply(network_output, gt, 1.0, 0.0, 0.0)[0]


def ssim(img1, img2, window_size=11, size_average=True):
    """utils/loss_utils.py:61-86 (window 11, sigma 1.5, mean over all pixels)."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("the fused SSIM covers the reference's only call pattern: window 11, size_average")
    return _L1SSIM.apply(img1, img2, 0.0, 1.0, 0.0)[1]


def l1_ssim_loss(image, gt_image, lambda_dssim=0.2, return_terms=False):
    """train.py:235-243: (1 - lambda) * l1_loss + lambda * (1 - ssim), both terms from one pass.
    With return_terms also returns (Ll1, ssim) as detached device scalars (train.py logs Ll1)."""
    lam = float(lambda_dssim)
    out3 = _L1SSIM.apply(image, gt_image, 1.0 - lam, -lam, lam)
    return (out3[2], out3[0].detach(), out3[1].detach()) if return_terms else out3[2]


class _FeatureNormReg(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat):
        if not feat.is_cuda:
            raise RuntimeError("trase_b200: the fused feature-norm regulariser runs on CUDA tensors only")
        if feat.dim() != 3:
            raise RuntimeError("trase_b200: rendered features must be [C,H,W]")
        lib = _lib.load()
        f = feat.detach().contiguous().float()
        C, N = f.shape[0], f.shape[1] * f.shape[2]
        dev = f.device
        with _cuda.on_device(dev):
            norm = torch.empty(N, dtype=torch.float32, device=dev)
            scratch = torch.empty(lib.trase_b200_feature_norm_scratch_bytes(N) // 4, dtype=torch.float32, device=dev)
            out2 = torch.empty(2, dtype=torch.float32, device=dev)
            _lib.check(lib.trase_b200_feature_norm_reg_forward(C, N, f.data_ptr(), norm.data_ptr(), scratch.data_ptr(),
                                                               out2.data_ptr(), _stream(dev)))
        ctx.save_for_backward(f, norm, out2)
        ctx.shape = feat.shape
        return out2                                   # {mean norm, (1 - mean norm)^2}

    @staticmethod
    def backward(ctx, g_out2):
        f, norm, out2 = ctx.saved_tensors
        lib = _lib.load()
        dev = f.device
        with _cuda.on_device(dev):
            d = torch.empty_like(f)
            go = g_out2.detach().contiguous().float()
            _lib.check(lib.trase_b200_feature_norm_reg_backward(f.shape[0], norm.numel(), f.data_ptr(), norm.data_ptr(),
                                                                out2.data_ptr(), go.data_ptr(), d.data_ptr(),
                                                                _stream(dev)))
        return d.view(ctx.shape)


def feature_norm_reg(rendered_features, return_norm=False):
    """train.py:280-281: (1 - rendered_features.norm(dim=0, p=2).mean()) ** 2, one kernel per direction."""
    out2 = _FeatureNormReg.apply(rendered_features)
    return (out2[1], out2[0]) if return_norm else out2[1]


# ---------------------------------------------------------------------------------------------------------
# FEATURE-state contrastive losses in their default "soft" mining mode (utils/loss_utils.py:304-349) and the
# style-transfer loop's nearest-neighbour feature loss (:223-228).  These sit outside the rasterizer path and
# work on the N x N similarity matrix of the ~5000 sampled pixels: plain torch, written with masked sums so that
# no boolean-index gather and no host synchronisation is needed (the reference syncs twice per loss).

def _soft_mined_pair_loss(C, C_F, mined, same_mask, weights, per_pair):
    """Shared form of the two losses.  `mined[i, j]`: pair (i, j) triggers mining; a column j is kept when any row
    mined it; kept pairs are the strictly-upper-triangle entries of kept columns; the loss sums `per_pair` over
    the kept pairs whose mask relation is `same_mask` and divides by the number of kept pairs."""
    n = C_F.shape[0]
    col_kept = mined.any(dim=0)                                               # [n]
    upper = torch.ones(n, n, dtype=torch.bool, device=C_F.device).triu(1)     # i < j (the diagonal is excluded)
    kept = upper & col_kept[None, :]
    n_kept = kept.sum()
    sel = kept & same_mask
    # the weights are masked first (not the products): an unselected pair must not see its weight at all, in
    # either direction — degenerate inputs give NaN weights (0/0 in the reference's min-max mapping)
    w_sel = sel.to(per_pair.dtype) if weights is None else torch.where(sel, weights, torch.zeros_like(weights))
    total = (w_sel * per_pair).sum()
    # no selected pair: the reference returns the Python float 0.0 (and prints a warning)
    return torch.where(sel.any(), total / n_kept.clamp_min(1).to(total.dtype), torch.zeros_like(total))


def pixel_mask_correspondence_loss_soft_hard_positive(C, C_F, positive_th=0.75, weights=None, **_unused):
    """utils/loss_utils.py:304-325: pull together pixel pairs of the same mask, mined by 'some same-mask pair in
    this column is still below positive_th'."""
    same = C == 1
    return _soft_mined_pair_loss(C, C_F, (C_F < positive_th) & same, same, weights, -C_F)


def pixel_mask_correspondence_loss_soft_negative(C, C_F, negative_th=0.5, weights=None, **_unused):
    """utils/loss_utils.py:327-349: push apart pixel pairs of different masks (hinge at 0), mined by 'some
    different-mask pair in this column is still above negative_th'."""
    diff = C == 0
    return _soft_mined_pair_loss(C, C_F, (C_F > negative_th) & diff, diff, weights, torch.relu(C_F))


def loss_nnfm_style(feat1, feats2):
    """utils/loss_utils.py:223-228: mean over the columns of feat1 [C, N1] of the cosine distance to the nearest
    column of feats2 [C, N2]."""
    a = feat1 / torch.linalg.norm(feat1, dim=0)
    b = feats2 / torch.linalg.norm(feats2, dim=0)
    return (1.0 - a.t() @ b).amin(dim=1).mean()
