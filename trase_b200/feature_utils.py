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

from . import _cuda, _lib


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
        with _cuda.on_device(dev):
            out = torch.empty((N, C), dtype=torch.float32, device=dev)
            inv = torch.empty((N,), dtype=torch.float32, device=dev)
            _lib.check(lib.trase_b200_sample_features_forward(C, H, W, Hm, Wm, N, f.data_ptr(), idx.data_ptr(),
                                                              int(normalize), out.data_ptr(), inv.data_ptr(),
                                                              _cuda.raw_stream(dev)))
        ctx.save_for_backward(idx, out, inv)
        ctx.meta = (C, H, W, Hm, Wm, N, int(normalize))
        return out

    @staticmethod
    def backward(ctx, g_out):
        idx, out, inv = ctx.saved_tensors
        C, H, W, Hm, Wm, N, normalize = ctx.meta
        lib = _lib.load()
        dev = out.device
        with _cuda.on_device(dev):
            d = torch.zeros((C, H, W), dtype=torch.float32, device=dev)
            g = g_out.detach().contiguous().float()
            _lib.check(lib.trase_b200_sample_features_backward(C, H, W, Hm, Wm, N, idx.data_ptr(), normalize,
                                                               out.data_ptr(), inv.data_ptr(), g.data_ptr(), d.data_ptr(),
                                                               _cuda.raw_stream(dev)))
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


# ---------------------------------------------------------------------------------------------------------
# No-grad front end of the FEATURE state (utils/feature_utils.py:17-48): which pixels and masks take part in
# this step's contrastive loss, the pixel-pair weights and the mask-correspondence matrix.  Plain torch on the
# caller's device (the reference hard-codes .cuda(); these follow the masks' device and take an optional
# generator so that tests are reproducible).

@torch.no_grad()
def get_sample_pixel_and_mask(sam_masks, num_sampled_pixels, num_sampled_masks, generator=None):
    """utils/feature_utils.py:17-26: Bernoulli samples of the masks (rate num_sampled_masks / N) and of the pixels
    (rate num_sampled_pixels / (H*W)); pixels that no mask covers are dropped."""
    dev = sam_masks.device
    n_masks, Hm, Wm = sam_masks.shape[0], sam_masks.shape[-2], sam_masks.shape[-1]
    u_mask = torch.rand(n_masks, generator=generator, device=dev if generator is None else generator.device).to(dev)
    u_pix = torch.rand(Hm, Wm, generator=generator, device=dev if generator is None else generator.device).to(dev)
    sampled_mask = u_mask < num_sampled_masks / n_masks
    covered = sam_masks.sum(dim=0) != 0
    sampled_pixel = (u_pix < num_sampled_pixels / (Hm * Wm)) & covered
    return sampled_pixel, sampled_mask


@torch.no_grad()
def get_pixel_weights(sam_masks, sampled_pixel):
    """utils/feature_utils.py:28-38: pairs of pixels that lie in small masks weigh more.  Per pixel the mean area of
    the masks covering it; per pair the product of the two; weight = max product / product clamped to >= 1, then
    mapped affinely onto [1, 10]."""
    area = sam_masks.sum(dim=(-2, -1))                                        # [n_masks]
    cover = sam_masks.sum(dim=0)                                              # masks per pixel
    mean_area = (sam_masks * area[:, None, None]).sum(dim=0) / (cover + 1e-9)
    a = mean_area[sampled_pixel]                                              # [N]
    prod = a[:, None] * a[None, :]
    top = prod.max()
    prod = torch.where(prod == 0, torch.full_like(prod, 1e10), prod)
    w = (top / prod).clamp_min(1.0)
    lo, hi = w.min(), w.max()
    return (w - lo) / (hi - lo) * 9.0 + 1.0


@torch.no_grad()
def get_pixel_mask_correspondence_matrix(sam_masks, sampled_pixel, sampled_mask):
    """utils/feature_utils.py:40-48: C[h, j] = 1 when sampled pixels h and j share at least one sampled mask."""
    member = sam_masks[:, sampled_pixel][sampled_mask, :].float()             # [masks, N]
    shared = member.t() @ member
    return (shared != 0).to(shared.dtype)


def build_feature_smooth_map(xyz, K=16):
    """Index map of the K nearest Gaussians of every Gaussian (self included), [P, K] int64 — what
    scene/gaussian_model.py:86-92 obtains from pytorch3d.ops.knn_points.  The kNN search itself is outside the
    accelerated path; this helper uses scipy's k-d tree on the host so that callers without pytorch3d can still
    fill `GaussianParams.feature_smooth_map`."""
    from scipy.spatial import cKDTree
    pts = xyz.detach().cpu().double().numpy()
    _, idx = cKDTree(pts).query(pts, k=K)
    return torch.from_numpy(idx.reshape(len(pts), K)).to(device=xyz.device, dtype=torch.int64)


# ---------------------------------------------------------------------------------------------------------
# The whole 'soft' contrastive loss of the FEATURE state on the sampled pixels, without any N x N tensor
# (csrc/contrastive.cu): mask relation, pair weights and similarities are recomputed from per-pixel data.
LAST_NUM_SAMPLES = 0


class _SoftContrastive(torch.autograd.Function):
    @staticmethod
    def forward(ctx, f, bits, a, a_minmax, pos_th, neg_th):
        lib = _lib.load()
        dev = f.device
        N, nwords = int(f.shape[0]), int(bits.shape[1])
        fc = f.detach().contiguous().float()
        with _cuda.on_device(dev):
            scratch = torch.empty(max(1, lib.trase_b200_contrastive_scratch_bytes(N)), dtype=torch.uint8, device=dev)
            result = torch.zeros(8, dtype=torch.float32, device=dev)
            _lib.check(lib.trase_b200_contrastive_soft_forward(
                N, nwords, fc.data_ptr(), bits.data_ptr(), None if a is None else a.data_ptr(),
                None if a is None else a_minmax.data_ptr(), float(pos_th), float(neg_th), scratch.data_ptr(),
                result.data_ptr(), _cuda.raw_stream(dev)))
        ctx.save_for_backward(fc, bits, a, a_minmax, scratch, result)
        ctx.th = (float(pos_th), float(neg_th))
        return result[0:2].clone()

    @staticmethod
    def backward(ctx, g):
        lib = _lib.load()
        fc, bits, a, a_minmax, scratch, result = ctx.saved_tensors
        dev = fc.device
        N, nwords = int(fc.shape[0]), int(bits.shape[1])
        with _cuda.on_device(dev):
            up = g.detach().contiguous().float()
            df = torch.zeros_like(fc)
            _lib.check(lib.trase_b200_contrastive_soft_backward(
                N, nwords, fc.data_ptr(), bits.data_ptr(), None if a is None else a.data_ptr(),
                None if a is None else a_minmax.data_ptr(), ctx.th[0], ctx.th[1], scratch.data_ptr(), result.data_ptr(),
                up.data_ptr(), df.data_ptr(), _cuda.raw_stream(dev)))
        return df, None, None, None, None, None


def sampled_pixel_indices(sampled_pixel):
    """Flat indices of the True entries of the [Hm, Wm] sample mask, in row-major order (what boolean indexing
    yields).  This is the one host synchronisation of the loss front end (the count has to reach the host); it only
    depends on the sampling, not on the render, so a training loop can take it BEFORE launching the render — or on a
    side stream — and pass the result to soft_contrastive_losses(..., pix=...), instead of stalling between the
    forward pass and the loss."""
    return sampled_pixel.reshape(-1).nonzero().reshape(-1)


def soft_contrastive_losses(rendered_features, sam_masks, sampled_pixel, sampled_mask, positive_th=0.75,
                            negative_th=0.5, use_weights=True, pix=None):
    """(positive loss, negative loss) of train.py:283-292 in the default contrastive_mode 'soft', i.e.

        C   = get_pixel_mask_correspondence_matrix(sam_masks, sampled_pixel, sampled_mask)
        C_F = get_features_correspondence_matrix(F.interpolate(rendered_features, mask resolution), sampled_pixel)
        w   = get_pixel_weights(sam_masks, sampled_pixel)
        pixel_mask_correspondence_loss_soft_hard_positive(C, C_F, positive_th, w),  ..._soft_negative(C, C_F, negative_th, w)

    evaluated by four kernels that never build an N x N tensor.  `rendered_features` is the FULL-resolution [32,H,W]
    feature image (the bilinear resize is folded into the sampling).  CUDA only; up to 256 masks."""
    lib = _lib.load()
    if not rendered_features.is_cuda:
        raise RuntimeError("trase_b200: soft_contrastive_losses runs on CUDA tensors only")
    dev = rendered_features.device
    n_masks, Hm, Wm = int(sam_masks.shape[0]), int(sam_masks.shape[-2]), int(sam_masks.shape[-1])
    if n_masks > 256:
        raise RuntimeError("trase_b200: soft_contrastive_losses handles at most 256 masks per view")
    if pix is None:
        pix = sampled_pixel_indices(sampled_pixel)                # one host sync, like the reference's boolean index
    f = _SampleFeatures.apply(rendered_features, pix, Hm, Wm, True)           # [N, 32] normalised, differentiable
    N = int(pix.numel())
    global LAST_NUM_SAMPLES
    LAST_NUM_SAMPLES = N                    # for reports (bench.py): how many pixels the last call sampled
    if N == 0:
        z = f.sum() * 0.0
        return z, z
    with torch.no_grad(), _cuda.on_device(dev):
        m8 = sam_masks.to(device=dev, dtype=torch.uint8).contiguous()
        area = m8.sum(dim=(-2, -1), dtype=torch.int64)
        sm8 = sampled_mask.to(device=dev, dtype=torch.uint8).contiguous()
        bits = torch.empty((N, (n_masks + 63) // 64), dtype=torch.int64, device=dev)
        a = torch.empty((N,), dtype=torch.float32, device=dev)
        _lib.check(lib.trase_b200_pixel_mask_info(N, n_masks, Hm * Wm, m8.data_ptr(), area.data_ptr(), sm8.data_ptr(),
                                                  pix.data_ptr(), bits.data_ptr(), a.data_ptr(),
                                                  _cuda.raw_stream(dev)))
        a_minmax = torch.stack([a.min(), a.max()]) if use_weights else None
    out = _SoftContrastive.apply(f, bits, a if use_weights else None, a_minmax, positive_th, negative_th)
    return out[0], out[1]
