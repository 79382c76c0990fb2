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
