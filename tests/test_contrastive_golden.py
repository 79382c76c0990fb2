"""FEATURE-state loss front end against golden vectors produced by the reference's own functions
(tests/golden/make_golden_contrastive.py imports utils/feature_utils.py and utils/loss_utils.py from /root/reference).

CPU: the torch restatements (mask-correspondence matrix, pixel weights, the two soft-mined losses, loss_nnfm_style).
GPU: the whole FEATURE-state loss through the fused front end (feature-norm kernel + sampled bilinear resize kernel)
with its gradient w.r.t. the rendered feature image."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from trase_b200 import feature_utils as FU
from trase_b200 import loss_utils as LU

CASES = ["mixed", "no_negative"]


def _case(g, tag, dev="cpu"):
    t = lambda k: torch.from_numpy(g[tag + "_" + k]).to(dev)
    return t("masks").long(), t("feat"), t("sampled_pixel"), t("sampled_mask")


@pytest.mark.parametrize("tag", CASES)
def test_mask_matrix_and_weights_match_the_reference(tag):
    g = load_golden("contrastive")
    masks, _, sp, sm = _case(g, tag)
    C = FU.get_pixel_mask_correspondence_matrix(masks, sp, sm)
    assert np.array_equal(C.numpy(), g[tag + "_C"])
    w = FU.get_pixel_weights(masks, sp)
    assert np.allclose(w.numpy(), g[tag + "_weights"], rtol=1e-6, atol=1e-6, equal_nan=True)


@pytest.mark.parametrize("tag", CASES)
def test_soft_mined_losses_match_the_reference(tag):
    g = load_golden("contrastive")
    C, w = torch.from_numpy(g[tag + "_C"]), torch.from_numpy(g[tag + "_weights"])
    C_F = torch.from_numpy(g[tag + "_C_F"]).requires_grad_(True)
    pos = LU.pixel_mask_correspondence_loss_soft_hard_positive(C=C, C_F=C_F, positive_th=0.75, weights=w)
    neg = LU.pixel_mask_correspondence_loss_soft_negative(C=C, C_F=C_F, negative_th=0.5, weights=w)
    assert abs(float(pos) - g[tag + "_vals"][0]) <= 1e-6 * max(1.0, abs(g[tag + "_vals"][0]))
    assert abs(float(neg) - g[tag + "_vals"][1]) <= 1e-6 * max(1.0, abs(g[tag + "_vals"][1]))
    (pos + neg).backward()                       # differentiable, and finite when nothing was mined
    assert torch.isfinite(C_F.grad).all()


def test_sample_pixel_and_mask_rates_and_coverage():
    g = torch.Generator().manual_seed(3)
    masks = torch.zeros(50, 60, 80, dtype=torch.int64)
    for i in range(50):
        masks[i, i:i + 10, 2 * (i % 30):2 * (i % 30) + 20] = 1
    sp, sm = FU.get_sample_pixel_and_mask(masks, 1500, 25, generator=g)
    assert sp.shape == (60, 80) and sm.shape == (50,)
    assert not (sp & (masks.sum(0) == 0)).any()              # uncovered pixels are never sampled
    assert 10 <= int(sm.sum()) <= 40
    covered = float((masks.sum(0) != 0).float().mean())
    assert abs(int(sp.sum()) - 1500 * covered) < 200


def test_nnfm_style_loss_matches_the_reference():
    g = load_golden("contrastive")
    f1 = torch.from_numpy(g["nnfm_f1"]).requires_grad_(True)
    l = LU.loss_nnfm_style(f1, torch.from_numpy(g["nnfm_f2"]))
    l.backward()
    assert abs(float(l) - g["nnfm_val"][0]) <= 1e-6
    assert np.abs(f1.grad.numpy() - g["nnfm_grad"]).max() <= 1e-6 * max(1.0, np.abs(g["nnfm_grad"]).max())


def test_smooth_map_helper_returns_self_first():
    xyz = torch.randn(200, 3, generator=torch.Generator().manual_seed(1))
    idx = FU.build_feature_smooth_map(xyz, K=4)
    assert idx.shape == (200, 4) and torch.equal(idx[:, 0], torch.arange(200))
    d = (xyz[:, None] - xyz[None]).norm(dim=-1)
    assert torch.equal(idx.sort(1).values, d.topk(4, largest=False).indices.sort(1).values)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", CASES)
def test_feature_state_loss_and_gradient_on_gpu(native_lib, tag):
    """train.py:280-296 end to end on the GPU: fused feature-norm regulariser + fused sampled resize + the losses."""
    g = load_golden("contrastive")
    dev = torch.device("cuda:0")
    masks, feat, sp, sm = _case(g, tag, dev)
    feat.requires_grad_(True)
    C = FU.get_pixel_mask_correspondence_matrix(masks, sp, sm)
    w = FU.get_pixel_weights(masks, sp)
    reg = LU.feature_norm_reg(feat)
    C_F = FU.get_features_correspondence_matrix(feat, sp)          # resize folded in
    assert np.abs(C_F.detach().cpu().numpy() - g[tag + "_C_F"]).max() <= 2e-6
    pos = LU.pixel_mask_correspondence_loss_soft_hard_positive(C=C, C_F=C_F, positive_th=0.75, weights=w)
    neg = LU.pixel_mask_correspondence_loss_soft_negative(C=C, C_F=C_F, negative_th=0.5, weights=w)
    loss = pos + neg + 1.0 * reg
    loss.backward()
    vals = g[tag + "_vals"]
    for got, ref in zip((pos, neg, reg, loss), vals):
        assert abs(float(got) - ref) <= 2e-5 * max(1.0, abs(ref))
    gref = g[tag + "_grad"]
    assert np.abs(feat.grad.cpu().numpy() - gref).max() <= 1e-4 * np.abs(gref).max()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", CASES)
def test_fused_soft_contrastive_losses_on_gpu(native_lib, tag):
    """The same loss through the fused kernels that never build an N x N tensor (csrc/contrastive.cu): values and the
    gradient w.r.t. the rendered feature image against the reference's golden vectors, and the per-pixel mask data
    (membership bits, mean mask area) against the torch formulation."""
    g = load_golden("contrastive")
    dev = torch.device("cuda:0")
    masks, feat, sp, sm = _case(g, tag, dev)
    feat.requires_grad_(True)
    reg = LU.feature_norm_reg(feat)
    pos, neg = FU.soft_contrastive_losses(feat, masks, sp, sm, positive_th=0.75, negative_th=0.5)
    loss = pos + neg + 1.0 * reg
    loss.backward()
    vals = g[tag + "_vals"]
    for got, ref in zip((pos, neg, reg, loss), vals):
        assert abs(float(got.detach()) - ref) <= 2e-5 * max(1.0, abs(ref)), (float(got.detach()), ref)
    gref = g[tag + "_grad"]
    assert np.abs(feat.grad.cpu().numpy() - gref).max() <= 1e-4 * np.abs(gref).max()


@pytest.mark.gpu
def test_fused_soft_contrastive_matches_torch_formulation_at_training_size(native_lib):
    """~4000 sampled pixels, 25 masks, unweighted and weighted: fused kernels against the torch restatement (which is
    pinned to the reference by the golden vectors above).  A column whose only mining pair sits within rounding of a
    threshold may flip between the two evaluations of C_F; the values must still agree to 1e-4."""
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(11)
    H, W, Hm, Wm, n_masks = 240, 136, 120, 68, 25
    masks = torch.zeros(n_masks, Hm, Wm, dtype=torch.uint8)
    for i in range(n_masks):
        h, w = int(torch.randint(10, 60, (1,), generator=gen)), int(torch.randint(8, 34, (1,), generator=gen))
        y, x = int(torch.randint(0, Hm - h, (1,), generator=gen)), int(torch.randint(0, Wm - w, (1,), generator=gen))
        masks[i, y:y + h, x:x + w] = 1
    masks = masks.to(dev)
    base = torch.randn(32, 1, 1, generator=gen)
    feat0 = (1.5 * torch.randn(32, H, W, generator=gen) + base).to(dev)   # similarities straddle both thresholds
    sp, sm = FU.get_sample_pixel_and_mask(masks, 4000, 20, generator=torch.Generator(device=dev).manual_seed(3))
    for use_w in (False, True):
        f1 = feat0.clone().requires_grad_(True)
        C = FU.get_pixel_mask_correspondence_matrix(masks, sp, sm)
        C_F = FU.get_features_correspondence_matrix(f1, sp)
        w = FU.get_pixel_weights(masks, sp) if use_w else None
        p1 = LU.pixel_mask_correspondence_loss_soft_hard_positive(C=C, C_F=C_F, positive_th=0.75, weights=w)
        n1 = LU.pixel_mask_correspondence_loss_soft_negative(C=C, C_F=C_F, negative_th=0.5, weights=w)
        (p1 + 2.0 * n1).backward()
        f2 = feat0.clone().requires_grad_(True)
        p2, n2 = FU.soft_contrastive_losses(f2, masks, sp, sm, 0.75, 0.5, use_weights=use_w)
        (p2 + 2.0 * n2).backward()
        assert C_F.shape[0] > 2000 and float(n1) != 0.0 and float(p1) != 0.0
        assert abs(float(p1) - float(p2)) <= 1e-4 * max(1e-3, abs(float(p1))), (float(p1), float(p2))
        assert abs(float(n1) - float(n2)) <= 1e-4 * max(1e-3, abs(float(n1))), (float(n1), float(n2))
        gmax = f1.grad.abs().max()
        assert (f1.grad - f2.grad).abs().max() <= 2e-4 * gmax


@pytest.mark.gpu
def test_fused_contrastive_many_masks_and_empty_sample(native_lib):
    """130 masks (three 64-bit membership words per pixel) against the torch formulation, and the degenerate call
    with no sampled pixel at all (both losses 0, zero gradient, no launch with N = 0)."""
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(21)
    H, W, Hm, Wm, n_masks = 96, 80, 48, 40, 130
    masks = torch.zeros(n_masks, Hm, Wm, dtype=torch.uint8)
    for i in range(n_masks):
        h, w = int(torch.randint(4, 20, (1,), generator=gen)), int(torch.randint(4, 16, (1,), generator=gen))
        y, x = int(torch.randint(0, Hm - h, (1,), generator=gen)), int(torch.randint(0, Wm - w, (1,), generator=gen))
        masks[i, y:y + h, x:x + w] = 1
    masks = masks.to(dev)
    feat0 = (1.2 * torch.randn(32, H, W, generator=gen) + torch.randn(32, 1, 1, generator=gen)).to(dev)
    sp, sm = FU.get_sample_pixel_and_mask(masks, 600, 100, generator=torch.Generator(device=dev).manual_seed(5))
    assert int(sm.sum()) > 64                              # sampled masks beyond the first membership word
    f1 = feat0.clone().requires_grad_(True)
    C = FU.get_pixel_mask_correspondence_matrix(masks, sp, sm)
    C_F = FU.get_features_correspondence_matrix(f1, sp)
    w = FU.get_pixel_weights(masks, sp)
    p1 = LU.pixel_mask_correspondence_loss_soft_hard_positive(C=C, C_F=C_F, positive_th=0.75, weights=w)
    n1 = LU.pixel_mask_correspondence_loss_soft_negative(C=C, C_F=C_F, negative_th=0.5, weights=w)
    (p1 + n1).backward()
    f2 = feat0.clone().requires_grad_(True)
    p2, n2 = FU.soft_contrastive_losses(f2, masks, sp, sm, 0.75, 0.5)
    (p2 + n2).backward()
    assert abs(float(p1) - float(p2)) <= 1e-4 * max(1e-3, abs(float(p1))), (float(p1), float(p2))
    assert abs(float(n1) - float(n2)) <= 1e-4 * max(1e-3, abs(float(n1))), (float(n1), float(n2))
    assert (f1.grad - f2.grad).abs().max() <= 2e-4 * f1.grad.abs().max()
    f3 = feat0.clone().requires_grad_(True)
    p3, n3 = FU.soft_contrastive_losses(f3, masks, torch.zeros_like(sp), sm, 0.75, 0.5)
    (p3 + n3).backward()
    assert float(p3) == 0.0 and float(n3) == 0.0 and not f3.grad.any()
    with pytest.raises(RuntimeError):
        FU.soft_contrastive_losses(feat0, torch.zeros(300, Hm, Wm, dtype=torch.uint8, device=dev), sp,
                                   torch.ones(300, dtype=torch.bool, device=dev))
