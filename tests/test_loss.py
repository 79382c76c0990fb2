"""Fused L1 + SSIM loss (trase_b200/loss_utils.py, csrc/loss.cu).

CPU: the torch restatement (oracle/loss_oracle.py) against vectors produced by the reference's own
`l1_loss` / `ssim` (tests/golden/l1_ssim.npz).  GPU: the fused kernels against the same vectors and
against the restatement evaluated in float64.

Tolerances: loss values 1e-5 absolute (the SSIM map divides differences of fp32 blurred moments, and the
reference's own fp32 result moves by ~1e-6 with the convolution algorithm); image gradients 1e-3 of the
largest gradient magnitude (north_star's gradient tolerance)."""
import os

import numpy as np
import pytest
import torch

from oracle import loss_oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "l1_ssim.npz")
VAL_TOL, GRAD_TOL = 1e-5, 1e-3


def _gold(tag):
    z = np.load(GOLD)
    return (torch.from_numpy(z[f"{tag}_img"]), torch.from_numpy(z[f"{tag}_gt"]), z[f"{tag}_vals"],
            torch.from_numpy(z[f"{tag}_grad"]))


def _oracle(img, gt, lam=0.2, dtype=torch.float64):
    x = img.detach().to("cpu", dtype).requires_grad_(True)
    y = gt.detach().to("cpu", dtype)
    l1, ss = loss_oracle.l1_ref(x, y), loss_oracle.ssim_ref(x, y)
    loss = (1 - lam) * l1 + lam * (1 - ss)
    loss.backward()
    return l1.item(), ss.item(), loss.item(), x.grad


@pytest.mark.parametrize("tag", ["ragged", "flat"])
def test_oracle_matches_reference_vectors(tag):
    img, gt, vals, grad = _gold(tag)
    l1, ss, loss, g = _oracle(img, gt, dtype=torch.float32)
    assert abs(l1 - vals[0]) < VAL_TOL and abs(ss - vals[1]) < VAL_TOL and abs(loss - vals[2]) < VAL_TOL
    assert (g - grad).abs().max() <= GRAD_TOL * grad.abs().max()
    assert abs(loss_oracle.l1_ssim_ref(img, gt).item() - vals[2]) < VAL_TOL


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["ragged", "flat"])
def test_fused_loss_matches_reference_vectors(tag):
    from trase_b200 import loss_utils
    img, gt, vals, grad = _gold(tag)
    x = img.cuda().requires_grad_(True)
    loss, l1, ss = loss_utils.l1_ssim_loss(x, gt.cuda(), 0.2, return_terms=True)
    loss.backward()
    assert abs(l1.item() - vals[0]) < VAL_TOL and abs(ss.item() - vals[1]) < VAL_TOL
    assert abs(loss.item() - vals[2]) < VAL_TOL
    assert (x.grad.cpu() - grad).abs().max() <= GRAD_TOL * grad.abs().max()


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3, 32, 32), (3, 7, 9), (1, 100, 131), (2, 3, 64, 48), (3, 1014, 1352)])
def test_fused_loss_matches_float64_oracle(shape):
    from trase_b200 import loss_utils
    g = torch.Generator().manual_seed(sum(shape))
    gt = torch.rand(*shape, generator=g)
    img = (gt + 0.2 * torch.randn(*shape, generator=g)).clamp(0, 1)
    x = img.cuda().requires_grad_(True)
    # the two reference-named entry points separately, then the fused combination
    l1 = loss_utils.l1_loss(x, gt.cuda())
    ss = loss_utils.ssim(x, gt.cuda())
    (0.8 * l1 + 0.2 * (1.0 - ss)).backward()
    g_sep = x.grad.clone()
    x.grad = None
    loss = loss_utils.l1_ssim_loss(x, gt.cuda(), 0.2)
    (3.0 * loss).backward()                     # upstream gradient other than 1
    # the oracle sees a batch as B*C channels, like the unbatched reference call
    o_l1, o_ss, o_loss, o_grad = _oracle(img.reshape(-1, *shape[-2:]), gt.reshape(-1, *shape[-2:]))
    o_grad = o_grad.reshape(shape)
    assert abs(l1.item() - o_l1) < VAL_TOL and abs(ss.item() - o_ss) < VAL_TOL and abs(loss.item() - o_loss) < VAL_TOL
    scale = o_grad.abs().max()
    assert (g_sep.cpu().double() - o_grad).abs().max() <= GRAD_TOL * scale
    assert (x.grad.cpu().double() / 3.0 - o_grad).abs().max() <= GRAD_TOL * scale


@pytest.mark.gpu
def test_fused_loss_is_deterministic_and_rejects_bad_input():
    from trase_b200 import loss_utils
    g = torch.Generator().manual_seed(1)
    a, b = torch.rand(3, 200, 300, generator=g).cuda(), torch.rand(3, 200, 300, generator=g).cuda()
    v = [loss_utils.l1_ssim_loss(a, b).item() for _ in range(3)]
    assert v[0] == v[1] == v[2]
    with pytest.raises(RuntimeError):
        loss_utils.l1_loss(a.cpu(), b.cpu())
    with pytest.raises(RuntimeError):
        loss_utils.l1_loss(a, b[:, :10])
    with pytest.raises(NotImplementedError):
        loss_utils.ssim(a, b, window_size=7)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(32, 5, 7), (32, 64, 100), (4, 301, 513)])
def test_feature_norm_regulariser_matches_torch(shape):
    """train.py:280-281 in float64 torch against the fused kernels (value 1e-6 abs, gradient 1e-3 rel)."""
    from trase_b200 import loss_utils
    g = torch.Generator().manual_seed(shape[1])
    f = torch.randn(*shape, generator=g) * 0.3
    f[:, 0, :3] = 0.0                                  # zero-norm pixels: gradient must be 0 there, not NaN
    x = f.cuda().requires_grad_(True)
    reg, mean_norm = loss_utils.feature_norm_reg(x, return_norm=True)
    (2.0 * reg).backward()
    xr = f.double().requires_grad_(True)
    m = xr.norm(dim=0, p=2).mean()
    r = (1 - m) ** 2
    (2.0 * r).backward()
    assert abs(reg.item() - r.item()) < 1e-6 and abs(mean_norm.item() - m.item()) < 1e-6
    assert torch.isfinite(x.grad).all()
    assert (x.grad.cpu().double() - xr.grad).abs().max() <= GRAD_TOL * xr.grad.abs().max()


@pytest.mark.gpu
@pytest.mark.parametrize("shape,mask_hw", [((32, 101, 135), (64, 80)), ((32, 64, 64), (64, 64)), ((8, 40, 50), (90, 77))])
def test_sampled_feature_front_end_matches_reference_formulation(shape, mask_hw):
    """train.py:283 + utils/feature_utils.py:50-55 in torch (interpolate the whole image, boolean-index, normalize,
    einsum) against the fused sampler: correspondence matrix 1e-5 abs, image gradient 1e-3 of its max."""
    import torch.nn.functional as F
    from trase_b200 import feature_utils
    g = torch.Generator().manual_seed(shape[1])
    feat = torch.randn(*shape, generator=g)
    sampled = torch.rand(*mask_hw, generator=g) < 0.05
    up = torch.randn(int(sampled.sum()), int(sampled.sum()), generator=g)
    x = feat.cuda().requires_grad_(True)
    cf = feature_utils.get_features_correspondence_matrix(x, sampled.cuda(), mask_hw)
    (cf * up.cuda()).sum().backward()
    xr = feat.double().requires_grad_(True)
    r = F.interpolate(xr.unsqueeze(0), mask_hw, mode="bilinear").squeeze(0)
    s = F.normalize(r[:, sampled].permute(1, 0), dim=-1, p=2)
    cr = torch.einsum("hc,jc->hj", s, s)
    (cr * up.double()).sum().backward()
    assert cf.shape == cr.shape
    assert (cf.cpu().double() - cr).abs().max() <= 1e-5
    assert (x.grad.cpu().double() - xr.grad).abs().max() <= GRAD_TOL * xr.grad.abs().max()
