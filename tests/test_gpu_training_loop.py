"""Drop-in check at the level a TRASE user sees: a few iterations of the training step
(DeformNetwork -> render() -> L1+SSIM and feature-norm losses -> backward -> Adam) run twice —
once on this package (native rasterizer, fused glue, fused losses, ArenaAdam over the flat arenas) and once
on the reference's path (the unmodified compiled reference rasterizer from oracle/_ref, its elementwise torch
glue, the torch formulation of its losses, torch.optim.Adam configured as scene/gaussian_model.py:259-289) —
and the trained parameters are compared.  The DeformNetwork runs in fp32 torch on both sides so the
comparison isolates the rasterizer path.  Tolerance: 2e-3 of each tensor's largest magnitude after 4 steps
(gradient parity is 1e-3 per step; Adam's normalisation amplifies differences of tiny gradients)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _build(dev, seed=5):
    from trase_b200 import synth, render as R
    from trase_b200.deform import DeformNetwork
    sc = synth.make_scene(3000, 96, 64, seed)
    pc = R.GaussianParams.from_activated(sc, dev)
    torch.manual_seed(seed)
    net = DeformNetwork(backend="torch").to(dev)
    with torch.no_grad():
        for head in (net.gaussian_warp, net.gaussian_rotation, net.gaussian_scaling):
            head.weight.mul_(1e-2)
            head.bias.mul_(1e-2)
    view = R.make_view(sc["cam"], dev)
    g = torch.Generator().manual_seed(seed + 1)
    gt = torch.rand(3, 64, 96, generator=g).to(dev)
    return sc, pc, net, view, gt


LRS = {"_xyz": 1.6e-4, "_features_dc": 2.5e-3, "_features_rest": 2.5e-3 / 20, "_opacity": 5e-2, "_scaling": 5e-3,
       "_rotation": 1e-3, "_gaussian_features": 2.5e-3}


def _groups(pc):
    return [{"params": [getattr(pc, k)], "lr": lr, "name": k} for k, lr in LRS.items()]


def test_four_training_steps_match_the_reference_path():
    from oracle import loss_oracle, ref_loader
    if not ref_loader.available():
        pytest.skip("oracle/_ref not built")
    from trase_b200 import render as R, loss_utils
    from trase_b200.distributed import GradArena
    from trase_b200.optim import ArenaAdam
    dev = torch.device("cuda:0")
    bg = torch.zeros(3, device=dev)
    t_val = 0.3

    # ---- ours
    sc, pc, net, view, gt = _build(dev)
    arena = GradArena(list(pc.parameters()))
    opt = ArenaAdam(arena, _groups(pc), eps=1e-15)
    ours_losses = []
    for _ in range(4):
        opt.zero_grad()
        net.zero_grad()
        t = torch.full((3000, 1), t_val, device=dev)
        d_xyz, d_rot, d_scale = net(pc.get_xyz.detach(), t)
        out = R.render(view, pc, R.PIPE_DEFAULT, bg, d_xyz, d_rot, d_scale)
        loss = loss_utils.l1_ssim_loss(out["render"], gt, 0.2) + loss_utils.feature_norm_reg(out["render_gaussian_features"])
        loss.backward()
        opt.step()
        ours_losses.append(loss.item())

    # ---- the reference's path
    ref = ref_loader.load()
    sc, pc_r, net_r, view, gt = _build(dev)
    topt = torch.optim.Adam(_groups(pc_r), lr=0.0, eps=1e-15)
    saved = R.GaussianRasterizer, R.GaussianRasterizationSettings, R.FUSED_GLUE
    R.GaussianRasterizer, R.GaussianRasterizationSettings, R.FUSED_GLUE = (ref.GaussianRasterizer,
                                                                            ref.GaussianRasterizationSettings, False)
    ref_losses = []
    try:
        for _ in range(4):
            topt.zero_grad()
            net_r.zero_grad()
            t = torch.full((3000, 1), t_val, device=dev)
            d_xyz, d_rot, d_scale = net_r(pc_r.get_xyz.detach(), t)
            out = R.render(view, pc_r, R.PIPE_DEFAULT, bg, d_xyz, d_rot, d_scale)
            feats = out["render_gaussian_features"]
            loss = loss_oracle.l1_ssim_ref(out["render"], gt, 0.2) + (1.0 - feats.norm(dim=0, p=2).mean()) ** 2
            loss.backward()
            topt.step()
            ref_losses.append(loss.item())
    finally:
        R.GaussianRasterizer, R.GaussianRasterizationSettings, R.FUSED_GLUE = saved

    for a, b in zip(ours_losses, ref_losses):
        assert abs(a - b) <= 1e-4 * max(1.0, abs(b)), (ours_losses, ref_losses)
    assert ours_losses[-1] < ours_losses[0]                     # it trains
    for name in LRS:
        a, b = getattr(pc, name).detach(), getattr(pc_r, name).detach()
        assert (a - b).abs().max().item() <= 2e-3 * b.abs().max().item(), name
