"""Fused Adam over the arenas and the densification statistics (trase_b200/optim.py, csrc/optim.cu) against
torch.optim.Adam configured as the reference does (scene/gaussian_model.py:259-289: per-attribute groups,
lr=0.0 default, eps=1e-15, learning rates rewritten between steps) and the reference's three-line
add_densification_stats (scene/gaussian_model.py:637-639).  Tolerance: 2e-6 relative to the largest parameter
(fp32 op-for-op restatement; torch fuses differently)."""
import copy

import pytest
import torch


def _params(dev):
    g = torch.Generator().manual_seed(0)
    shapes = {"xyz": (500, 3), "f_dc": (500, 1, 3), "f_rest": (500, 15, 3), "opacity": (500, 1), "scaling": (500, 3),
              "rotation": (500, 4), "gaussian_feats": (500, 1, 32)}
    return {k: torch.nn.Parameter(torch.randn(*s, generator=g).to(dev)) for k, s in shapes.items()}


@pytest.mark.gpu
def test_arena_adam_matches_torch_adam():
    from trase_b200.distributed import GradArena
    from trase_b200.optim import ArenaAdam
    dev = torch.device("cuda:0")
    ours, ref = _params(dev), None
    ref = {k: torch.nn.Parameter(v.detach().clone()) for k, v in ours.items()}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 5e-2, "scaling": 5e-3, "rotation": 1e-3,
           "gaussian_feats": 2.5e-3}
    groups = lambda d: [{"params": [d[k]], "lr": lrs[k], "name": k} for k in d]
    topt = torch.optim.Adam(groups(ref), lr=0.0, eps=1e-15)
    arena = GradArena(list(ours.values()))
    opt = ArenaAdam(arena, groups(ours), eps=1e-15)
    for k in ours:                                   # parameters are now views of the flat buffer, values kept
        assert torch.equal(ours[k].detach(), ref[k].detach())
    g = torch.Generator(device=dev).manual_seed(1)
    for it in range(6):
        new_lr = 1.6e-4 * (0.9 ** it)                # the reference rewrites the xyz learning rate every iteration
        for o in (topt, opt):
            for grp in o.param_groups:
                if grp["name"] == "xyz":
                    grp["lr"] = new_lr
        arena.zero()
        for k in ours:
            gr = torch.randn(ours[k].shape, device=dev, generator=g) * (10.0 ** (it - 3))
            ours[k].grad.copy_(gr)
            ref[k].grad = gr.clone()
        topt.step()
        opt.step()
    scale = max(v.detach().abs().max().item() for v in ref.values())
    for k in ours:
        assert (ours[k].detach() - ref[k].detach()).abs().max().item() <= 2e-6 * scale, k


@pytest.mark.gpu
def test_arena_adam_steps_named_groups_only():
    """TRASE alternates a GAUSSIAN and a FEATURE optimizer; groups not named in step() keep parameters and moments."""
    from trase_b200.distributed import GradArena
    from trase_b200.optim import ArenaAdam
    dev = torch.device("cuda:0")
    ps = _params(dev)
    arena = GradArena(list(ps.values()))
    opt = ArenaAdam(arena, [{"params": [ps[k]], "lr": 1e-2, "name": k} for k in ps])
    before = {k: v.detach().clone() for k, v in ps.items()}
    arena.flat.fill_(0.5)
    opt.step(groups=("gaussian_feats",))
    for k in ps:
        changed = not torch.equal(ps[k].detach(), before[k])
        assert changed == (k == "gaussian_feats"), k
    assert opt.exp_avg[: ps["xyz"].numel()].abs().max().item() == 0.0


@pytest.mark.gpu
def test_densification_stats_match_reference_lines():
    from trase_b200.optim import add_densification_stats
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(3)
    P = 10007
    vs = torch.zeros(P, 3, device=dev, requires_grad=True)
    vs.grad = torch.randn(P, 3, device=dev, generator=g)
    filt = torch.rand(P, device=dev, generator=g) > 0.4
    acc, den = torch.rand(P, 1, device=dev, generator=g), torch.ones(P, 1, device=dev)
    acc_ref, den_ref = acc.clone(), den.clone()
    acc_ref[filt] += torch.norm(vs.grad[filt, :2], dim=-1, keepdim=True)      # scene/gaussian_model.py:638
    den_ref[filt] += 1                                                          # :639
    add_densification_stats(acc, den, vs, filt)
    assert torch.allclose(acc, acc_ref, rtol=1e-6, atol=1e-7) and torch.equal(den, den_ref)
