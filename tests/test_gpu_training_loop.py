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


# ---------------------------------------------------------------------------------------------------------
# Densification over the flat arenas against the reference's per-tensor procedure, restated here on
# torch.optim.Adam exactly as scene/gaussian_model.py does it (cat_tensors_to_optimizer :513-534,
# _prune_optimizer :472-492, densify_and_clone :594-615, densify_and_split :563-592, densify_and_prune :617-635,
# reset_opacity :378-381 with replace_tensor_to_optimizer :457-470).
class _RefModel:
    NAMES = ["_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation", "_gaussian_features"]

    def __init__(self, tensors, lrs):
        self.p = {k: torch.nn.Parameter(v.clone()) for k, v in tensors.items()}
        self.opt = torch.optim.Adam([{"params": [self.p[k]], "lr": lrs[k], "name": k} for k in self.NAMES], lr=0.0,
                                    eps=1e-15)
        self.percent_dense = 0.01

    def _replace(self, fn_param, fn_state):
        for group in self.opt.param_groups:
            old = group["params"][0]
            st = self.opt.state.get(old, None)
            new = torch.nn.Parameter(fn_param(group["name"], old).requires_grad_(True))
            if st is not None:
                st["exp_avg"] = fn_state(group["name"], st["exp_avg"])
                st["exp_avg_sq"] = fn_state(group["name"], st["exp_avg_sq"])
                del self.opt.state[old]
                self.opt.state[new] = st
            group["params"][0] = new
            self.p[group["name"]] = new

    def cat(self, ext):
        self._replace(lambda n, t: torch.cat((t, ext[n]), dim=0),
                      lambda n, s: torch.cat((s, torch.zeros_like(ext[n])), dim=0))
        P = self.p["_xyz"].shape[0]
        dev = self.p["_xyz"].device
        self.accum, self.denom = torch.zeros((P, 1), device=dev), torch.zeros((P, 1), device=dev)
        self.max_radii2D = torch.zeros(P, device=dev)

    def prune(self, mask):
        keep = ~mask
        self._replace(lambda n, t: t[keep], lambda n, s: s[keep])
        self.accum, self.denom, self.max_radii2D = self.accum[keep], self.denom[keep], self.max_radii2D[keep]

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, N=2):
        from trase_b200.densify import build_rotation
        grads = self.accum / self.denom
        grads[grads.isnan()] = 0.0
        scaling = lambda: torch.exp(self.p["_scaling"])
        sel = (torch.norm(grads, dim=-1) >= max_grad) & (scaling().max(dim=1).values <= self.percent_dense * extent)
        n_clone = int(sel.sum())
        self.cat({k: self.p[k][sel] for k in self.NAMES})
        P = self.p["_xyz"].shape[0]
        padded = torch.zeros(P, device=grads.device)
        padded[:grads.shape[0]] = grads.squeeze()
        sel = (padded >= max_grad) & (scaling().max(dim=1).values > self.percent_dense * extent)
        n_split = int(sel.sum())
        stds = scaling()[sel].repeat(N, 1)
        samples = torch.normal(mean=torch.zeros((stds.size(0), 3), device=stds.device), std=stds)
        rots = build_rotation(self.p["_rotation"][sel]).repeat(N, 1, 1)
        ext = {k: self.p[k][sel].repeat(*([N] + [1] * (self.p[k].dim() - 1))) for k in self.NAMES}
        ext["_xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self.p["_xyz"][sel].repeat(N, 1)
        ext["_scaling"] = torch.log(scaling()[sel].repeat(N, 1) / (0.8 * N))
        self.cat(ext)
        self.prune(torch.cat((sel, torch.zeros(N * n_split, device=sel.device, dtype=torch.bool))))
        prune = (torch.sigmoid(self.p["_opacity"]) < min_opacity).squeeze()
        if max_screen_size:
            prune = prune | (self.max_radii2D > max_screen_size) | (scaling().max(dim=1).values > 0.1 * extent)
        self.prune(prune)
        return n_clone, n_split

    def reset_opacity(self):
        o = torch.min(torch.sigmoid(self.p["_opacity"]), torch.ones_like(self.p["_opacity"]) * 0.01)
        new = torch.log(o / (1 - o))
        for group in self.opt.param_groups:
            if group["name"] == "_opacity":
                old = group["params"][0]
                st = self.opt.state.get(old)
                st["exp_avg"], st["exp_avg_sq"] = torch.zeros_like(new), torch.zeros_like(new)
                del self.opt.state[old]
                group["params"][0] = torch.nn.Parameter(new.detach().requires_grad_(True))
                self.opt.state[group["params"][0]] = st
                self.p["_opacity"] = group["params"][0]


def test_densify_prune_and_opacity_reset_match_the_per_tensor_procedure(native_lib, tmp_path):
    """A few training steps, then densify_and_prune and reset_opacity on the flat arenas against the reference's
    per-tensor procedure started from the SAME state: Gaussian counts, every parameter and both Adam moments must be
    identical (they are copies and the same torch formulas); training then continues on the resized arenas; the
    checkpoint tuple round-trips and its optimizer halves load into stock torch.optim.Adam."""
    from trase_b200 import render as R, loss_utils, densify as D
    from trase_b200.distributed import GradArena
    from trase_b200.optim import ArenaAdam
    dev = torch.device("cuda:0")
    bg = torch.zeros(3, device=dev)
    sc, pc, net, view, gt = _build(dev)
    P0 = pc._xyz.shape[0]
    arena = GradArena(list(pc.parameters()))
    opt = ArenaAdam(arena, _groups(pc), eps=1e-15)
    state = D.DensifyState(P0, dev)

    def step():
        opt.zero_grad()
        P = pc._xyz.shape[0]
        t = torch.full((P, 1), 0.3, device=dev)
        with torch.no_grad():
            d = net(pc.get_xyz.detach(), t)
        out = R.render(view, pc, R.PIPE_DEFAULT, bg, d[0], d[1], d[2])
        loss = loss_utils.l1_ssim_loss(out["render"], gt, 0.2) + loss_utils.feature_norm_reg(out["render_gaussian_features"])
        loss.backward()
        state.update_max_radii(out["radii"], out["visibility_filter"])
        D.add_densification_stats(state.xyz_gradient_accum, state.denom, out["viewspace_points"], out["visibility_filter"])
        opt.step()
        return loss.item()

    for _ in range(5):
        step()
    # the reference-style model starts from exactly this state
    ref = _RefModel({k: getattr(pc, k).detach() for k in _RefModel.NAMES}, LRS)
    seg = {id(p): (off, p.numel()) for p, off in zip(arena.params, arena.offsets)}
    for group in ref.opt.param_groups:
        p = group["params"][0]
        off, n = seg[id(getattr(pc, group["name"]))]
        ref.opt.state[p] = {"step": torch.tensor(5.0), "exp_avg": opt.exp_avg[off:off + n].view_as(p).clone(),
                            "exp_avg_sq": opt.exp_avg_sq[off:off + n].view_as(p).clone()}
    ref.accum, ref.denom = state.xyz_gradient_accum.clone(), state.denom.clone()
    ref.max_radii2D = state.max_radii2D.clone()
    thr = float((state.xyz_gradient_accum / state.denom.clamp_min(1)).quantile(0.9))     # top 10 % densify
    extent = 100.0 * float(torch.exp(pc._scaling.detach()).max(dim=1).values.median())   # percent_dense * extent = the
    with torch.no_grad():                              # median size: clone and split both fire; and the two prune rules
        big = torch.arange(P0, device=dev) % 7 == 0
        pc._scaling.data[big] += 2.5
        ref.p["_scaling"].data[big] += 2.5
        pc._opacity.data[::11] = -8.0
        ref.p["_opacity"].data[::11] = -8.0
    torch.manual_seed(77)
    n_ref = ref.densify_and_prune(thr, 0.005, extent, 20)
    torch.manual_seed(77)
    n_ours = D.densify_and_prune(pc, opt, state, thr, 0.005, extent, 20)
    assert n_ours == n_ref and n_ours[0] > 0 and n_ours[1] > 0
    P1 = pc._xyz.shape[0]
    assert P1 == ref.p["_xyz"].shape[0] and P1 != P0
    seg = {id(p): (off, p.numel()) for p, off in zip(arena.params, arena.offsets)}
    for group in ref.opt.param_groups:
        k = group["name"]
        mine, theirs = getattr(pc, k), group["params"][0]
        assert mine.shape == theirs.shape, k
        assert torch.equal(mine.detach(), theirs.detach()), k
        off, n = seg[id(mine)]
        st = ref.opt.state[theirs]
        assert torch.equal(opt.exp_avg[off:off + n].view_as(mine), st["exp_avg"]), k
        assert torch.equal(opt.exp_avg_sq[off:off + n].view_as(mine), st["exp_avg_sq"]), k
        assert mine.grad.shape == mine.shape and mine.grad.data_ptr() == arena.flat.data_ptr() + 4 * off
    assert state.xyz_gradient_accum.shape == (P1, 1) and not state.max_radii2D.any()
    # opacity reset
    ref.reset_opacity()
    D.reset_opacity(pc, opt)
    assert torch.equal(pc._opacity.detach(), ref.p["_opacity"].detach())
    off, n = seg[id(pc._opacity)]
    assert not opt.exp_avg[off:off + n].any() and not opt.exp_avg_sq[off:off + n].any()
    # prune_points alone
    mask = torch.arange(P1, device=dev) % 5 == 0
    ref.prune(mask)
    D.prune_points(pc, opt, state, mask)
    assert torch.equal(pc._rotation.detach(), ref.p["_rotation"].detach())
    # training goes on over the resized arenas
    losses = [step() for _ in range(3)]
    assert all(l == l for l in losses) and pc._xyz.shape[0] == state.denom.shape[0]
    # checkpoint tuple (train.py:394-396) round trip
    path = tmp_path / "chkpnt7.pth"
    torch.save((D.capture(pc, opt, state, 1.5), 7), path)
    model_args, it = torch.load(path, weights_only=False)
    assert it == 7 and len(model_args) == 13
    sc2, pc2, net2, view2, gt2 = _build(dev)
    arena2 = GradArena(list(pc2.parameters()))
    opt2 = ArenaAdam(arena2, _groups(pc2), eps=1e-15)
    state2 = D.DensifyState(pc2._xyz.shape[0], dev)
    assert D.restore(pc2, opt2, state2, model_args) == 1.5
    for k in _RefModel.NAMES:
        assert torch.equal(getattr(pc2, k).detach(), getattr(pc, k).detach()), k
    assert torch.equal(opt2.exp_avg, opt.exp_avg) and torch.equal(opt2.exp_avg_sq, opt.exp_avg_sq)
    assert opt2._steps == opt._steps
    stock = torch.optim.Adam([{"params": [torch.nn.Parameter(getattr(pc, k).detach().clone())], "lr": 0.0, "name": k}
                              for k in LRS if k != "_gaussian_features"], lr=0.0, eps=1e-15)
    stock.load_state_dict(model_args[11]["GAUSSIAN"])           # each half is a stock Adam state_dict
    assert len(stock.state) == 6


def test_prune_keeps_other_arena_tensors_intact(native_lib):
    """An arena that also holds the DeformNetwork (as bench.py builds it): resizing the per-Gaussian segments must
    carry the network's parameters and moments over unchanged and leave every .grad a view of the new arena."""
    from trase_b200 import densify as D
    from trase_b200.deform import DeformNetwork
    from trase_b200.distributed import GradArena
    from trase_b200.optim import ArenaAdam
    dev = torch.device("cuda:0")
    sc, pc, _, _, _ = _build(dev)
    torch.manual_seed(1)
    net = DeformNetwork(backend="native").to(dev)
    params = list(pc.parameters()) + list(net.parameters())
    arena = GradArena(params, early=list(pc.parameters()))
    groups = _groups(pc) + [{"params": list(net.parameters()), "lr": 1e-3, "name": "deform"}]
    opt = ArenaAdam(arena, groups, eps=1e-15)
    opt.exp_avg.uniform_(-1, 1)
    opt.exp_avg_sq.uniform_(0, 1)
    before_net = [p.detach().clone() for p in net.parameters()]
    seg = {id(p): (off, p.numel()) for p, off in zip(arena.params, arena.offsets)}
    before_m = [opt.exp_avg[seg[id(p)][0]:seg[id(p)][0] + p.numel()].clone() for p in net.parameters()]
    xyz_before = pc._xyz.detach().clone()
    state = D.DensifyState(pc._xyz.shape[0], dev)
    mask = torch.arange(pc._xyz.shape[0], device=dev) % 3 == 0
    D.prune_points(pc, opt, state, mask)
    assert torch.equal(pc._xyz.detach(), xyz_before[~mask])
    seg = {id(p): (off, p.numel()) for p, off in zip(arena.params, arena.offsets)}
    for p, b, m in zip(net.parameters(), before_net, before_m):
        off, n = seg[id(p)]
        assert torch.equal(p.detach(), b)
        assert torch.equal(opt.exp_avg[off:off + n], m)
        assert p.grad.data_ptr() == arena.flat.data_ptr() + 4 * off
    assert arena.n_early == sum(p.numel() for p in pc.parameters())
    # the resized model still trains through the native path
    P = pc._xyz.shape[0]
    outs = net(pc.get_xyz.detach(), torch.full((P, 1), 0.2, device=dev))
    sum(o.sum() for o in outs).backward()
    opt.step()
    assert torch.isfinite(opt.flat_p).all()


@pytest.mark.gpu
def test_direct_gaussian_gradient_accumulation_matches_autograd(native_lib):
    """With the Gaussian parameters' `.grad` in a GradArena, the activation-glue backward adds their gradients straight
    into the arena slots (xyz, scaling, rotation, opacity, and the SH block split into f_dc / f_rest); autograd still
    fires the parameters' post-accumulate hooks (the arena's overlap trigger is one).  Same arena contents as
    autograd's own accumulation, also on top of existing contents and over two views; same hook count either way."""
    from trase_b200 import _assemble, synth
    from trase_b200 import render as R
    from trase_b200.distributed import GradArena
    dev = torch.device("cuda:0")
    P, W, H = 20000, 160, 128
    sc = synth.make_scene(P, W, H, 7)
    view = R.make_view(sc["cam"], dev)
    ups = [synth.make_upstream_grads(W, H, 7 + v) for v in range(2)]
    res, seen = {}, {}
    saved = _assemble.DIRECT_GRAD_ACCUM
    try:
        for direct in (False, True):
            _assemble.DIRECT_GRAD_ACCUM = direct
            pc = R.GaussianParams.from_activated(sc, dev)
            pc._gaussian_features.requires_grad_(False)                    # GAUSSIAN state
            params = [p for p in pc.parameters() if p.requires_grad]
            arena = GradArena(params, early=params)
            calls = []
            for p in params:                        # a second post-accumulate hook per parameter: must fire either way
                p.register_post_accumulate_grad_hook(lambda q: calls.append(float(q.grad.abs().sum()) > 0))
            arena.flat.fill_(0.125)
            for v in range(2):
                d = [torch.full((P, k), 1e-3 * (v + 1), device=dev, requires_grad=True) for k in (3, 4, 3)]
                out = R.render(view, pc, R.PIPE_DEFAULT, torch.zeros(3, device=dev), d[0], d[1], d[2])
                torch.autograd.backward([out["render"], out["render_gaussian_features"]],
                                        [ups[v]["dL_dcolor"].to(dev), ups[v]["dL_dobjects"].to(dev)])
            torch.cuda.synchronize()
            res[direct] = arena.flat.clone()
            seen[direct] = len(calls)
    finally:
        _assemble.DIRECT_GRAD_ACCUM = saved
    scale = (res[False] - 0.125).abs().max()
    assert scale > 0
    assert (res[False] - res[True]).abs().max() <= 1e-4 * scale + 1e-7      # atomics: the two runs differ in last bits
    assert seen[True] == seen[False] == 2 * 6            # six parameters, two views: the hooks fire on both paths
