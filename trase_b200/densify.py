"""Adaptive density control and checkpoints over the flat arenas (second part).

Mirrors GaussianModel's densification API (scene/gaussian_model.py:457-639): `densify_and_prune`, `prune_points`,
`reset_opacity`, `add_densification_stats`, and `capture` / `restore` for `chkpnt{N}.pth` (:106-151, train.py:394-396).

The reference re-allocates every parameter tensor and both Adam moments one tensor at a time (torch.cat for clones and
split children, boolean indexing for pruning, re-keying the optimizer state by Parameter object each time).  Here the
whole clone -> split -> prune sequence is reduced to ONE row map `src` (output Gaussian r is a copy of input Gaussian
src[r]) plus a `fresh` flag (clones and split children start with zero moments, scene/gaussian_model.py:517-518), and one
gather launch rebuilds all per-Gaussian segments of the parameter arena and of both moment arenas
(csrc/optim.cu arena_gather_rows_kernel).  The Parameter objects stay the same (only their storage moves), so nothing has
to be re-keyed and gradient hooks survive.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _cuda, _lib
from .optim import ArenaAdam, add_densification_stats  # noqa: F401  (re-exported: same place as in the reference)


class DensifyState:
    """xyz_gradient_accum [P,1], denom [P,1], max_radii2D [P] (scene/gaussian_model.py:255-256, 161)."""

    def __init__(self, P: int, device):
        self.percent_dense = 0.01
        self.reset(P, device)

    def reset(self, P: int, device):
        self.xyz_gradient_accum = torch.zeros((P, 1), device=device)
        self.denom = torch.zeros((P, 1), device=device)
        self.max_radii2D = torch.zeros((P,), device=device)

    def update_max_radii(self, radii, visibility_filter):
        """train.py:364."""
        self.max_radii2D[visibility_filter] = torch.max(self.max_radii2D[visibility_filter],
                                                        radii[visibility_filter].to(self.max_radii2D.dtype))


def build_rotation(r: torch.Tensor) -> torch.Tensor:
    """Rotation matrices of normalised quaternions (r, x, y, z) (utils/general_utils.py:122-143)."""
    q = r / r.norm(dim=1, keepdim=True)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)


def _gaussian_tensors(pc):
    return [pc._xyz, pc._features_dc, pc._features_rest, pc._scaling, pc._rotation, pc._opacity, pc._gaussian_features]


@torch.no_grad()
def regather(pc, opt: Optional[ArenaAdam], src: torch.Tensor, fresh: torch.Tensor) -> None:
    """Rebuild every per-Gaussian tensor (and, with an ArenaAdam, its moments and gradient slot) so that output row r
    is input row src[r]; moments of rows with fresh[r] start at zero.  Other tensors of the arena (e.g. the
    DeformNetwork) are carried over unchanged."""
    lib = _lib.load()
    P_old, P_new = int(pc._xyz.shape[0]), int(src.numel())
    dev = pc._xyz.device
    src = src.to(device=dev, dtype=torch.int64).contiguous()
    fresh = fresh.to(device=dev, dtype=torch.uint8).contiguous()
    mine = {id(p) for p in _gaussian_tensors(pc)}
    in_arena = set()
    if opt is not None:
        arena = opt.arena
        in_arena = {id(p) for p in arena.params}
        sizes = [(p.numel() // P_old) * P_new if id(p) in mine else p.numel() for p in arena.params]
        offs, n = [], 0
        for sz in sizes:
            offs.append(n)
            n += sz
        with _cuda.on_device(dev):
            new_p = torch.empty(n, device=dev)
            new_m = torch.empty(n, device=dev)
            new_v = torch.empty(n, device=dev)
            new_g = torch.zeros(n, device=dev)
            o_old, o_new, width = [], [], []
            for p, oo, on in zip(arena.params, arena.offsets, offs):
                if id(p) in mine:
                    o_old.append(oo); o_new.append(on); width.append(p.numel() // P_old)
                else:                                   # not per-Gaussian: plain copy
                    k = p.numel()
                    new_p[on:on + k].copy_(opt.flat_p[oo:oo + k])
                    new_m[on:on + k].copy_(opt.exp_avg[oo:oo + k])
                    new_v[on:on + k].copy_(opt.exp_avg_sq[oo:oo + k])
            if o_old and P_new > 0:
                ns = len(o_old)
                _lib.check(lib.trase_b200_arena_gather_rows(
                    ns, (C.c_int64 * ns)(*o_old), (C.c_int64 * ns)(*o_new), (C.c_int * ns)(*width), P_new,
                    src.data_ptr(), fresh.data_ptr(), opt.flat_p.data_ptr(), opt.exp_avg.data_ptr(),
                    opt.exp_avg_sq.data_ptr(), new_p.data_ptr(), new_m.data_ptr(), new_v.data_ptr(),
                    _cuda.raw_stream(dev)))
        ends = []
        for p, on, sz in zip(arena.params, offs, sizes):
            shape = (P_new,) + tuple(p.shape[1:]) if id(p) in mine else tuple(p.shape)
            p.data = new_p[on:on + sz].view(shape)
            p.grad = new_g[on:on + sz].view(shape)
            p._trase_arena_grad_ptr = p.grad.data_ptr()
            ends.append(on + sz)
        arena.offsets, arena.flat = offs, new_g
        arena.n_early = sum(sz for sz in sizes[:arena.n_early_params])
        opt.flat_p, opt.exp_avg, opt.exp_avg_sq, opt._seg_end = new_p, new_m, new_v, ends
    for p in _gaussian_tensors(pc):                     # per-Gaussian tensors outside the arena (frozen ones)
        if id(p) not in in_arena:
            p.data = p.data[src].contiguous()
            p.grad = None
    pc.feature_smooth_map = None                        # the neighbour map is stale (scene/gaussian_model.py:632)


@torch.no_grad()
def prune_points(pc, opt, state: DensifyState, mask: torch.Tensor) -> None:
    """GaussianModel.prune_points (scene/gaussian_model.py:494-511): drop the Gaussians where `mask` is True."""
    keep = (~mask).nonzero().reshape(-1)
    regather(pc, opt, keep, torch.zeros_like(keep, dtype=torch.uint8))
    state.xyz_gradient_accum = state.xyz_gradient_accum[keep]
    state.denom = state.denom[keep]
    state.max_radii2D = state.max_radii2D[keep]


@torch.no_grad()
def densify_and_prune(pc, opt, state: DensifyState, max_grad: float, min_opacity: float, extent: float,
                      max_screen_size, N: int = 2, generator=None) -> Tuple[int, int]:
    """GaussianModel.densify_and_prune (scene/gaussian_model.py:617-635) = densify_and_clone (:594-615) +
    densify_and_split (:563-592) + prune_points, evaluated as index arithmetic on the CURRENT tensors followed by one
    gather.  Returns (number cloned, number split) like the reference."""
    dev = pc._xyz.device
    P = int(pc._xyz.shape[0])
    grads = state.xyz_gradient_accum / state.denom
    grads[grads.isnan()] = 0.0
    scal = torch.exp(pc._scaling.detach())                                   # get_scaling
    big = scal.max(dim=1).values > state.percent_dense * extent
    # clone: small Gaussians with a large view-space gradient; the copies are appended (:594-613)
    clone_idx = ((grads.norm(dim=-1) >= max_grad) & ~big).nonzero().reshape(-1)
    n_clone = int(clone_idx.numel())
    src1 = torch.cat([torch.arange(P, device=dev), clone_idx])
    fresh1 = torch.cat([torch.zeros(P, dtype=torch.bool, device=dev), torch.ones(n_clone, dtype=torch.bool, device=dev)])
    # split: large Gaussians with a large gradient, selected over originals + clones with the clones' gradient
    # padded with zeros (:565-570); N children sampled from each, the selected ones removed afterwards (:589-590)
    padded = torch.cat([grads.reshape(-1), torch.zeros(n_clone, device=dev)])
    split_mask = (padded >= max_grad) & big[src1]
    sel = src1[split_mask]                                                  # rows of the current tensors
    n_split = int(sel.numel())
    stds = scal[sel].repeat(N, 1)
    samples = torch.normal(mean=torch.zeros_like(stds), std=stds, generator=generator)
    rots = build_rotation(pc._rotation.detach()[sel]).repeat(N, 1, 1)
    child_xyz = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + pc._xyz.detach()[sel].repeat(N, 1)
    child_scaling = torch.log(scal[sel].repeat(N, 1) / (0.8 * N))
    n_child = N * n_split
    src2 = torch.cat([src1, sel.repeat(N)])
    fresh2 = torch.cat([fresh1, torch.ones(n_child, dtype=torch.bool, device=dev)])
    child2 = torch.cat([torch.full((P + n_clone,), -1, dtype=torch.int64, device=dev),
                        torch.arange(n_child, device=dev)])
    keep2 = ~torch.cat([split_mask, torch.zeros(n_child, dtype=torch.bool, device=dev)])
    src3, fresh3, child3 = src2[keep2], fresh2[keep2], child2[keep2]
    # prune: transparent Gaussians, and with a size limit those too large in world space; max_radii2D was reset to
    # zero by both densification_postfix calls (:560), so the screen-size test cannot fire here (:624-627)
    prune = torch.sigmoid(pc._opacity.detach()[src3]).reshape(-1) < min_opacity
    if max_screen_size:
        smax = scal[src3].max(dim=1).values
        is_child = child3 >= 0
        smax = torch.where(is_child, torch.exp(child_scaling[child3.clamp_min(0)]).max(dim=1).values, smax)
        prune = prune | (smax > 0.1 * extent)
    keep3 = ~prune
    src4, fresh4, child4 = src3[keep3], fresh3[keep3], child3[keep3]
    regather(pc, opt, src4, fresh4)
    rows = (child4 >= 0).nonzero().reshape(-1)
    if rows.numel():
        pc._xyz.data[rows] = child_xyz[child4[rows]]
        pc._scaling.data[rows] = child_scaling[child4[rows]]
    state.reset(int(src4.numel()), dev)
    return n_clone, n_split


@torch.no_grad()
def reset_opacity(pc, opt) -> None:
    """GaussianModel.reset_opacity (scene/gaussian_model.py:378-381): opacity <- min(opacity, 0.01) in logit space,
    its Adam moments zeroed (replace_tensor_to_optimizer, :457-470)."""
    o = torch.sigmoid(pc._opacity.detach()).clamp_max(0.01)
    pc._opacity.data.copy_(torch.log(o / (1 - o)))
    if opt is not None:
        for p, off in zip(opt.arena.params, opt.arena.offsets):
            if p is pc._opacity:
                opt.exp_avg[off:off + p.numel()].zero_()
                opt.exp_avg_sq[off:off + p.numel()].zero_()


# ---------------------------------------------------------------------------------------------------------
def capture(pc, opt: Optional[ArenaAdam], state: DensifyState, spatial_lr_scale: float = 1.0):
    """GaussianModel.capture (scene/gaussian_model.py:106-121): the tuple train.py:394-396 saves as
    `torch.save((gaussians.capture(), iteration), model_path + "/chkpnt" + str(iteration) + ".pth")`.
    Element 11 is the optimizer state.  (The reference calls `.state_dict()` on its {"GAUSSIAN", "FEATURE"} dict of
    optimizers there, which raises; written here as the dict of the two torch.optim.Adam state_dicts that call was
    meant to produce, so each half loads into a stock torch.optim.Adam.)"""
    return (pc.active_sh_degree, pc._xyz, pc._features_dc, pc._features_rest, pc._scaling, pc._rotation, pc._opacity,
            pc._gaussian_features, state.max_radii2D, state.xyz_gradient_accum, state.denom,
            optimizer_state_dict(opt) if opt is not None else None, spatial_lr_scale)


FEATURE_GROUPS = ("gaussian_feats", "_gaussian_features", "cls")


def optimizer_state_dict(opt: ArenaAdam) -> dict:
    """{"GAUSSIAN": sd, "FEATURE": sd} in torch.optim.Adam's state_dict layout (one parameter per group, in group
    order, as scene/gaussian_model.py:259-286 builds them)."""
    out = {"GAUSSIAN": {"state": {}, "param_groups": []}, "FEATURE": {"state": {}, "param_groups": []}}
    seg = {id(p): (off, p.numel()) for p, off in zip(opt.arena.params, opt.arena.offsets)}
    for gi, g in enumerate(opt.param_groups):
        half = out["FEATURE" if g.get("name") in FEATURE_GROUPS else "GAUSSIAN"]
        idxs = []
        first = sum(len(x["params"]) for x in half["param_groups"])          # parameters are numbered per optimizer
        for p in g["params"]:
            k = first + len(idxs)
            idxs.append(k)
            if id(p) in seg and opt._steps[gi] > 0:
                off, n = seg[id(p)]
                half["state"][k] = {"step": torch.tensor(float(opt._steps[gi])),
                                    "exp_avg": opt.exp_avg[off:off + n].view_as(p).clone(),
                                    "exp_avg_sq": opt.exp_avg_sq[off:off + n].view_as(p).clone()}
        pg = {k: v for k, v in g.items() if k != "params"}
        pg.update(params=idxs, betas=opt.betas, eps=opt.eps, weight_decay=0, amsgrad=False, maximize=False,
                  foreach=None, capturable=False, differentiable=False, fused=None)
        half["param_groups"].append(pg)
    return out


@torch.no_grad()
def load_optimizer_state_dict(opt: ArenaAdam, sd: dict) -> None:
    cursor = {"GAUSSIAN": 0, "FEATURE": 0}
    seg = {id(p): (off, p.numel()) for p, off in zip(opt.arena.params, opt.arena.offsets)}
    for gi, g in enumerate(opt.param_groups):
        which = "FEATURE" if g.get("name") in FEATURE_GROUPS else "GAUSSIAN"
        half = sd[which]
        for p in g["params"]:
            st = half["state"].get(cursor[which])
            cursor[which] += 1
            if st is None or id(p) not in seg:
                continue
            off, n = seg[id(p)]
            opt.exp_avg[off:off + n].copy_(st["exp_avg"].reshape(-1))
            opt.exp_avg_sq[off:off + n].copy_(st["exp_avg_sq"].reshape(-1))
            opt._steps[gi] = int(float(st["step"]))
    # learning rates travel with the groups
    for which in ("GAUSSIAN", "FEATURE"):
        names = {pg.get("name"): pg for pg in sd[which]["param_groups"]}
        for g in opt.param_groups:
            if g.get("name") in names:
                g["lr"] = names[g["name"]]["lr"]


@torch.no_grad()
def restore(pc, opt: Optional[ArenaAdam], state: DensifyState, model_args) -> float:
    """GaussianModel.restore (scene/gaussian_model.py:123-151): load a captured tuple (e.g. torch.load(chkpnt)[0]).
    Tensors are resized to the checkpoint's Gaussian count first.  Returns spatial_lr_scale."""
    (pc.active_sh_degree, xyz, f_dc, f_rest, scaling, rotation, opacity, feats, max_radii2D, accum, denom, opt_dict,
     spatial_lr_scale) = model_args
    P_new = int(xyz.shape[0])
    dev = pc._xyz.device
    if P_new != pc._xyz.shape[0]:
        idx = torch.arange(P_new, device=dev) % max(1, int(pc._xyz.shape[0]))
        regather(pc, opt, idx, torch.zeros(P_new, dtype=torch.uint8, device=dev))
    for p, v in zip(_gaussian_tensors(pc), (xyz, f_dc, f_rest, scaling, rotation, opacity, feats)):
        p.data.copy_(v.detach().to(dev))
    state.max_radii2D = max_radii2D.to(dev).clone()
    state.xyz_gradient_accum = accum.to(dev).clone()
    state.denom = denom.to(dev).clone()
    if opt is not None and opt_dict is not None:
        load_optimizer_state_dict(opt, opt_dict)
    return spatial_lr_scale
