#!/usr/bin/env python
"""bench.py — fwd+bwd RGB+feature render steps/s on BASELINE.json's configs.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME]

Workloads (BASELINE.json `configs`):
  neu3d       configs[1] (default, the headline line): 300k Gaussians, 1352x1014, DeformNetwork MLP + RGB+feature
              fwd+bwd, GAUSSIAN training state.  One step = one view.
  hypernerf8  configs[2]: 200k Gaussians, 536x960, FEATURE training state: MLP under no_grad, kNN-smoothed features,
              soft-mined contrastive feature loss on ~5000 sampled pixels; one step = 8 views per GPU.
  nnfm        configs[4]: the style-transfer (NNFM) loop shape of config 5: MLP under no_grad, SH-colour render,
              feature extractor, nearest-neighbour feature loss; only dL/dSH is kept and all-reduced.
  inference2m configs[3]: 2M Gaussians 1920x1080 forward only (views/s).
  tiny        configs[0].

At N > 1 every rank renders its own views (weak scaling) and the gradient arena of the state's trainable parameters
is all-reduced with NCCL inside the step.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "fwd+bwd RGB+feature render steps/sec"
UNIT = "steps/s"
DTYPE = "f32 rasterizer; DeformNetwork MLP tf32 operands / fp32 accumulate, weight gradients from bf16 side copies"
WORKLOADS = {  # BASELINE.json configs
    "neu3d": dict(P=300_000, W=1352, H=1014, seed=2, views=1, state="GAUSSIAN",
                  desc="configs[1]: Neu3D-shaped 300k Gaussians 1352x1014, DeformNetwork MLP + RGB+feature fwd+bwd"),
    "tiny": dict(P=20_000, W=128, H=128, seed=1, views=1, state="GAUSSIAN", desc="configs[0]: tiny 20k Gaussians 128x128"),
    "inference2m": dict(P=2_000_000, W=1920, H=1080, seed=4, views=1, state="INFERENCE",
                        desc="configs[3]: 2M Gaussians 1920x1080 feature-field inference (DeformNetwork + "
                             "RGB+feature+depth forward, no backward)"),
    "hypernerf8": dict(P=200_000, W=536, H=960, seed=3, views=8, state="FEATURE",
                       desc="configs[2]: HyperNeRF-shaped 200k Gaussians 536x960, FEATURE state: MLP under no_grad, "
                            "kNN-smoothed features (K=16, dropout 0.5), soft-mined contrastive loss on ~5000 sampled "
                            "pixels of 25 synthetic 268x480 masks + feature-norm regulariser, 8 jittered views per step"),
    "nnfm": dict(P=300_000, W=1352, H=1014, seed=2, views=1, state="STYLE",
                 desc="configs[4]: NNFM style-transfer loop shape (train_style_transfer_nnfm.py:141-283): MLP under "
                      "no_grad, SH-colour render, fixed random 3->512 stride-8 conv stack standing in for VGG16 "
                      "(no weights offline), loss_nnfm_style against a 256x256 style image, only dL/dSH kept"),
}


# --------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the run: NVML in-process every 10 ms (nvidia-smi every 0.2 s as a
    fallback).  Started before warm-up and stopped after the last timed region, so even a 0.1 s timed region is
    covered by tens of samples; `mark()` brackets the timed regions so the line can say how many fell inside."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
               ("sw_power_cap", 0x4))

    def __init__(self, device: torch.device):
        self.rows, self._stop, self._t = [], threading.Event(), None
        self._in_region = False
        self.h, self.nv = None, None
        self.index = device.index or 0
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device).uuid)
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.nv = pynvml
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h, self.nv, self.sm_max = None, None, None

    def _sample_nvml(self):
        nv = self.nv
        sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
        try:
            mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
        except Exception:
            mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
        self.rows.append((sm, mask, self._in_region))

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5)
        if out.returncode == 0 and out.stdout.strip():
            r = [x.strip() for x in out.stdout.strip().split(",")]
            self.sm_max = float(r[1])
            mask = sum(bit for (n, bit), v in zip(self.REASONS, r[2:6]) if v.lower().startswith("active"))
            self.rows.append((float(r[0]), mask, self._in_region))

    def _run(self):
        while not self._stop.is_set():
            try:
                if self.h is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop.wait(0.01 if self.h is not None else 0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def stop(self):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=6)

    def __enter__(self):           # `with sampler:` brackets a timed region
        self._in_region = True
        return self

    def __exit__(self, *a):
        self._in_region = False
        return False

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"], "samples": 0}
        inside = [r for r in self.rows if r[2]] or self.rows
        sm = sorted(r[0] for r in inside)
        mask = 0
        for r in inside:
            mask |= r[1]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.sm_max, "reasons": [n for n, b in self.REASONS if mask & b],
                "samples": len(self.rows), "samples_in_timed_regions": len([r for r in self.rows if r[2]]),
                "source": "nvml 10 ms" if self.h is not None else "nvidia-smi 0.2 s"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1354.9), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1354.9, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------
def rect_masks(n, Hm, Wm, gen):
    m = torch.zeros(n, Hm, Wm, dtype=torch.uint8)
    for i in range(n):
        h = int(torch.randint(Hm // 8, Hm // 2, (1,), generator=gen))
        w = int(torch.randint(Wm // 8, Wm // 2, (1,), generator=gen))
        y = int(torch.randint(0, Hm - h, (1,), generator=gen))
        x = int(torch.randint(0, Wm - w, (1,), generator=gen))
        m[i, y:y + h, x:x + w] = 1
    return m


class Workload:
    """Synthetic scene, raw parameters, DeformNetwork, cameras, and the per-state step.

    `ref` (reference arm only) = namespace(rasterizer_cls, utils): the unmodified reference rasterizer and the
    reference's own Python (DeformNetwork, losses), all loaded from oracle/_ref."""

    def __init__(self, name, device, rank=0, mlp_backend="native", ref=None):
        from trase_b200 import synth
        from trase_b200.deform import DeformNetwork
        from trase_b200.render import GaussianParams, make_view
        cfg = WORKLOADS[name]
        self.name, self.cfg, self.device, self.ref, self.rank = name, cfg, device, ref, rank
        self.state = cfg["state"]
        P, W, H = cfg["P"], cfg["W"], cfg["H"]
        self.nviews = cfg["views"]
        self.scene = synth.make_scene(P, W, H, cfg["seed"])
        self.pc = GaussianParams.from_activated(self.scene, device)
        torch.manual_seed(1234)
        ours = DeformNetwork(backend="native" if mlp_backend == "native" else "torch").to(device)
        with torch.no_grad():  # keep the deformation small, as in a trained model (random heads would blow R up)
            for head in (ours.gaussian_warp, ours.gaussian_rotation, ours.gaussian_scaling):
                head.weight.mul_(1e-2)
                head.bias.mul_(1e-2)
        if ref is not None and ref.utils is not None:
            # reference arm: the reference's own DeformNetwork class (utils/time_utils.py), same weights
            self.deform = ref.utils.time_utils.DeformNetwork().to(device)
            self.deform.load_state_dict(ours.state_dict())
        else:
            self.deform = ours
        # cameras: rank r's view v; config 2/5 yaw a little per rank, config 3 jitters +-0.2 rad about the centre
        self.cams, self.views, self.times = [], [], []
        for v in range(self.nviews):
            gv = rank * self.nviews + v
            if name == "hypernerf8":
                g = torch.Generator().manual_seed(300 + gv)
                yaw, pitch = ((torch.rand(2, generator=g) - 0.5) * 0.4).tolist()
                wv = synth.orbit_world_view(yaw, pitch) if gv else torch.eye(4)
            else:
                wv = synth.orbit_world_view(0.02 * gv, 0.0) if gv else torch.eye(4)
            cam = synth.make_camera(W, H, wv)
            self.cams.append(cam)
            self.views.append(make_view(cam, device))
            self.times.append(0.37 + 0.1 * gv)
        self.cam, self.view = self.cams[0], self.views[0]
        # training: a materialised [P,1] tensor (train.py:202 adds per-iteration noise to the expanded time stamp);
        # no_grad states and inference: the reference's own broadcast view, fid.unsqueeze(0).expand(N, -1)
        if self.state == "GAUSSIAN":
            self.time_t = [torch.full((P, 1), t, device=device) for t in self.times]
        else:
            self.time_t = [torch.full((1, 1), t, device=device).expand(P, 1) for t in self.times]
        self.time = self.time_t[0]
        self.bg = torch.zeros(3, device=device)
        self.copy_stream = torch.cuda.Stream(device=device)
        self.loss_host = torch.zeros(max(1, self.nviews)).pin_memory()
        self.loss_event = torch.cuda.Event()
        cam_blocks = [torch.cat([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1), c["campos"].reshape(-1),
                                 torch.tensor([t])]) for c, t in zip(self.cams, self.times)]
        self.cam_host = [b.pin_memory() for b in cam_blocks]
        self.cam_dev = [torch.empty_like(b, device=device) for b in cam_blocks]
        getattr(self, "_setup_" + self.state.lower())()
        if self.state != "INFERENCE":
            from trase_b200.distributed import GradArena
            # one flat gradient arena over the state's trainable parameters; every .grad is a view into it
            self.grad_arena = GradArena(self.params, early=[p for p in self.pc.parameters() if p.requires_grad])
            self.arena = self.grad_arena.flat

    # ---- per-state set-up ------------------------------------------------
    def _setup_inference(self):
        self.params = []

    def _setup_gaussian(self):
        """GAUSSIAN state (scene/gaussian_model.py:303-315): everything but the segmentation feature trains; the
        feature parameter has requires_grad=False, so its 32 floats/Gaussian are neither accumulated nor reduced."""
        from trase_b200 import synth
        W, H, P = self.cfg["W"], self.cfg["H"], self.cfg["P"]
        up = synth.make_upstream_grads(W, H, self.cfg["seed"] + self.rank)
        self.dL_dcolor, self.dL_dobjects = up["dL_dcolor"].to(self.device), up["dL_dobjects"].to(self.device)
        if not self.full_arena:
            self.pc._gaussian_features.requires_grad_(False)
        self.params = [p for p in self.pc.parameters() if p.requires_grad] + list(self.deform.parameters())
        g = torch.Generator().manual_seed(99 + self.rank)
        self.gt_host = torch.rand(3, H, W, generator=g).pin_memory()
        self.gt_dev = torch.empty(3, H, W, device=self.device)

    full_arena = False     # True: keep the round-1 arena (feature gradient included) for comparison runs
    fused_loss = True      # FEATURE state: the fused contrastive-loss kernels (False: the torch N x N formulation)

    def _setup_feature(self):
        """FEATURE state (train.py:244-296): only the per-Gaussian segmentation feature trains."""
        from trase_b200.feature_utils import build_feature_smooth_map
        for p in list(self.pc.parameters()) + list(self.deform.parameters()):
            p.requires_grad_(False)
        self.pc._gaussian_features.requires_grad_(True)
        self.params = [self.pc._gaussian_features]
        self.knn = build_feature_smooth_map(self.scene["means3D"], K=16).to(self.device)   # host k-d tree, set-up only
        self.pc.set_feature_smooth_map(self.knn)
        Hm, Wm = self.cfg["H"] // 2, self.cfg["W"] // 2
        self.mask_hw = (Hm, Wm)
        self.masks_host = [rect_masks(25, Hm, Wm, torch.Generator().manual_seed(700 + self.rank * self.nviews + v)
                                      ).pin_memory() for v in range(self.nviews)]
        self.masks_dev = [m.to(self.device) for m in self.masks_host]
        self.masks_stage = [torch.empty_like(m, device=self.device) for m in self.masks_host]
        self.num_sampled_pixels, self.num_sampled_masks, self.rfn = 5000, 25, 1.0
        self.sample_gen = torch.Generator(device=self.device).manual_seed(4242 + self.rank)
        self.sample_stream = torch.cuda.Stream(device=self.device)
        self.mask_free = {}
        self.last_samples = 0

    def _setup_style(self):
        """Style-transfer loop (train_style_transfer_nnfm.py; scene/gaussian_model.py:266-271 'style_transfer'
        optimizer groups): only f_dc / f_rest are optimised, so only dL/dSH is kept (192 B/Gaussian)."""
        for p in list(self.pc.parameters()) + list(self.deform.parameters()):
            p.requires_grad_(False)
        self.pc._features_dc.requires_grad_(True)
        self.pc._features_rest.requires_grad_(True)
        self.params = [self.pc._features_dc, self.pc._features_rest]
        torch.manual_seed(77)
        C = torch.nn.Conv2d
        # stand-in for VGG16 up to a stride-8 block (no weights offline): fixed random 3 -> 512, stride 8
        self.fx = torch.nn.Sequential(C(3, 64, 3, 2, 1), torch.nn.ReLU(), C(64, 128, 3, 2, 1), torch.nn.ReLU(),
                                      C(128, 256, 3, 2, 1), torch.nn.ReLU(), C(256, 512, 3, 1, 1), torch.nn.ReLU()
                                      ).to(self.device)
        for p in self.fx.parameters():
            p.requires_grad_(False)
        self.style_host = torch.rand(3, 256, 256, generator=torch.Generator().manual_seed(5)).pin_memory()
        self.style_dev = self.style_host.to(self.device)
        self.style_stage = torch.empty_like(self.style_dev)
        self.img_mean = torch.tensor([0.485, 0.456, 0.406], device=self.device).view(3, 1, 1)
        self.img_std = torch.tensor([0.229, 0.224, 0.225], device=self.device).view(3, 1, 1)
        self.seg_mask = (torch.rand(self.cfg["P"], generator=torch.Generator().manual_seed(6)) < 0.5).to(self.device)

    # ---- pieces ------------------------------------------------------------
    def mlp(self, time_t):
        xyz = self.pc.get_xyz.detach()
        if self.state == "GAUSSIAN":
            return self.deform(xyz, time_t)
        with torch.no_grad():                                   # train.py:200-202, train_style_transfer_nnfm.py:186
            return self.deform(xyz, time_t)

    def render(self, view, time_t, **kw):
        from trase_b200 import render as R
        d_xyz, d_rot, d_scale = self.mlp(time_t)
        if self.ref is None:
            return R.render(view, self.pc, R.PIPE_DEFAULT, self.bg, d_xyz, d_rot, d_scale, **kw)
        # reference arm: the reference's own GaussianRasterizer and its elementwise torch glue (render() mirrors
        # gaussian_renderer/__init__.py:37-155, whose own module cannot be imported without the scene package)
        saved = R.GaussianRasterizer, R.GaussianRasterizationSettings, R.FUSED_GLUE
        R.GaussianRasterizer, R.GaussianRasterizationSettings = self.ref.rasterizer_cls
        R.FUSED_GLUE = False
        try:
            return R.render(view, self.pc, R.PIPE_DEFAULT, self.bg, d_xyz, d_rot, d_scale, **kw)
        finally:
            R.GaussianRasterizer, R.GaussianRasterizationSettings, R.FUSED_GLUE = saved

    def _host_view(self, v):
        """Camera block of view v copied from pinned host memory; returns (view namespace, time tensor)."""
        self.cam_dev[v].copy_(self.cam_host[v], non_blocking=True)
        c, base = self.cam_dev[v], self.views[v]
        view = SimpleNamespace(FoVx=base.FoVx, FoVy=base.FoVy, image_height=base.image_height,
                               image_width=base.image_width, world_view_transform=c[0:16].view(4, 4),
                               full_proj_transform=c[16:32].view(4, 4), camera_center=c[32:35])
        return view, c[35:36].expand(self.cfg["P"], 1)

    # ---- GAUSSIAN state ------------------------------------------------------
    def _gaussian_resident(self):
        self.arena.zero_()
        out = self.render(self.view, self.time)
        self.grad_arena.arm()
        torch.autograd.backward([out["render"], out["render_gaussian_features"]], [self.dL_dcolor, self.dL_dobjects])
        self.grad_arena.all_reduce()
        return out

    def _gaussian_e2e(self):
        """H2D of this view's camera block and ground-truth image, the reference's L1 + SSIM colour loss and
        feature-norm regulariser (train.py:235-238, 280-281), backward, D2H of the loss."""
        main = torch.cuda.current_stream()
        view, t = self._host_view(0)
        # the ground-truth image is only read by the loss: its copy runs on a side stream under the forward pass
        self.copy_stream.wait_stream(main)      # the previous step's loss has finished reading gt_dev
        with torch.cuda.stream(self.copy_stream):
            self.gt_dev.copy_(self.gt_host, non_blocking=True)
        self.arena.zero_()
        out = self.render(view, t)
        main.wait_stream(self.copy_stream)
        if self.ref is None:
            from trase_b200 import loss_utils          # fused kernels (csrc/loss.cu)
            loss = loss_utils.l1_ssim_loss(out["render"], self.gt_dev, 0.2) + \
                loss_utils.feature_norm_reg(out["render_gaussian_features"])
        else:
            lu = self.ref.utils.loss_utils             # the reference's own l1_loss / ssim
            loss = 0.8 * lu.l1_loss(out["render"], self.gt_dev) + 0.2 * (1.0 - lu.ssim(out["render"], self.gt_dev)) + \
                (1.0 - out["render_gaussian_features"].norm(dim=0, p=2).mean()) ** 2
        # the loss goes back to the host as soon as it exists; the host waits for THAT copy, not for the backward
        # pass, exactly as a training loop that logs its loss (the timed region still ends with a full synchronize)
        self.loss_host[0:1].copy_(loss.detach().reshape(1), non_blocking=True)
        self.loss_event.record()
        self.grad_arena.arm()
        loss.backward()
        self.grad_arena.all_reduce()
        self.loss_event.synchronize()
        return float(self.loss_host[0])

    # ---- FEATURE state (config 3) ---------------------------------------------
    def _feature_sample(self, masks, masks_host=None):
        """train.py:253 for one view: the sampled pixels and masks.  It does not depend on the render, so it runs on a
        side stream and is issued one view AHEAD (right before the previous view's backward is launched): the host
        synchronisation inside it — the sample count comes back to size the loss kernels — is then waited out while
        the GPU still has that view's forward queued, not with the GPU idle."""
        if self.ref is not None:
            return None                                     # the reference arm samples inline, as train.py does
        from trase_b200 import feature_utils as FU
        main = torch.cuda.current_stream()
        with torch.cuda.stream(self.sample_stream):
            if masks_host is not None:                      # e2e: this view's masks come from pinned host memory
                ev = self.mask_free.get(masks.data_ptr())   # (the previous step's loss has finished reading them)
                if ev is not None:
                    self.sample_stream.wait_event(ev)
                masks.copy_(masks_host, non_blocking=True)
            sp, sm = FU.get_sample_pixel_and_mask(masks, self.num_sampled_pixels, self.num_sampled_masks,
                                                  generator=self.sample_gen)
            pix = FU.sampled_pixel_indices(sp) if self.fused_loss else None
            done = torch.cuda.Event()
            done.record(self.sample_stream)
        for t_ in (sp, sm, pix):
            if t_ is not None:
                t_.record_stream(main)
        return sp, sm, pix, done

    def _feature_view_loss(self, view, time_t, masks, masks_host=None, samples=None):
        """train.py:244-296 for one view; returns the loss tensor."""
        if self.ref is None:
            from trase_b200 import feature_utils as FU, loss_utils as LU
            main = torch.cuda.current_stream()
            sp, sm, pix, done = samples if samples is not None else self._feature_sample(masks, masks_host)
            out = self.render(view, time_t, norm_gaussian_features=True, is_smooth_gaussian_features=True, smooth_K=16,
                              want_viewspace_grad=False)   # past densify_until_iter nobody reads viewspace_points.grad
            feats = out["render_gaussian_features"]
            reg = LU.feature_norm_reg(feats)
            main.wait_event(done)                          # the samples (and, end to end, the masks) are read from here on
            if self.fused_loss:
                # mask matrix, pair weights, similarities and both losses without any N x N tensor (csrc/contrastive.cu)
                pos, neg = FU.soft_contrastive_losses(feats, masks, sp, sm, positive_th=0.75, negative_th=0.5, pix=pix)
                self.last_samples = FU.LAST_NUM_SAMPLES
                if masks_host is not None:
                    self.mask_free.setdefault(masks.data_ptr(), torch.cuda.Event()).record(main)
                return pos + neg + self.rfn * reg
            C = FU.get_pixel_mask_correspondence_matrix(masks, sp, sm)
            C_F = FU.get_features_correspondence_matrix(feats, sp)        # bilinear resize folded into the gather
            w = FU.get_pixel_weights(masks, sp)
            loss = LU.pixel_mask_correspondence_loss_soft_hard_positive(C=C, C_F=C_F, positive_th=0.75, weights=w) + \
                LU.pixel_mask_correspondence_loss_soft_negative(C=C, C_F=C_F, negative_th=0.5, weights=w) + \
                self.rfn * reg
            self.last_samples = C_F.shape[0]
            return loss
        ru = self.ref.utils
        FU, LU = ru.feature_utils, ru.loss_utils
        if masks_host is not None:
            masks.copy_(masks_host, non_blocking=True)
        sp, sm = FU.get_sample_pixel_and_mask(masks, self.num_sampled_pixels, self.num_sampled_masks)
        C = FU.get_pixel_mask_correspondence_matrix(masks, sp, sm)
        out = self.render(view, time_t, norm_gaussian_features=True, is_smooth_gaussian_features=True, smooth_K=16,
                              want_viewspace_grad=False)   # past densify_until_iter nobody reads viewspace_points.grad
        feats = out["render_gaussian_features"]
        reg = (1 - feats.norm(dim=0, p=2).mean()) ** 2
        resized = torch.nn.functional.interpolate(feats.unsqueeze(0), masks.shape[-2:], mode="bilinear").squeeze(0)
        C_F = FU.get_features_correspondence_matrix(resized, sp)
        w = FU.get_pixel_weights(masks, sp)
        loss = LU.pixel_mask_correspondence_loss_soft_hard_positive(C=C, C_F=C_F, positive_th=0.75, weights=w) + \
            LU.pixel_mask_correspondence_loss_soft_negative(C=C, C_F=C_F, negative_th=0.5, weights=w) + self.rfn * reg
        self.last_samples = C_F.shape[0]
        return loss

    def _feature_resident(self):
        self.arena.zero_()
        samples = self._feature_sample(self.masks_dev[0])
        for v in range(self.nviews):
            loss = self._feature_view_loss(self.views[v], self.time_t[v], self.masks_dev[v], samples=samples)
            if v + 1 < self.nviews:
                samples = self._feature_sample(self.masks_dev[v + 1])
            else:
                self.grad_arena.arm()
            loss.backward()
        self.grad_arena.all_reduce()

    def _feature_e2e(self):
        """Per view: H2D of the camera block and of the view's 25 masks (uint8, 268x480), D2H of the loss."""
        self.arena.zero_()
        samples = self._feature_sample(self.masks_stage[0], self.masks_host[0])
        for v in range(self.nviews):
            view, t = self._host_view(v)
            loss = self._feature_view_loss(view, t, self.masks_stage[v], self.masks_host[v], samples=samples)
            self.loss_host[v:v + 1].copy_(loss.detach().reshape(1), non_blocking=True)
            if v + 1 < self.nviews:
                samples = self._feature_sample(self.masks_stage[v + 1], self.masks_host[v + 1])
            else:
                self.loss_event.record()
                self.grad_arena.arm()
            loss.backward()
        self.grad_arena.all_reduce()
        self.loss_event.synchronize()
        return float(self.loss_host[self.nviews - 1])

    # ---- style-transfer loop (config 5) ------------------------------------------
    def _style_loss(self, view, time_t, style_img):
        out = self.render(view, time_t)
        image = out["render"]
        f_r = self.fx(((image - self.img_mean) / self.img_std).unsqueeze(0)).squeeze(0)
        f_s = self.fx(((style_img - self.img_mean) / self.img_std).unsqueeze(0)).squeeze(0)   # recomputed per step, as the reference does
        nnfm = self.ref.utils.loss_utils.loss_nnfm_style if self.ref is not None else self._nnfm()
        return nnfm(f_r.view(f_r.shape[0], -1), f_s.view(f_s.shape[0], -1))

    @staticmethod
    def _nnfm():
        from trase_b200.loss_utils import loss_nnfm_style
        return loss_nnfm_style

    def _style_finish(self, loss):
        self.grad_arena.arm()
        loss.backward()
        # gaussians.set_background_zero_grad(segmented_mask) (scene/gaussian_model.py:155-157)
        keep = self.seg_mask.view(-1, 1, 1)
        self.pc._features_dc.grad.mul_(keep)
        self.pc._features_rest.grad.mul_(keep)
        self.grad_arena.all_reduce()

    def _style_resident(self):
        self.arena.zero_()
        self._style_finish(self._style_loss(self.view, self.time, self.style_dev))

    def _style_e2e(self):
        view, t = self._host_view(0)
        self.style_stage.copy_(self.style_host, non_blocking=True)
        self.arena.zero_()
        loss = self._style_loss(view, t, self.style_stage)
        self.loss_host[0:1].copy_(loss.detach().reshape(1), non_blocking=True)
        self.loss_event.record()
        self._style_finish(loss)
        self.loss_event.synchronize()
        return float(self.loss_host[0])

    # ---- dispatch -----------------------------------------------------------
    def step_resident(self):
        return getattr(self, "_" + self.state.lower() + "_resident")()

    def step_e2e(self):
        return getattr(self, "_" + self.state.lower() + "_e2e")()

    def h2d_bytes(self):
        cam = sum(c.numel() * 4 for c in self.cam_host[:self.nviews])
        if self.state == "GAUSSIAN":
            return self.gt_host.numel() * 4 + cam
        if self.state == "FEATURE":
            return sum(m.numel() for m in self.masks_host) + cam
        return self.style_host.numel() * 4 + cam

    def d2h_bytes(self):
        return (4 + 4) * self.nviews  # per view: loss + the rasterizer's instance count


# --------------------------------------------------------------------------
def timed_region(fn, steps, warmup, world_size, device, sampler=None):
    """W untimed steps, then exactly K steps bracketed by barrier + synchronize;
    device time from CUDA events on the launching stream; MAX over ranks."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(device)
    if world_size > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx = sampler if sampler is not None else _Null()
    with ctx:
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize(device)
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize(device)
    ms = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
    if world_size > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


# SMs the persistent MLP kernels leave free when a gradient all-reduce runs under the MLP backward (N > 1)
# (N=8 sweep, profiles/r2_l_*: 0 SMs / default NCCL 1493.9 steps/s; 16 SMs / default NCCL 1443.9; 8 SMs with NCCL held
# to 8 CTAs 1507.7; 16 / 16 1496.4; 0 SMs with 8 CTAs 1404.4: the all-reduce has 1.3 ms of MLP backward to hide under, so
# a narrow collective on SMs of its own beats a wide one that time-shares the SMs of the persistent kernels)
RESERVE_SMS_MULTI = 8
NCCL_CTAS_MULTI = "8"
MLP_MAC_PER_ROW = 504_320          # 84*256 + 6*256*256 + 340*256 + 256*10


def algorithmic_bytes(stage, P, R, N, K=16, n_pass=6):
    """SURVEY.md §8(d).2 per-stage compulsory traffic (bytes per launch group)."""
    return {
        "preprocess_fwd": P * (44 + 12 * K) + 83 * P,
        "scan_duplicate": 8 * P + 12 * R,
        "sort": 24 * R * n_pass,
        "ranges": 16 * R,
        "render_fwd": 172 * R + 152 * N,
        "render_bwd": 148 * N + 172 * R + 172 * P,
        "preprocess_bwd": P * (44 + 12 * K + 24 + 172) + P * (12 + 12 * K + 12 + 16),
    }[stage]


def mlp_flops(stage, P):
    """DeformNetwork algorithmic flops per launch group: forward 2*504320*P; data-gradient chain = the seven hidden
    256x256 products plus the 10-wide heads (the first layer's input carries no gradient); weight gradients = one
    product per weight, same count as the forward."""
    return {"mlp_fwd": 2 * MLP_MAC_PER_ROW * P, "mlp_bwd_chain": 2 * (7 * 256 * 256 + 256 * 10) * P,
            "mlp_wgrad": 2 * MLP_MAC_PER_ROW * P}[stage]


def cpu_baseline_port(name, budget_s=30.0):
    """The CPU oracle (oracle/, a C port of the reference's algorithm) timed on this box's host cores: one full
    fwd+bwd rasterizer step of the same workload's first view plus the MLP in torch CPU fp32 (forward only where the
    state runs it under no_grad), scaled to the step's view count."""
    from oracle import oracle
    from trase_b200 import synth
    from trase_b200.deform import DeformNetwork
    cfg = WORKLOADS[name]
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sc = synth.make_scene(P, W, H, cfg["seed"])
    up = synth.make_upstream_grads(W, H, cfg["seed"])
    oracle.lib()
    t0 = time.perf_counter()
    f = oracle.forward_scene(sc, [0, 0, 0])
    if cfg["state"] != "INFERENCE":
        oracle.backward(f, up["dL_dcolor"].numpy(), up["dL_dobjects"].numpy())
    t_raster = time.perf_counter() - t0
    rows = min(P, 60_000)            # MLP on a bounded row sample, scaled to P rows
    net = DeformNetwork()
    x = sc["means3D"][:rows]
    t = torch.full((rows, 1), 0.37)
    t0 = time.perf_counter()
    if cfg["state"] == "GAUSSIAN":
        outs = net(x, t)
        torch.autograd.backward(list(outs), [torch.ones_like(o) for o in outs])
    else:
        with torch.no_grad():
            net(x, t)
    t_mlp = (time.perf_counter() - t0) * (P / rows)
    total = (t_raster + t_mlp) * cfg["views"]
    return {"value": 1.0 / total, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "1 rasterizer %s view of %s with the C oracle (OpenMP, %.2f s) + DeformNetwork %s in torch CPU "
                      "fp32 on %d of %d rows scaled to all rows (%.2f s), x %d views per step; losses not included" %
                      ("forward" if cfg["state"] == "INFERENCE" else "fwd+bwd", name, t_raster,
                       "fwd+bwd" if cfg["state"] == "GAUSSIAN" else "forward", rows, P, t_mlp, cfg["views"])}


# --------------------------------------------------------------------------
def scene_statistics(wl):
    """V (visible Gaussians), R, contributors per pixel and the (pixel, Gaussian) pair count of one view, read from
    the workspaces of one extra untimed forward (the per-pixel contributor count is an output of the forward
    kernel, no counting kernel needed)."""
    from trase_b200 import _raster
    captured = {}
    real = _raster.rasterize_forward

    def spy(*a, **k):
        res = real(*a, **k)
        captured["res"], captured["P"] = res, int(a[1].shape[0])
        return res

    _raster.rasterize_forward = spy
    try:
        with torch.no_grad():
            wl.render(wl.view, wl.time)
    finally:
        _raster.rasterize_forward = real
    res, P = captured["res"], captured["P"]
    W, H = wl.cfg["W"], wl.cfg["H"]
    v = _raster.workspace_views(res.geom, res.img, res.binning, P, W, H, res.capacity, res.num_rendered)
    nc = v["n_contrib"].long()
    return {"V": int((res.radii > 0).sum()), "R": int(res.num_rendered), "contributors_per_pixel_mean":
            round(float(nc.float().mean()), 2), "contributors_per_pixel_max": int(nc.max()),
            "pairs_visited": int(nc.sum()),
            "pairs_note": "pairs_visited = sum over pixels of the index of the last contributing list entry "
                          "(img workspace n_contrib): the trip count of the backward kernel's per-pixel loop and, up to "
                          "the block-wide early exit, of the forward kernel's"}


def run_ours(args, rank, world, device):
    from trase_b200 import _lib, build as tb_build, _deform_native
    if not os.path.exists(_lib.LIB_PATH):
        tb_build.build()
    _lib.load()  # fails loudly when the CUDA library is missing: there is no fallback
    if args.mlp == "torch-tf32":
        torch.backends.cuda.matmul.allow_tf32 = True
    Workload.full_arena = bool(args.full_arena)
    Workload.fused_loss = not args.torch_feature_loss
    for key in ("fwd_ppt", "bwd_ppt", "fwd_mma"):                      # kernel tuning experiments: TRASE_B200_BWD_PPT=1 ...
        if os.environ.get("TRASE_B200_" + key.upper()):
            _lib.set_tuning(key, int(os.environ["TRASE_B200_" + key.upper()]))
    sampler = ClockSampler(device).start() if rank == 0 else None
    wl = Workload(args.workload, device, rank, mlp_backend="native" if args.mlp == "native" else "torch")
    cfg = wl.cfg
    step = wl.step_resident
    # launches of OUR kernels in the timed region (counted by the library itself); warm-up excluded
    for _ in range(args.warmup):
        step()
    l0 = _lib.launch_count()
    ms = timed_region(step, args.steps, 0, world, device, sampler)
    launches = _lib.launch_count() - l0
    value = world * args.steps / (ms / 1e3)

    # end to end: host buffers -> public API -> loss on the host
    ms_e2e = timed_region(wl.step_e2e, args.steps, max(3, args.warmup), world, device, sampler)
    e2e = world * args.steps / (ms_e2e / 1e3)
    if sampler:
        sampler.stop()
    clocks = sampler.summary() if sampler else None

    # per-stage device times inside the step (second pass with event pairs around each stage); every rank
    # runs it because the step contains the collective
    _lib.profile_enable(True)
    _deform_native.profile_enable(True)
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize(device)
    prof = _lib.profile_read()
    mlp_prof = _deform_native.profile_read()
    _lib.profile_enable(False)
    _deform_native.profile_enable(False)
    if rank != 0:
        return None
    P, N = cfg["P"], cfg["W"] * cfg["H"]
    T = ((cfg["W"] + 15) // 16) * ((cfg["H"] + 15) // 16)
    stats = scene_statistics(wl)
    R = stats["R"]
    # SURVEY §8(d).2 counts the reference's sort: 8-bit passes over all 32 + tile bits, whatever we launch
    n_pass = (32 + max(1, T.bit_length()) + 7) // 8
    peak, tc_peak, peak_src = measured_peaks()
    launches_per_stage = wl.nviews * args.steps          # one launch group of each raster stage per view
    stages = {}
    for k, (tot, cnt) in prof.items():
        if cnt:
            per = tot / launches_per_stage
            ab = algorithmic_bytes(k, P, R, N, 16, n_pass)
            stages[k] = {"ms": round(per, 4), "GBps": round(ab / per / 1e6, 1), "algorithmic_bytes": ab}
    for k in ("render_fwd", "render_bwd"):
        if k in stages:
            stages[k]["pairs_per_s"] = round(stats["pairs_visited"] / (stages[k]["ms"] * 1e-3), 1)
    for k, (tot, cnt) in mlp_prof.items():
        if cnt and k in ("mlp_fwd", "mlp_bwd_chain", "mlp_wgrad"):
            per = tot / launches_per_stage
            fl = mlp_flops(k, P)
            stages[k] = {"ms": round(per, 4), "flops": fl, "TFLOPs": round(fl / per / 1e9, 1),
                         "frac_of_bf16_sustained": round(fl / per / 1e9 / tc_peak, 4)}
    mlp_keys = [k for k in ("mlp_fwd", "mlp_bwd_chain", "mlp_wgrad") if k in stages]
    mlp_total = None
    if mlp_keys:
        t = sum(stages[k]["ms"] for k in mlp_keys)
        fl = sum(stages[k]["flops"] for k in mlp_keys)
        mlp_total = {"ms": round(t, 4), "flops": fl, "TFLOPs": round(fl / t / 1e9, 1),
                     "frac_of_bf16_sustained": round(fl / t / 1e9 / tc_peak, 4), "peak_TFLOPs": tc_peak,
                     "note": "tf32 operands: the tf32 tensor peak is half the bf16 figure used as the denominator"}
    raster = [k for k in stages if not k.startswith("mlp_")]
    dom = max(raster, key=lambda k: stages[k]["ms"]) if raster else None
    roof, traffic = None, None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(dom, {}).get("bytes")
    except Exception:
        pass
    if dom:
        ach = stages[dom]["GBps"]
        roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4), "traffic": traffic if args.workload == "neu3d" else None,
                "traffic_unit": "bytes per launch (ncu --set full, profiles/traffic.json)",
                "peak_source": peak_src, "algorithmic_bytes": stages[dom]["algorithmic_bytes"],
                "launch_ms": stages[dom]["ms"],
                "note": "compositing kernels are FP32/shuffle-bound on (pixel,Gaussian) pairs, not HBM-bound "
                        "(SURVEY.md §8d); HBM fraction is low by construction; see stages[*].pairs_per_s"}
    arena_desc = "%d-float gradient arena (%s state: %s)" % (
        wl.arena.numel(), wl.state, {"GAUSSIAN": "xyz, SH, opacity, scaling, rotation + MLP; the segmentation feature "
                                     "has requires_grad=False in this state", "FEATURE": "segmentation feature only",
                                     "STYLE": "SH only"}[wl.state] if not args.full_arena else "all parameters")
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": DTYPE,
        "data": "synthetic (SURVEY.md §8(d).1 generator, seed %d; random-init DeformNetwork with output heads "
                "scaled 1e-2)" % cfg["seed"],
        "config": {"workload": cfg["desc"], "P": P, "W": cfg["W"], "H": cfg["H"], "num_rendered": R,
                   "views_per_step_per_gpu": wl.nviews, "views_per_s": round(value * wl.nviews, 2), "mlp": args.mlp,
                   "state": wl.state, "parallelism": ("view-parallel dp%d, NCCL all-reduce of the " % world if
                                                      world > 1 else "single GPU; ") + arena_desc,
                   "l2": "no flush: per-step working set (parameters + 36 image planes x2 + workspaces) exceeds "
                         "the 126 MB L2"},
        "e2e": {"value": round(e2e, 3), "unit": UNIT, "h2d_bytes_per_step": wl.h2d_bytes(),
                "d2h_bytes_per_step": wl.d2h_bytes(), "ms_per_step": round(ms_e2e / args.steps, 4)},
        "gpu_launches": launches,
        "clocks": clocks, "roofline": roof, "stages": stages, "mlp": mlp_total, "scene": stats,
        "raster_only_ms": round(sum(stages[k]["ms"] for k in raster), 4),
    }
    if wl.state == "FEATURE":
        line["config"]["sampled_pixels_last_view"] = wl.last_samples
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_port(args.workload)
    return line


def run_inference(args, rank, world, device):
    """Extra, not the headline line: views/s of inference rendering (configs[3] shape); views shard
    round-robin over the ranks with no collective."""
    from trase_b200 import _lib
    _lib.load()
    wl = Workload(args.workload, device, rank, mlp_backend="native" if args.mlp == "native" else "torch")

    def step():
        with torch.no_grad():
            return wl.render(wl.view, wl.time)

    ms = timed_region(step, args.steps, args.warmup, world, device)
    if rank != 0:
        return None
    return {"metric": "inference views/sec (RGB+feature+depth)", "value": round(world * args.steps / (ms / 1e3), 3),
            "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": DTYPE, "data": "synthetic", "config": {"workload": wl.cfg["desc"], "mlp": args.mlp}}


def reference_raster_only(wl, device, iters=10):
    """Event time of the reference's own rasterize_gaussians / rasterize_gaussians_backward on the workload's scene
    (activated parameters, no deformation) — the rasterizer-only yardstick beside `raster_only_ms` of our arm."""
    from oracle import ref_loader
    from trase_b200 import synth
    sc = synth.scene_to(wl.scene, device)
    up = {k: v.to(device) for k, v in synth.make_upstream_grads(wl.cfg["W"], wl.cfg["H"], wl.cfg["seed"]).items()}
    f = ref_loader.run_forward(sc, wl.bg)
    ref_loader.run_backward(sc, wl.bg, f, up["dL_dcolor"], up["dL_dobjects"])
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    tf = tb = 0.0
    for _ in range(iters):
        ev[0].record()
        f = ref_loader.run_forward(sc, wl.bg)
        ev[1].record()
        ref_loader.run_backward(sc, wl.bg, f, up["dL_dcolor"], up["dL_dobjects"])
        ev[2].record()
        torch.cuda.synchronize(device)
        tf += ev[0].elapsed_time(ev[1])
        tb += ev[1].elapsed_time(ev[2])
    return {"fwd_ms": round(tf / iters, 4), "bwd_ms": round(tb / iters, 4), "total_ms": round((tf + tb) / iters, 4),
            "note": "reference _C.rasterize_gaussians + _backward on the static scene (includes its host sync and "
                    "torch allocations), random upstream gradients"}


def run_reference(args, rank, world, device):
    """Reference arm: the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built from /root/reference by
    oracle/build_ref.sh) through its own public API, the reference's own DeformNetwork class and loss functions
    (oracle/_ref/trase_ref/utils), on the same workload and the same view-parallel harness.  The elementwise glue
    between them is trase_b200.render with FUSED_GLUE off (gaussian_renderer/__init__.py itself needs the scene
    package).  --ref-mlp tf32 lets torch use TF32 for the reference MLP (the reference's default is fp32).
    Falls back to the CPU oracle port when no GPU or no oracle/_ref is available."""
    from oracle import ref_loader
    if device is None or not ref_loader.available():
        if rank != 0:
            return None
        cb = cpu_baseline_port(args.workload)
        return {"impl": "reference", "metric": METRIC, "value": round(cb["value"], 5), "unit": UNIT, "n_gpus": world,
                "steps": 1, "warmup": 0, "ms_per_step": round(1e3 / cb["value"], 2), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOADS[args.workload]["desc"]}, "cpu_baseline": cb,
                "e2e": {"value": round(cb["value"], 5), "unit": UNIT, "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}, "device": "cpu"}
    rmod = ref_loader.load()
    torch.backends.cuda.matmul.allow_tf32 = args.ref_mlp == "tf32"
    ref = SimpleNamespace(rasterizer_cls=(rmod.GaussianRasterizer, rmod.GaussianRasterizationSettings),
                          utils=ref_loader.load_utils() if ref_loader.utils_available() else None)
    Workload.full_arena = bool(args.full_arena)
    wl = Workload(args.workload, device, rank, mlp_backend="torch", ref=ref)
    if wl.state == "INFERENCE":
        def step():
            with torch.no_grad():
                return wl.render(wl.view, wl.time)
        step_e2e = step
    else:
        step, step_e2e = wl.step_resident, wl.step_e2e
    ms = timed_region(step, args.steps, args.warmup, world, device)
    value = world * args.steps / (ms / 1e3)
    ms_e2e = timed_region(step_e2e, args.steps, max(3, args.warmup), world, device)
    e2e = world * args.steps / (ms_e2e / 1e3)
    if rank != 0:
        return None
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 rasterizer; MLP torch %s" % ("tf32 (allow_tf32=True)" if args.ref_mlp == "tf32" else "fp32"),
            "data": "synthetic",
            "config": {"workload": wl.cfg["desc"], "rasterizer": "unmodified reference CUDA rasterizer compiled for "
                       "sm_100 (oracle/_ref)", "mlp": "the reference's utils/time_utils.py DeformNetwork, torch eager "
                       + args.ref_mlp if ref.utils is not None else "trase_b200.deform torch backend",
                       "losses": "the reference's utils/loss_utils.py / utils/feature_utils.py" if ref.utils is not None
                       else "restated", "views_per_step_per_gpu": wl.nviews, "state": wl.state},
            "device": "cuda",
            "e2e": {"value": round(e2e, 3), "unit": UNIT, "h2d_bytes_per_step": wl.h2d_bytes() if wl.state != "INFERENCE" else 0,
                    "d2h_bytes_per_step": wl.d2h_bytes() if wl.state != "INFERENCE" else 0,
                    "ms_per_step": round(ms_e2e / args.steps, 4)},
            "cpu_baseline": None}
    if world == 1 and wl.state != "INFERENCE":
        line["raster_only"] = reference_raster_only(wl, device)
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="neu3d", choices=list(WORKLOADS))
    ap.add_argument("--mlp", default=os.environ.get("TRASE_B200_MLP", "native"), choices=["torch", "torch-tf32", "native"])
    ap.add_argument("--ref-mlp", default="fp32", choices=["fp32", "tf32"],
                    help="reference arm: precision torch may use for the reference DeformNetwork's matmuls")
    ap.add_argument("--full-arena", action="store_true",
                    help="GAUSSIAN state: also accumulate and all-reduce the (untrained) feature gradient, as round 1 did")
    ap.add_argument("--torch-feature-loss", action="store_true",
                    help="FEATURE state: evaluate the contrastive loss with the torch N x N formulation instead of the "
                         "fused kernels")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reserve-sms", type=int, default=None,
                    help="SMs the persistent MLP kernels leave to the concurrent NCCL all-reduce (default: 0 at N=1, "
                         "RESERVE_SMS_MULTI at N>1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    have_gpu = torch.cuda.is_available()
    device = None
    if have_gpu:
        device = torch.device("cuda", local_rank)
        torch.cuda.set_device(device)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if args.impl == "ours" and WORKLOADS[args.workload]["state"] == "GAUSSIAN":
            # the state whose all-reduce runs under the MLP backward (the others have nothing to hide it under)
            os.environ.setdefault("NCCL_MAX_CTAS", NCCL_CTAS_MULTI)   # see RESERVE_SMS_MULTI
        dist.init_process_group("nccl" if have_gpu else "gloo", device_id=device if have_gpu else None)
    try:
        if args.impl == "reference":
            line = run_reference(args, rank, world, device)
        else:
            if not have_gpu:
                raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback "
                                 "(use --impl reference for the CPU port)")
            from trase_b200 import _lib
            reserve = args.reserve_sms if args.reserve_sms is not None else (
                RESERVE_SMS_MULTI if world > 1 and WORKLOADS[args.workload]["state"] == "GAUSSIAN" else 0)
            if args.workload != "inference2m":        # inference has no collective
                _lib.set_tuning("reserve_sms", reserve)
            line = run_inference(args, rank, world, device) if args.workload == "inference2m" else \
                run_ours(args, rank, world, device)
            if line is not None and args.workload != "inference2m":
                line["config"]["reserve_sms"] = reserve
                if world > 1:
                    line["config"]["nccl_max_ctas"] = os.environ.get("NCCL_MAX_CTAS")
        if rank == 0 and line is not None:
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
