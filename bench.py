#!/usr/bin/env python
"""bench.py — fwd+bwd RGB+feature render steps/s on BASELINE.json's config 2
(Neu3D-shaped: 300k Gaussians, 1352x1014, DeformNetwork MLP + rasterizer).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One "step" = one training view: DeformNetwork forward -> parameter activations
-> rasterizer forward (RGB + 32-ch feature + depth) -> backward through the
rasterizer, the activations and the MLP.  At N > 1 every rank renders its own
view (weak scaling) and the per-Gaussian + MLP gradient arena is all-reduced
with NCCL inside the step.  Prints ONE JSON line on rank 0.
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

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "fwd+bwd RGB+feature render steps/sec"
UNIT = "steps/s"
WORKLOADS = {  # BASELINE.json configs
    "neu3d": dict(P=300_000, W=1352, H=1014, seed=2, desc="configs[1]: Neu3D-shaped 300k Gaussians 1352x1014, "
                  "DeformNetwork MLP + RGB+feature fwd+bwd"),
    "tiny": dict(P=20_000, W=128, H=128, seed=1, desc="configs[0]: tiny 20k Gaussians 128x128"),
    "inference2m": dict(P=2_000_000, W=1920, H=1080, seed=4, desc="configs[3]: 2M Gaussians 1920x1080 feature-field "
                        "inference (DeformNetwork + RGB+feature+depth forward, no backward)"),
}


# --------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.rows.append([x.strip() for x in out.stdout.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------
class Workload:
    """Synthetic scene, raw parameters, DeformNetwork, camera."""

    def __init__(self, name, device, rank=0, mlp_backend="torch"):
        from trase_b200 import synth
        from trase_b200.deform import DeformNetwork
        from trase_b200.render import GaussianParams, make_view
        cfg = WORKLOADS[name]
        self.name, self.cfg, self.device = name, cfg, device
        P, W, H = cfg["P"], cfg["W"], cfg["H"]
        self.scene = synth.make_scene(P, W, H, cfg["seed"])
        self.pc = GaussianParams.from_activated(self.scene, device)
        torch.manual_seed(1234)
        self.deform = DeformNetwork(backend=mlp_backend).to(device)
        with torch.no_grad():  # keep the deformation small, as in a trained model (random heads would blow R up)
            for head in (self.deform.gaussian_warp, self.deform.gaussian_rotation, self.deform.gaussian_scaling):
                head.weight.mul_(1e-2)
                head.bias.mul_(1e-2)
        # each rank gets its own view: a small yaw about the scene centre + its own time stamp
        ang = 0.02 * rank
        c, s = math.cos(ang), math.sin(ang)
        wv = torch.eye(4)
        if rank:
            Rm = torch.tensor([[c, 0, s], [0, 1, 0], [-s, 0, c]])
            centre = torch.tensor([0.0, 0.0, 4.75])
            wv[:3, :3] = Rm.T                      # row-vector convention (transposed world->view)
            wv[3, :3] = centre - centre @ Rm.T
        self.cam = synth.make_camera(W, H, wv)
        self.view = make_view(self.cam, device)
        # training: a materialised [P,1] tensor (train.py:202 adds per-iteration noise to the expanded time stamp);
        # inference: the reference's own broadcast view, fid.unsqueeze(0).expand(N, -1) (render.py / train.py:450)
        self.time = torch.full((P, 1), 0.37 + 0.1 * rank, device=device)
        if name.startswith("inference"):
            self.time = torch.full((1, 1), 0.37 + 0.1 * rank, device=device).expand(P, 1)
        up = synth.make_upstream_grads(W, H, cfg["seed"] + rank)
        self.dL_dcolor, self.dL_dobjects = up["dL_dcolor"].to(device), up["dL_dobjects"].to(device)
        self.bg = torch.zeros(3, device=device)
        self.params = list(self.pc.parameters()) + list(self.deform.parameters())
        # one flat gradient arena; every .grad is a view into it (single memset, single all-reduce)
        from trase_b200.distributed import GradArena
        # the Gaussians' own gradients are final before the MLP backward starts: early segment, reduced under it
        self.grad_arena = GradArena(self.params, early=list(self.pc.parameters()))
        self.arena = self.grad_arena.flat
        # e2e leg: ground-truth image + camera block in pinned host memory
        g = torch.Generator().manual_seed(99 + rank)
        self.gt_host = torch.rand(3, H, W, generator=g).pin_memory()
        cam_block = torch.cat([self.cam["viewmatrix"].reshape(-1), self.cam["projmatrix"].reshape(-1),
                               self.cam["campos"].reshape(-1), torch.tensor([0.37 + 0.1 * rank])])
        self.cam_host = cam_block.pin_memory()
        self.gt_dev = torch.empty(3, H, W, device=device)
        self.cam_dev = torch.empty_like(cam_block, device=device)
        self.loss_host = torch.zeros(1).pin_memory()
        self.copy_stream = torch.cuda.Stream(device=device)
        self.loss_event = torch.cuda.Event()
        self.last_R = None

    def render(self, view, time_t, rasterizer_cls=None):
        from trase_b200 import render as R
        d_xyz, d_rot, d_scale = self.deform(self.pc.get_xyz.detach(), time_t)
        if rasterizer_cls is None:
            return R.render(view, self.pc, R.PIPE_DEFAULT, self.bg, d_xyz, d_rot, d_scale)
        # reference arm: identical glue, the reference's own GaussianRasterizer
        saved = R.GaussianRasterizer, R.GaussianRasterizationSettings, R.FUSED_GLUE
        R.GaussianRasterizer, R.GaussianRasterizationSettings = rasterizer_cls
        R.FUSED_GLUE = False          # the reference's own elementwise torch glue: none of our kernels on this arm
        try:
            return R.render(view, self.pc, R.PIPE_DEFAULT, self.bg, d_xyz, d_rot, d_scale)
        finally:
            R.GaussianRasterizer, R.GaussianRasterizationSettings, R.FUSED_GLUE = saved

    def step_resident(self, world_size=1, rasterizer_cls=None):
        """Inputs already in HBM: fwd + bwd with fixed upstream gradients."""
        self.arena.zero_()
        out = self.render(self.view, self.time, rasterizer_cls)
        self.grad_arena.arm()
        torch.autograd.backward([out["render"], out["render_gaussian_features"]],
                                [self.dL_dcolor, self.dL_dobjects])
        self.grad_arena.all_reduce()
        return out

    def step_e2e(self, world_size=1, rasterizer_cls=None):
        """Through the public API with HOST buffers: H2D of this view's camera
        block and ground-truth image, the reference's L1 + SSIM colour loss and
        feature-norm regulariser (train.py:235-238, 280-281), backward, D2H of the loss."""
        from types import SimpleNamespace
        # the camera block is needed at once; the ground-truth image is only read by the loss, so its
        # copy runs on a side stream under this step's forward pass and the loss waits for it
        main = torch.cuda.current_stream()
        self.cam_dev.copy_(self.cam_host, non_blocking=True)
        self.copy_stream.wait_stream(main)      # the previous step's loss has finished reading gt_dev
        with torch.cuda.stream(self.copy_stream):
            self.gt_dev.copy_(self.gt_host, non_blocking=True)
        v = self.view
        view = SimpleNamespace(FoVx=v.FoVx, FoVy=v.FoVy, image_height=v.image_height, image_width=v.image_width,
                               world_view_transform=self.cam_dev[0:16].view(4, 4),
                               full_proj_transform=self.cam_dev[16:32].view(4, 4), camera_center=self.cam_dev[32:35])
        t = self.cam_dev[35:36].expand(self.cfg["P"], 1)
        self.arena.zero_()
        out = self.render(view, t, rasterizer_cls)
        main.wait_stream(self.copy_stream)
        # the reference's two losses on the path's outputs: GAUSSIAN state (train.py:235-238)
        # (1 - l) * L1 + l * (1 - SSIM), l = 0.2, and the FEATURE state's norm regulariser (train.py:280-281)
        if rasterizer_cls is None:
            from trase_b200 import loss_utils          # fused kernels (csrc/loss.cu)
            loss = loss_utils.l1_ssim_loss(out["render"], self.gt_dev, 0.2) + \
                loss_utils.feature_norm_reg(out["render_gaussian_features"])
        else:
            from oracle import loss_oracle             # reference arm: the reference's torch formulation
            loss = loss_oracle.l1_ssim_ref(out["render"], self.gt_dev, 0.2) + \
                (1.0 - out["render_gaussian_features"].norm(dim=0, p=2).mean()) ** 2
        # the step's result goes back to the host as soon as it exists (after the loss forward); the host waits for
        # THAT copy, not for the backward pass, so the next step's launches queue up behind this step's backward
        # exactly as in a training loop that logs its loss (the timed region still ends with a full synchronize)
        self.loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        self.loss_event.record()
        self.grad_arena.arm()
        loss.backward()
        self.grad_arena.all_reduce()
        self.loss_event.synchronize()
        return float(self.loss_host[0])

    def h2d_bytes(self):
        return self.gt_host.numel() * 4 + self.cam_host.numel() * 4

    def d2h_bytes(self):
        return 4 + 4  # loss + the rasterizer's instance count


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
RESERVE_SMS_MULTI = 0


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


def cpu_baseline_port(name, budget_s=30.0):
    """The CPU oracle (oracle/, a C port of the reference's algorithm) timed on
    this box's host cores: one full fwd+bwd rasterizer step of the same workload
    plus the MLP fwd+bwd in torch CPU fp32."""
    import numpy as np
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
    oracle.backward(f, up["dL_dcolor"].numpy(), up["dL_dobjects"].numpy())
    t_raster = time.perf_counter() - t0
    # MLP on a bounded row sample, scaled to P rows
    rows = min(P, 60_000)
    net = DeformNetwork()
    x = sc["means3D"][:rows]
    t = torch.full((rows, 1), 0.37)
    t0 = time.perf_counter()
    outs = net(x, t)
    torch.autograd.backward(list(outs), [torch.ones_like(o) for o in outs])
    t_mlp = (time.perf_counter() - t0) * (P / rows)
    total = t_raster + t_mlp
    return {"value": 1.0 / total, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "1 full rasterizer fwd+bwd step of %s with the C oracle (OpenMP, %.2f s) + DeformNetwork "
                      "fwd+bwd in torch CPU fp32 on %d of %d rows scaled to all rows (%.2f s)" %
                      (name, t_raster, rows, P, t_mlp)}


# --------------------------------------------------------------------------
def run_ours(args, rank, world, device):
    from trase_b200 import _lib, build as tb_build
    if not os.path.exists(_lib.LIB_PATH):
        tb_build.build()
    _lib.load()  # fails loudly when the CUDA library is missing: there is no fallback
    if args.mlp == "torch-tf32":
        torch.backends.cuda.matmul.allow_tf32 = True
    wl = Workload(args.workload, device, rank, mlp_backend="native" if args.mlp == "native" else "torch")
    cfg = wl.cfg
    step = lambda: wl.step_resident(world)
    sampler = ClockSampler(device.index) if rank == 0 else None
    # launches of OUR kernels in the timed region (counted by the library itself); warm-up excluded
    for _ in range(args.warmup):
        step()
    l0 = _lib.launch_count()
    ms = timed_region(step, args.steps, 0, world, device, sampler)
    launches = _lib.launch_count() - l0
    value = world * args.steps / (ms / 1e3)
    clocks = sampler.summary() if sampler else None

    # end to end: host buffers -> public API -> loss on the host
    ms_e2e = timed_region(lambda: wl.step_e2e(world), args.steps, max(3, args.warmup), world, device)
    e2e = world * args.steps / (ms_e2e / 1e3)

    line = None
    # per-stage device times inside the step (second pass with event pairs around each stage); every rank
    # runs it because the step contains the collective
    _lib.profile_enable(True)
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize(device)
    prof = _lib.profile_read()
    _lib.profile_enable(False)
    if rank == 0:
        from trase_b200 import _raster
        R = _raster._capacity_hint.get((device.index or 0, cfg["P"], cfg["W"], cfg["H"]), 0)
        P, N = cfg["P"], cfg["W"] * cfg["H"]
        T = ((cfg["W"] + 15) // 16) * ((cfg["H"] + 15) // 16)
        # SURVEY §8(d).2 counts the reference's sort: 8-bit passes over all 32 + tile bits, whatever we launch
        n_pass = (32 + max(1, T.bit_length()) + 7) // 8
        stages = {}
        for k, (tot, cnt) in prof.items():
            if cnt:
                per_step = tot / args.steps
                ab = algorithmic_bytes(k, P, R, N, 16, n_pass)
                stages[k] = {"ms": round(per_step, 4), "GBps": round(ab / per_step / 1e6, 1)}
        dom = max(stages, key=lambda k: stages[k]["ms"]) if stages else None
        peak, peak_src = measured_peaks()
        roof = None
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(dom, {}).get("bytes")
        except Exception:
            pass
        if dom:
            ach = stages[dom]["GBps"]
            roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": round(ach / peak, 4), "traffic": traffic, "traffic_unit": "bytes per launch (ncu --set full, profiles/traffic.json)",
                    "peak_source": peak_src,
                    "algorithmic_bytes": algorithmic_bytes(dom, P, R, N, 16, n_pass),
                    "launch_ms": stages[dom]["ms"],
                    "note": "compositing kernels are FP32/shuffle-bound on (pixel,Gaussian) pairs, not HBM-bound "
                            "(SURVEY.md §8d); HBM fraction is low by construction"}
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic (SURVEY.md §8(d).1 generator, "
            "seed 2; random-init DeformNetwork with output heads scaled 1e-2)",
            "config": {"workload": cfg["desc"], "P": P, "W": cfg["W"], "H": cfg["H"], "num_rendered": R,
                       "views_per_step_per_gpu": 1, "mlp": args.mlp, "parallelism": "view-parallel dp%d, NCCL "
                       "all-reduce of the %d-float gradient arena" % (world, wl.arena.numel()) if world > 1 else
                       "single GPU", "l2": "no flush: per-step working set (params 109 MB + 36 image planes x2 "
                       "+ workspaces) exceeds the 126 MB L2"},
            "e2e": {"value": round(e2e, 3), "unit": UNIT, "h2d_bytes_per_step": wl.h2d_bytes(),
                    "d2h_bytes_per_step": wl.d2h_bytes(), "ms_per_step": round(ms_e2e / args.steps, 4)},
            "gpu_launches": launches,
            "clocks": clocks, "roofline": roof, "stages": stages,
        }
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
            "dtype": "f32", "data": "synthetic", "config": {"workload": wl.cfg["desc"], "mlp": args.mlp}}


def run_reference(args, rank, world, device):
    """Reference arm: the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built
    from /root/reference by oracle/build_ref.sh) through its own public API, same
    workload, same glue and MLP (the reference's MLP is eager torch fp32, which
    is exactly trase_b200.deform's "torch" backend).  Falls back to the CPU
    oracle port when no GPU or no oracle/_ref is available."""
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
    ref = ref_loader.load()
    cls = (ref.GaussianRasterizer, ref.GaussianRasterizationSettings)
    wl = Workload(args.workload, device, rank, mlp_backend="torch")
    ms = timed_region(lambda: wl.step_resident(world, cls), args.steps, args.warmup, world, device)
    value = world * args.steps / (ms / 1e3)
    ms_e2e = timed_region(lambda: wl.step_e2e(world, cls), args.steps, max(3, args.warmup), world, device)
    e2e = world * args.steps / (ms_e2e / 1e3)
    if rank != 0:
        return None
    return {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl.cfg["desc"], "rasterizer": "unmodified reference CUDA rasterizer compiled for "
                       "sm_100 (oracle/_ref), mlp torch fp32"}, "device": "cuda",
            "e2e": {"value": round(e2e, 3), "unit": UNIT, "h2d_bytes_per_step": wl.h2d_bytes(),
                    "d2h_bytes_per_step": wl.d2h_bytes(), "ms_per_step": round(ms_e2e / args.steps, 4)},
            "cpu_baseline": None}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="neu3d", choices=list(WORKLOADS))
    ap.add_argument("--mlp", default=os.environ.get("TRASE_B200_MLP", "native"), choices=["torch", "torch-tf32", "native"])
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
        dist.init_process_group("nccl" if have_gpu else "gloo", device_id=device if have_gpu else None)
    try:
        if args.impl == "reference":
            line = run_reference(args, rank, world, device)
        else:
            if not have_gpu:
                raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback "
                                 "(use --impl reference for the CPU port)")
            from trase_b200 import _lib
            reserve = args.reserve_sms if args.reserve_sms is not None else (RESERVE_SMS_MULTI if world > 1 else 0)
            if args.workload != "inference2m":        # inference has no collective
                _lib.set_tuning("reserve_sms", reserve)
            line = run_inference(args, rank, world, device) if args.workload == "inference2m" else \
                run_ours(args, rank, world, device)
            if line is not None and args.workload != "inference2m":
                line["config"]["reserve_sms"] = reserve
        if rank == 0 and line is not None:
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
