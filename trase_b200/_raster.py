"""Host-side driver of the native rasterizer: buffer allocation (torch), the
ctypes calls, and the capacity/retry policy that replaces the reference's
blocking `cudaMemcpy` of the instance count (DGR/cuda_rasterizer/rasterizer_impl.cu:288).

Mirrors RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA
(DGR/rasterize_points.cu:40-214): same inputs, same outputs, torch tensors own
every buffer; the native side only sees raw pointers.
"""
from __future__ import annotations

import ctypes as C
import math
import threading
from typing import Dict, Optional, Tuple

import torch

from . import _cuda, _lib

NUM_CHANNELS = 3
NUM_OBJECTS = 32

# last instance count seen per (device, P, W, H): sizes the next call's binning workspace
_capacity_hint: Dict[Tuple[int, int, int, int], int] = {}
_pinned: Dict[int, torch.Tensor] = {}
# the tables below are shared by every host thread that renders; each entry is keyed per (device, stream), so two
# threads on two streams never share an entry — the lock only protects the dictionaries themselves
_state_lock = threading.Lock()


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    """Device pointer or NULL for 'not provided' (empty tensor convention,
    DGR/diff_gaussian_rasterization/__init__.py:206-219)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t: torch.Tensor, device: torch.device) -> torch.Tensor:
    if t.numel() == 0:
        return t
    if t.device != device:
        raise RuntimeError("trase_b200: all inputs must live on the same CUDA device (%s vs %s)" % (t.device, device))
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    if t.data_ptr() % 16 != 0:  # 16-byte vector loads need aligned bases
        t = t.clone()
    return t


def _pinned_int(device: torch.device, stream=None) -> torch.Tensor:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx, stream.cuda_stream if stream is not None else 0)     # per stream, like the readback event
    with _state_lock:
        t = _pinned.get(key)
        if t is None:
            t = torch.zeros(2, dtype=torch.int32).pin_memory()
            _pinned[key] = t
    return t


def _initial_capacity(P: int, W: int, H: int) -> int:
    return max(4 * P, 1 << 16)


class ForwardResult:
    __slots__ = ("num_rendered", "color", "objects", "depth", "radii", "geom", "binning", "img", "capacity")


def rasterize_forward(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3D_precomp,
                      viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, sh, sh_objs, degree,
                      campos, prefiltered, debug) -> ForwardResult:
    """Same argument order as `_C.rasterize_gaussians` (DGR/ext.cpp:16)."""
    lib = _lib.load()
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:63-66
    if not means3D.is_cuda:
        raise RuntimeError("trase_b200: the rasterizer runs on CUDA tensors only (no CPU fallback)")
    dev = means3D.device
    P, H, W = int(means3D.shape[0]), int(image_height), int(image_width)

    with _cuda.on_device(dev):
        means3D = _f32c(means3D, dev)
        bg = _f32c(bg, dev); viewmatrix = _f32c(viewmatrix, dev); projmatrix = _f32c(projmatrix, dev)
        campos = _f32c(campos, dev); opacities = _f32c(opacities, dev)
        colors_precomp = _f32c(colors_precomp, dev) if colors_precomp.numel() else colors_precomp
        scales = _f32c(scales, dev) if scales.numel() else scales
        rotations = _f32c(rotations, dev) if rotations.numel() else rotations
        cov3D_precomp = _f32c(cov3D_precomp, dev) if cov3D_precomp.numel() else cov3D_precomp
        sh = _f32c(sh, dev) if sh.numel() else sh
        sh_objs = _f32c(sh_objs, dev) if sh_objs.numel() else sh_objs
        M = int(sh.shape[1]) if sh.numel() != 0 else 0

        f32 = dict(dtype=torch.float32, device=dev)
        res = ForwardResult()
        res.color = torch.empty((NUM_CHANNELS, H, W), **f32)
        res.objects = torch.empty((NUM_OBJECTS, H, W), **f32)
        res.depth = torch.empty((1, H, W), **f32)
        res.radii = torch.empty((P,), dtype=torch.int32, device=dev)
        u8 = dict(dtype=torch.uint8, device=dev)
        res.geom = torch.empty((lib.trase_b200_geom_bytes(P),), **u8)
        res.img = torch.empty((lib.trase_b200_img_bytes(W, H),), **u8)

        key = (dev.index or 0, P, W, H)
        with _state_lock:
            hint = _capacity_hint.get(key)
        capacity = _initial_capacity(P, W, H) if hint is None else int(hint * 1.25) + 4096
        capacity = max(capacity, P)      # the depth sort of the P Gaussians borrows the instance buffers
        if P == 0:
            capacity = 0
        stream = torch.cuda.current_stream(dev)
        pinned = _pinned_int(dev, stream)

        a = _lib.FwdArgs()
        a.P, a.sh_degree, a.M, a.W, a.H = P, int(degree), M, W, H
        a.tan_fovx, a.tan_fovy, a.scale_modifier = float(tanfovx), float(tanfovy), float(scale_modifier)
        a.prefiltered, a.debug = int(bool(prefiltered)), int(bool(debug))
        a.background, a.means3D, a.shs, a.sh_objs = _ptr(bg), _ptr(means3D), _ptr(sh), _ptr(sh_objs)
        a.colors_precomp, a.opacities = _ptr(colors_precomp), _ptr(opacities)
        a.scales, a.rotations, a.cov3D_precomp = _ptr(scales), _ptr(rotations), _ptr(cov3D_precomp)
        a.viewmatrix, a.projmatrix, a.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
        a.out_color, a.out_objects, a.out_depth = res.color.data_ptr(), res.objects.data_ptr(), res.depth.data_ptr()
        a.radii = _ptr(res.radii)
        a.geom, a.geom_bytes = res.geom.data_ptr(), res.geom.numel()
        a.img, a.img_bytes = res.img.data_ptr(), res.img.numel()
        a.num_rendered_host = pinned.data_ptr()
        a.stream = stream.cuda_stream
        readback = _readback_event(dev, stream)
        a.readback_event = readback.cuda_event

        first = True
        while True:
            res.binning = torch.empty((lib.trase_b200_binning_bytes(capacity),), **u8)
            scratch = torch.empty((lib.trase_b200_scratch_bytes(P, capacity),), **u8)
            a.binning, a.binning_bytes = _ptr(res.binning), res.binning.numel()
            a.scratch, a.scratch_bytes = _ptr(scratch), scratch.numel()
            a.capacity = capacity
            fn = lib.trase_b200_rasterize_forward if first else lib.trase_b200_rasterize_forward_rebin
            _lib.check(fn(C.byref(a)))
            # One host wait per pass, on the event recorded right after the 4-byte readback of R was
            # queued (the GPU pipeline itself never stalls): R must reach Python as an int for ctx
            # (…/__init__.py:103).  The sort and the compositing kernel are still running when this
            # returns, so the caller's Python (loss, autograd start-up, backward launches) overlaps them.
            readback.synchronize()
            R = int(pinned[0].item()) if P > 0 else 0
            if R <= capacity:
                break
            capacity = max(int(R * 1.25) + 4096, P)  # workspace was too small: re-bin only
            first = False
        del scratch
        with _state_lock:
            _capacity_hint[key] = R
        res.num_rendered = R
        res.capacity = capacity
        return res


_events: Dict = {}


def _readback_event(dev, stream):
    """One reusable CUDA event per (device, stream), created by a first record, re-recorded by the library."""
    key = (dev, stream.cuda_stream)          # one event per stream: forwards on different streams may overlap
    with _state_lock:
        ev = _events.get(key)
        if ev is None:
            ev = torch.cuda.Event()
            ev.record(stream)
            _events[key] = ev
    return ev


def rasterize_backward(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier, cov3D_precomp,
                       viewmatrix, projmatrix, tanfovx, tanfovy, dL_dout_color, sh, dL_dout_objects, sh_objs,
                       degree, campos, geom, R, binning, img, debug, capacity=None, need_geom=True,
                       need_features=True):
    """Same argument order as `_C.rasterize_gaussians_backward` (DGR/ext.cpp:17);
    returns the same 9 gradients (DGR/rasterize_points.cu:213).

    need_geom / need_features say which of them anyone will read (autograd's needs_input_grad): with
    need_features=False the feature gradient slot is None, with need_geom=False the eight others are None and only
    the feature gradient is computed (no dL/dalpha chain, no preprocess backward).  dL_dout_objects (or, for the
    feature-only case, dL_dout_color) may be None = an all-zero upstream gradient."""
    lib = _lib.load()
    dev = means3D.device
    P = int(means3D.shape[0])
    if not need_geom and not need_features:
        return (None,) * 9
    some = dL_dout_color if dL_dout_color is not None else dL_dout_objects
    H, W = int(some.shape[1]), int(some.shape[2])
    if need_geom and dL_dout_color is None:
        dL_dout_color = torch.zeros((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
    if not need_geom and dL_dout_objects is None:          # feature gradient of a loss that never read the features
        return (None, None, torch.zeros((P, NUM_OBJECTS), dtype=torch.float32, device=dev)) + (None,) * 6
    with _cuda.on_device(dev):
        means3D = _f32c(means3D, dev)
        bg = _f32c(bg, dev); viewmatrix = _f32c(viewmatrix, dev); projmatrix = _f32c(projmatrix, dev)
        campos = _f32c(campos, dev)
        colors_precomp = _f32c(colors_precomp, dev) if colors_precomp.numel() else colors_precomp
        scales = _f32c(scales, dev) if scales.numel() else scales
        rotations = _f32c(rotations, dev) if rotations.numel() else rotations
        cov3D_precomp = _f32c(cov3D_precomp, dev) if cov3D_precomp.numel() else cov3D_precomp
        sh = _f32c(sh, dev) if sh.numel() else sh
        sh_objs = _f32c(sh_objs, dev) if sh_objs.numel() else sh_objs
        dL_dout_color = _f32c(dL_dout_color, dev) if dL_dout_color is not None else None
        dL_dout_objects = _f32c(dL_dout_objects, dev) if dL_dout_objects is not None else None
        M = int(sh.shape[1]) if sh.numel() != 0 else 0

        # all gradient buffers carved out of ONE zero-filled allocation (one memset instead of ten);
        # every sub-buffer starts on a 256-byte boundary (the compositing kernel adds float4s)
        shapes = [(P, 3), (P, 3), (P, NUM_CHANNELS), (P, NUM_OBJECTS), (P, 2, 2), (P, 1), (P, 6), (P, M, 3), (P, 3),
                  (P, 4)]
        wanted = [need_geom, need_geom, need_geom, need_features] + [need_geom] * 6
        shapes = [shp if w else (0,) for shp, w in zip(shapes, wanted)]     # unwanted gradients take no memory
        offs, n = [], 0
        for shp in shapes:
            offs.append(n)
            cnt = 1
            for d in shp:
                cnt *= d
            n += (cnt + 63) // 64 * 64
        flat = torch.zeros((n,), dtype=torch.float32, device=dev)
        (dL_dmeans3D, dL_dmeans2D, dL_dcolors, dL_dobjects, dL_dconic, dL_dopacity, dL_dcov3D, dL_dsh, dL_dscales,
         dL_drotations) = [flat[o:o + math.prod(shp)].view(shp) for o, shp in zip(offs, shapes)]
        if P != 0:
            if capacity is None:
                capacity = (binning.numel() // 12) if binning.numel() else 0
            a = _lib.BwdArgs()
            a.P, a.sh_degree, a.M, a.W, a.H = P, int(degree), M, W, H
            a.tan_fovx, a.tan_fovy, a.scale_modifier = float(tanfovx), float(tanfovy), float(scale_modifier)
            a.debug = int(bool(debug))
            a.background, a.means3D, a.shs, a.sh_objs = _ptr(bg), _ptr(means3D), _ptr(sh), _ptr(sh_objs)
            a.colors_precomp, a.scales, a.rotations = _ptr(colors_precomp), _ptr(scales), _ptr(rotations)
            a.cov3D_precomp = _ptr(cov3D_precomp)
            a.viewmatrix, a.projmatrix, a.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
            a.radii = _ptr(radii)
            a.geom, a.img, a.binning, a.capacity = _ptr(geom), _ptr(img), _ptr(binning), int(capacity)
            a.dL_dout_color = _ptr(dL_dout_color) if need_geom else None
            a.dL_dout_objects = _ptr(dL_dout_objects)
            a.dL_dmeans2D, a.dL_dconic, a.dL_dopacity = _ptr(dL_dmeans2D), _ptr(dL_dconic), _ptr(dL_dopacity)
            a.dL_dcolors, a.dL_dobjects, a.dL_dmeans3D = _ptr(dL_dcolors), _ptr(dL_dobjects), _ptr(dL_dmeans3D)
            a.dL_dcov3D, a.dL_dsh = _ptr(dL_dcov3D), _ptr(dL_dsh)
            a.dL_dscales, a.dL_drotations = _ptr(dL_dscales), _ptr(dL_drotations)
            a.stream = _cuda.raw_stream(dev)
            _lib.check(lib.trase_b200_rasterize_backward(C.byref(a)))
        out = (dL_dmeans2D, dL_dcolors, dL_dobjects, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
               dL_drotations)
        keep = [need_geom, need_geom, need_features] + [need_geom] * 6
        return tuple(g if k else None for g, k in zip(out, keep))


def mark_visible(means3D, viewmatrix, projmatrix) -> torch.Tensor:
    """`_C.mark_visible` (DGR/ext.cpp:18)."""
    lib = _lib.load()
    if not means3D.is_cuda:
        raise RuntimeError("trase_b200: mark_visible runs on CUDA tensors only")
    dev = means3D.device
    P = int(means3D.shape[0])
    with _cuda.on_device(dev):
        means3D = _f32c(means3D, dev)
        viewmatrix = _f32c(viewmatrix, dev); projmatrix = _f32c(projmatrix, dev)
        present = torch.zeros((P,), dtype=torch.bool, device=dev)
        if P:
            _lib.check(lib.trase_b200_mark_visible(P, means3D.data_ptr(), viewmatrix.data_ptr(),
                                                  projmatrix.data_ptr(), present.data_ptr(),
                                                  _cuda.raw_stream(dev)))
        return present


def workspace_views(res_geom, res_img, res_binning, P, W, H, capacity, R) -> dict:
    """Typed views into the three workspaces (what a test on the reference would
    read out of geomBuffer/binningBuffer/imgBuffer with the fromChunk layouts)."""
    out = {}
    go, io, bo = _lib.geom_offsets(P), _lib.img_offsets(W, H), _lib.binning_offsets(capacity)

    def view(buf, off, count, dtype):
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        return buf[off:off + nbytes].view(dtype)

    out["depths"] = view(res_geom, go["depths"], P, torch.float32)
    out["means2D"] = view(res_geom, go["means2D"], 2 * P, torch.float32).view(P, 2)
    out["conic_opacity"] = view(res_geom, go["conic_opacity"], 4 * P, torch.float32).view(P, 4)
    out["rgbd"] = view(res_geom, go["rgbd"], 4 * P, torch.float32).view(P, 4)
    out["cov3D"] = view(res_geom, go["cov3D"], 6 * P, torch.float32).view(P, 6)
    out["clamped"] = view(res_geom, go["clamped"], P, torch.uint8)
    out["tiles_touched"] = view(res_geom, go["tiles_touched"], P, torch.int32)
    out["point_offsets"] = view(res_geom, go["point_offsets"], P, torch.int32)
    out["counters"] = view(res_geom, go["counters"], 8, torch.int32)
    N = W * H
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out["final_T"] = view(res_img, io["final_T"], N, torch.float32)
    out["n_contrib"] = view(res_img, io["n_contrib"], N, torch.int32)
    out["ranges"] = view(res_img, io["ranges"], 2 * T, torch.int32).view(T, 2)
    if capacity > 0:
        out["point_list"] = view(res_binning, bo["point_list"], R, torch.int32)
        out["point_list_keys"] = view(res_binning, bo["point_list_keys"], R, torch.int64)
    else:
        out["point_list"] = torch.empty(0, dtype=torch.int32, device=res_geom.device)
        out["point_list_keys"] = torch.empty(0, dtype=torch.int64, device=res_geom.device)
    return out
