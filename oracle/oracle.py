"""TEST INFRASTRUCTURE — not part of the product (see oracle/trase_oracle.c).

numpy/ctypes front end of the C restatement.  Only tests/, bench.py's
cpu_baseline/reference legs and __graft_entry__.smoke() may import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None
NOBJ = 32


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "trase_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(src) > os.path.getmtime(_SO):
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared", "-o", _SO, src,
                               "-lm"])
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.oracle_bin.restype = C.c_int64
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def forward(means3D, opacities, sh_objs, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, bg, shs=None,
            sh_degree=3, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0):
    """Whole forward pass; returns a dict with every intermediate the parity
    tests compare (names follow the reference's GeometryState/BinningState/ImageState)."""
    L = lib()
    means3D = _f32(means3D).reshape(-1, 3)
    P = means3D.shape[0]
    opacities = _f32(opacities).reshape(-1)
    sh_objs = _f32(sh_objs).reshape(P, NOBJ)
    view, proj, cam, bg = _f32(viewmatrix).reshape(16), _f32(projmatrix).reshape(16), _f32(campos), _f32(bg)
    shs = _f32(shs)
    M = 0 if shs is None else (shs.shape[1] if shs.ndim == 3 else shs.size // max(3 * P, 1))
    colors_precomp, scales, rotations, cov3D_precomp = map(_f32, (colors_precomp, scales, rotations, cov3D_precomp))
    o = dict(P=P, W=W, H=H)
    o["radii"] = np.zeros(P, np.int32)
    o["means2D"] = np.zeros((P, 2), np.float32)
    o["depths"] = np.zeros(P, np.float32)
    o["cov3D"] = np.zeros((P, 6), np.float32)
    o["rgb"] = np.zeros((P, 3), np.float32)
    o["conic_opacity"] = np.zeros((P, 4), np.float32)
    o["tiles_touched"] = np.zeros(P, np.uint32)
    o["clamped"] = np.zeros((P, 3), np.uint8)
    L.oracle_preprocess_forward(P, int(sh_degree), M, _p(means3D), _p(scales), C.c_float(scale_modifier),
                                _p(rotations), _p(opacities), _p(shs), _p(cov3D_precomp), _p(colors_precomp),
                                _p(view), _p(proj), _p(cam), W, H, C.c_float(tanfovx), C.c_float(tanfovy),
                                _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]), _p(o["cov3D"]), _p(o["rgb"]),
                                _p(o["conic_opacity"]), _p(o["tiles_touched"]), _p(o["clamped"]))
    T = ((W + 15) // 16) * ((H + 15) // 16)
    R = int(o["tiles_touched"].sum(dtype=np.int64))
    o["num_rendered"] = R
    o["point_list_keys"] = np.zeros(max(R, 1), np.uint64)
    o["point_list"] = np.zeros(max(R, 1), np.uint32)
    o["ranges"] = np.zeros((T, 2), np.uint32)
    L.oracle_bin(P, _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]), _p(o["tiles_touched"]), W, H,
                 _p(o["point_list_keys"]), _p(o["point_list"]), _p(o["ranges"]))
    o["point_list_keys"] = o["point_list_keys"][:R]
    o["point_list"] = o["point_list"][:R]
    colors = colors_precomp if colors_precomp is not None else o["rgb"]
    o["colors"] = colors
    o["final_T"] = np.zeros(W * H, np.float32)
    o["n_contrib"] = np.zeros(W * H, np.uint32)
    o["color"] = np.zeros((3, H, W), np.float32)
    o["objects"] = np.zeros((NOBJ, H, W), np.float32)
    o["depth"] = np.zeros((1, H, W), np.float32)
    if P > 0:
        L.oracle_render_forward(W, H, _p(o["ranges"]), _p(o["point_list"]), _p(o["means2D"]), _p(colors),
                                _p(sh_objs), _p(o["depths"]), _p(o["conic_opacity"]), _p(bg), _p(o["final_T"]),
                                _p(o["n_contrib"]), _p(o["color"]), _p(o["objects"]), _p(o["depth"]))
    o["_in"] = dict(means3D=means3D, sh_objs=sh_objs, view=view, proj=proj, cam=cam, bg=bg, shs=shs, M=M,
                    sh_degree=int(sh_degree), colors_precomp=colors_precomp, scales=scales, rotations=rotations,
                    cov3D_precomp=cov3D_precomp, scale_modifier=float(scale_modifier), tanfovx=float(tanfovx),
                    tanfovy=float(tanfovy))
    return o


def backward(fwd: dict, dL_dcolor, dL_dobjects):
    """Gradients of the nine reference outputs (float64), from a forward() result."""
    L = lib()
    i = fwd["_in"]
    P, W, H = fwd["P"], fwd["W"], fwd["H"]
    dL_dcolor = _f32(dL_dcolor).reshape(3, H, W)
    dL_dobjects = _f32(dL_dobjects).reshape(NOBJ, H, W)
    g = dict(means2D=np.zeros((P, 3)), conic=np.zeros((P, 4)), opacities=np.zeros((P, 1)),
             colors_precomp=np.zeros((P, 3)), objects=np.zeros((P, NOBJ)), means3D=np.zeros((P, 3)),
             cov3D=np.zeros((P, 6)), sh=np.zeros((P, i["M"], 3)), scales=np.zeros((P, 3)),
             rotations=np.zeros((P, 4)))
    if P == 0:
        return g
    L.oracle_render_backward(W, H, _p(fwd["ranges"]), _p(fwd["point_list"]), _p(i["bg"]), _p(fwd["means2D"]),
                             _p(fwd["conic_opacity"]), _p(fwd["colors"]), _p(i["sh_objs"]), _p(fwd["final_T"]),
                             _p(fwd["n_contrib"]), _p(dL_dcolor), _p(dL_dobjects), _p(g["means2D"]), _p(g["conic"]),
                             _p(g["opacities"]), _p(g["colors_precomp"]), _p(g["objects"]))
    fy, fx = H / (2.0 * i["tanfovy"]), W / (2.0 * i["tanfovx"])
    cov3Ds = i["cov3D_precomp"] if i["cov3D_precomp"] is not None else fwd["cov3D"]
    use_sh = i["colors_precomp"] is None
    L.oracle_preprocess_backward(P, i["sh_degree"], i["M"], _p(i["means3D"]), _p(fwd["radii"]),
                                 _p(i["shs"]) if use_sh else None, _p(fwd["clamped"]),
                                 _p(i["scales"]) if i["cov3D_precomp"] is None else None,
                                 _p(i["rotations"]) if i["cov3D_precomp"] is None else None,
                                 C.c_float(i["scale_modifier"]), _p(cov3Ds), _p(i["view"]), _p(i["proj"]),
                                 C.c_float(fx), C.c_float(fy), C.c_float(i["tanfovx"]), C.c_float(i["tanfovy"]),
                                 _p(i["cam"]), _p(g["means2D"]), _p(g["conic"]), _p(g["colors_precomp"]),
                                 _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]))
    return g


def mark_visible(means3D, viewmatrix):
    L = lib()
    means3D = _f32(means3D).reshape(-1, 3)
    out = np.zeros(means3D.shape[0], np.uint8)
    L.oracle_mark_visible(means3D.shape[0], _p(means3D), _p(_f32(viewmatrix).reshape(16)), _p(out))
    return out.astype(bool)


def forward_scene(scene: dict, bg, **kw):
    """Convenience: scene dict from trase_b200.synth (CPU tensors)."""
    cam = scene["cam"]
    t = lambda x: x.detach().cpu().numpy()
    return forward(t(scene["means3D"]), t(scene["opacities"]), t(scene["sh_objs"]), t(cam["viewmatrix"]),
                   t(cam["projmatrix"]), t(cam["campos"]), cam["W"], cam["H"], cam["tanfovx"], cam["tanfovy"],
                   np.asarray(bg, np.float32), shs=t(scene["shs"]), sh_degree=scene["sh_degree"],
                   scales=t(scene["scales"]), rotations=t(scene["rotations"]), **kw)
