"""Reference-compatible `point_cloud.ply` files (on-disk format).

Same element / property layout as GaussianModel.save_ply / load_ply
(scene/gaussian_model.py:317-386, 388-455): one `vertex` element of float32
properties in the order
  x y z nx ny nz f_dc_0..2 f_rest_0..(3*(K-1)-1) opacity scale_0..2 rot_0..3 gaussian_feats_0..31 [cls]
with f_dc / f_rest / gaussian_feats stored channel-major (the reference transposes
[P,K,3] -> [P,3,K] before flattening).  Binary little-endian, read and written with
numpy only (the reference uses the `plyfile` package, which is not required here).
"""
from __future__ import annotations

import os
from typing import Dict, List, Tuple

import numpy as np
import torch


def attribute_names(n_rest: int, n_feat: int, with_cls: bool = False) -> List[str]:
    """construct_list_of_attributes (scene/gaussian_model.py:317-333)."""
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)]
    names += [f"f_rest_{i}" for i in range(3 * n_rest)]
    names += ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    names += [f"gaussian_feats_{i}" for i in range(n_feat)]
    return names + (["cls"] if with_cls else [])


def save_ply(path: str, pc, cls: torch.Tensor = None) -> None:
    """Write raw (pre-activation) parameters of a GaussianParams-like object (`_xyz`, `_features_dc`, ...)."""
    d = lambda t: t.detach().float().cpu()
    xyz = d(pc._xyz).numpy()
    P = xyz.shape[0]
    f_dc = d(pc._features_dc).transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    f_rest = d(pc._features_rest).transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    feats = d(pc._gaussian_features).transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, d(pc._opacity).numpy(), d(pc._scaling).numpy(), d(pc._rotation).numpy(),
            feats]
    if cls is not None:
        cols.append(d(cls).reshape(P, 1).numpy())
    names = attribute_names(pc._features_rest.shape[1], feats.shape[1], cls is not None)
    table = np.concatenate(cols, axis=1).astype("<f4")
    assert table.shape[1] == len(names)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(table).tobytes())


_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
              "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def read_vertex_table(path: str) -> Dict[str, np.ndarray]:
    """Minimal PLY reader: the first element of a binary_little_endian (or ascii) file as {property: array}."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fmt, count, props, in_first, seen = None, None, [], False, 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header without end_header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                seen += 1
                in_first = seen == 1
                if in_first:
                    count = int(tok[2])
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError("list properties are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        dt = np.dtype(props)
        if fmt == "binary_little_endian":
            data = np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)
        elif fmt == "ascii":
            data = np.loadtxt(f, dtype=dt, max_rows=count, ndmin=1)
        else:
            raise ValueError(f"unsupported PLY format {fmt}")
    return {name: np.asarray(data[name]) for name, _ in props}


def load_ply(path: str, device="cuda", max_sh_degree: int = 3, feature_dim: int = 32) -> Tuple[object, torch.Tensor]:
    """-> (GaussianParams, cls or None); mirrors GaussianModel.load_ply (scene/gaussian_model.py:388-455):
    properties are looked up by name, f_rest_* / scale_* / rot_* sorted by their numeric suffix."""
    from .render import GaussianParams
    v = read_vertex_table(path)
    P = v["x"].shape[0]
    by_suffix = lambda pre: sorted((k for k in v if k.startswith(pre)), key=lambda k: int(k.split("_")[-1]))
    col = lambda names: np.stack([v[n].astype(np.float32) for n in names], axis=1) if names else np.zeros((P, 0), np.float32)
    rest_names = by_suffix("f_rest_")
    n_rest = (max_sh_degree + 1) ** 2 - 1
    if len(rest_names) != 3 * n_rest:
        raise ValueError(f"expected {3 * n_rest} f_rest_* properties for SH degree {max_sh_degree}, found {len(rest_names)}")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
    f_dc = col(["f_dc_0", "f_dc_1", "f_dc_2"]).reshape(P, 3, 1)
    f_rest = col(rest_names).reshape(P, 3, n_rest)
    feat_names = [f"gaussian_feats_{i}" for i in range(feature_dim) if f"gaussian_feats_{i}" in v]
    feats = np.zeros((P, feature_dim, 1), np.float32)
    if feat_names:
        feats[:, :len(feat_names), 0] = col(feat_names)
    pc = GaussianParams(t(col(["x", "y", "z"])), t(f_dc).transpose(1, 2).contiguous(), t(f_rest).transpose(1, 2).contiguous(),
                        t(col(by_suffix("scale_"))), t(col(by_suffix("rot_"))), t(col(["opacity"])),
                        t(feats).transpose(1, 2).contiguous(), active_sh_degree=max_sh_degree, max_sh_degree=max_sh_degree)
    cls = t(v["cls"].astype(np.float32).reshape(P, 1)) if "cls" in v else None
    return pc, cls
