"""Generates the golden vectors in tests/golden/*.npz by running the UNMODIFIED
reference CUDA rasterizer (oracle/_ref, built by oracle/build_ref.sh) on a B200:

    gpurun -- python tests/golden/make_golden.py gpurun_out/golden
    cp gpurun_out/golden/*.npz tests/golden/

Each file holds the inputs (so it is self-contained) and every output and
intermediate the parity tests compare.  Gradients are the mean of 5 reference
runs (the reference accumulates with float atomics in nondeterministic order).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402
from trase_b200 import synth  # noqa: E402

CASES = {
    # name: (P, W, H, seed, bg, mode)
    "sh_deg3_bg0": (1200, 56, 40, 11, 0.0, "sh"),
    "sh_deg1_bg1": (700, 40, 48, 12, 1.0, "sh1"),
    "precomp_bg05": (600, 33, 47, 13, 0.5, "precomp"),
}


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    dev = torch.device("cuda:0")
    for name, (P, W, H, seed, bgv, mode) in CASES.items():
        sc_cpu = synth.make_scene(P, W, H, seed, target_radius_px=5.0)
        if mode == "sh1":
            sc_cpu["sh_degree"] = 1
        g = torch.Generator().manual_seed(seed + 100)
        colors_precomp = cov3D_precomp = None
        if mode == "precomp":
            colors_precomp = torch.rand(P, 3, generator=g)
            # a valid covariance: A A^T scaled to splat size
            A = torch.randn(P, 3, 3, generator=g) * 0.02
            S = A @ A.transpose(1, 2)
            cov3D_precomp = torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1)
            cov3D_precomp = cov3D_precomp.contiguous()
        sc = synth.scene_to(sc_cpu, dev)
        up = {k: v.to(dev) for k, v in synth.make_upstream_grads(W, H, seed).items()}
        bg = torch.full((3,), bgv, device=dev)
        cp = None if colors_precomp is None else colors_precomp.to(dev)
        c3 = None if cov3D_precomp is None else cov3D_precomp.to(dev)
        fw = ref_loader.run_forward(sc, bg, colors_precomp=cp, cov3D_precomp=c3)
        grads = [ref_loader.run_backward(sc, bg, fw, up["dL_dcolor"], up["dL_dobjects"], colors_precomp=cp,
                                         cov3D_precomp=c3) for _ in range(5)]
        out = dict(P=P, W=W, H=H, seed=seed, bg=bgv, sh_degree=sc_cpu["sh_degree"],
                   num_rendered=int(fw["num_rendered"]))
        cam = sc_cpu["cam"]
        for k in ("means3D", "scales", "rotations", "opacities", "shs", "sh_objs"):
            out["in_" + k] = sc_cpu[k].numpy()
        for k in ("viewmatrix", "projmatrix", "campos"):
            out["in_" + k] = cam[k].numpy()
        out["in_tanfovx"], out["in_tanfovy"] = cam["tanfovx"], cam["tanfovy"]
        if colors_precomp is not None:
            out["in_colors_precomp"] = colors_precomp.numpy()
            out["in_cov3D_precomp"] = cov3D_precomp.numpy()
        out["in_dL_dcolor"] = up["dL_dcolor"].cpu().numpy()
        out["in_dL_dobjects"] = up["dL_dobjects"].cpu().numpy()
        for k in ("color", "objects", "depth", "radii", "depths", "clamped", "means2D", "cov3D", "conic_opacity", "rgb",
                  "tiles_touched", "accum_alpha", "n_contrib", "ranges", "point_list", "point_list_keys"):
            out["ref_" + k] = fw[k].cpu().numpy()
        for k in grads[0]:
            out["ref_grad_" + k] = (sum(x[k].double() for x in grads) / len(grads)).float().cpu().numpy()
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **out)
        print(name, "R=%d visible=%d" % (out["num_rendered"], int((fw["radii"] > 0).sum())))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
