"""Generates tests/golden/deform_mlp.npz by importing the reference's own
DeformNetwork (utils/time_utils.py:60-131) from /root/reference in the build
container (CPU, fp32):  python tests/golden/make_golden_mlp.py

Holds a seeded state_dict, inputs, the three outputs and the parameter
gradients for a fixed upstream gradient — for the default model and the
`is_blender` variant."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.time_utils import DeformNetwork  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "deform_mlp.npz")


def main():
    out = {}
    for tag, blender in (("default", False), ("blender", True)):
        torch.manual_seed(11 if not blender else 12)
        net = DeformNetwork(is_blender=blender)
        M = 96
        g = torch.Generator().manual_seed(5)
        x = (torch.rand(M, 3, generator=g) - 0.5) * 3
        t = torch.full((M, 1), 0.41)
        go = [torch.randn(M, k, generator=g) for k in (3, 4, 3)]
        outs = net(x, t)
        torch.autograd.backward(list(outs), go)
        out[f"{tag}_x"], out[f"{tag}_t"] = x.numpy(), t.numpy()
        for i, o in enumerate(outs):
            out[f"{tag}_out{i}"] = o.detach().numpy()
            out[f"{tag}_go{i}"] = go[i].numpy()
        # Parameters are NOT stored: torch.manual_seed(seed) + construction order reproduce them (the test
        # checks the float64 sum of every tensor recorded here).  Gradients are stored for the layers with
        # distinct shapes (first, skip, one hidden, last, heads) and every bias.
        keep = ("linear.0.", "linear.3.weight", "linear.5.", "linear.7.", "gaussian_", "timenet", "bias")
        for n, p in net.named_parameters():
            out[f"{tag}_psum_{n}"] = np.float64(p.detach().double().sum().item())
            if any(k in n for k in keep):
                out[f"{tag}_grad_{n}"] = p.grad.numpy()
        out[f"{tag}_seed"] = 11 if not blender else 12
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT))


if __name__ == "__main__":
    main()
