"""Generates tests/golden/l1_ssim.npz by importing the reference's own loss
functions (utils/loss_utils.py:30-31, 46-86) from /root/reference in the build
container (CPU, fp32):  python tests/golden/make_golden_loss.py

loss_utils.py imports pytorch3d and torchvision at module level for helpers this
path never calls; neither ships in this image, so the script registers empty
stand-in modules for exactly those imports before importing the file.

Holds two seeded image pairs (one ragged size that is not a multiple of the
32-pixel tile, one with a flat region so sign(0) and sigma = 0 occur), the three
loss values and the image gradient of (1 - l) * L1 + l * (1 - SSIM), l = 0.2."""
import os
import sys
import types

import numpy as np
import torch

for name in ("pytorch3d", "pytorch3d.ops", "torchvision", "torchvision.utils"):
    if name not in sys.modules:
        try:
            __import__(name)
        except Exception:
            m = types.ModuleType(name)
            m.save_image = None
            sys.modules[name] = m
sys.modules["pytorch3d"].ops = sys.modules["pytorch3d.ops"]
sys.path.insert(0, "/root/reference")
from utils.loss_utils import l1_loss, ssim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "l1_ssim.npz")


def main():
    out = {}
    for tag, (C, H, W), seed in (("ragged", (3, 45, 70), 3), ("flat", (3, 33, 64), 4)):
        g = torch.Generator().manual_seed(seed)
        gt = torch.rand(C, H, W, generator=g)
        img = (gt + 0.15 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
        if tag == "flat":
            img[:, :20, :24] = 0.25
            gt[:, :20, :24] = 0.25          # identical flat patch: |x-y| = 0, variances 0
        img.requires_grad_(True)
        l1, ss = l1_loss(img, gt), ssim(img, gt)
        loss = 0.8 * l1 + 0.2 * (1.0 - ss)
        loss.backward()
        out[f"{tag}_img"], out[f"{tag}_gt"] = img.detach().numpy(), gt.numpy()
        out[f"{tag}_vals"] = np.array([l1.item(), ss.item(), loss.item()], dtype=np.float64)
        out[f"{tag}_grad"] = img.grad.numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
