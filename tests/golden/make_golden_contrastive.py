"""Generates tests/golden/contrastive.npz by importing the reference's own FEATURE-state functions from
/root/reference in the build container (CPU, fp32):   python tests/golden/make_golden_contrastive.py

  utils/feature_utils.py : get_pixel_weights, get_pixel_mask_correspondence_matrix,
                           get_features_correspondence_matrix            (:28-55)
  utils/loss_utils.py    : pixel_mask_correspondence_loss_soft_hard_positive / _soft_negative (:304-349),
                           loss_nnfm_style (:223-228)
  train.py:280-296       : feature-norm regulariser, bilinear resize to the mask resolution, the 'soft' loss sum

get_sample_pixel_and_mask (:17-26) hard-codes .cuda() and is a random draw; the sampled pixel / mask sets are
therefore drawn here with a seeded generator and stored as inputs.  loss_utils.py imports pytorch3d and torchvision
at module level for helpers this path never calls; neither ships in this image, so empty stand-in modules are
registered for exactly those imports first.

Two cases: one where both losses find pairs, one (features collapsed to one direction, all pixels in one mask)
where the negative loss finds none and returns 0.0."""
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
from utils.feature_utils import (get_features_correspondence_matrix, get_pixel_mask_correspondence_matrix,  # noqa: E402
                                 get_pixel_weights)
from utils.loss_utils import (loss_nnfm_style, pixel_mask_correspondence_loss_soft_hard_positive,  # noqa: E402
                              pixel_mask_correspondence_loss_soft_negative)

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "contrastive.npz")


def rect_masks(n, Hm, Wm, g):
    m = torch.zeros(n, Hm, Wm, dtype=torch.int64)
    for i in range(n):
        h = int(torch.randint(Hm // 8, Hm // 2, (1,), generator=g))
        w = int(torch.randint(Wm // 8, Wm // 2, (1,), generator=g))
        y = int(torch.randint(0, Hm - h, (1,), generator=g))
        x = int(torch.randint(0, Wm - w, (1,), generator=g))
        m[i, y:y + h, x:x + w] = 1
    return m


def main():
    out = {}
    for tag, seed in (("mixed", 5), ("no_negative", 6)):
        g = torch.Generator().manual_seed(seed)
        H, W, Hm, Wm, n_masks = 45, 70, 23, 35, 12
        masks = rect_masks(n_masks, Hm, Wm, g)
        feat = torch.randn(32, H, W, generator=g)
        if tag == "no_negative":
            masks[:] = 0
            masks[0] = 1                                              # every pixel shares mask 0: C == 1 everywhere
            feat = feat * 0.01 + torch.randn(32, 1, 1, generator=g)   # nearly collinear features
        sampled_mask = torch.rand(n_masks, generator=g) < 0.7
        if tag == "no_negative":
            sampled_mask[0] = True
        sampled_pixel = (torch.rand(Hm, Wm, generator=g) < 0.3) & (masks.sum(0) != 0)
        feat.requires_grad_(True)
        C = get_pixel_mask_correspondence_matrix(masks, sampled_pixel, sampled_mask)
        weights = get_pixel_weights(masks, sampled_pixel)
        norm_reg = (1 - feat.norm(dim=0, p=2).mean()) ** 2
        resized = torch.nn.functional.interpolate(feat.unsqueeze(0), (Hm, Wm), mode="bilinear").squeeze(0)
        C_F = get_features_correspondence_matrix(resized, sampled_pixel)
        pos = pixel_mask_correspondence_loss_soft_hard_positive(C=C, C_F=C_F, positive_th=0.75, weights=weights)
        neg = pixel_mask_correspondence_loss_soft_negative(C=C, C_F=C_F, negative_th=0.5, weights=weights)
        loss = pos + neg + 1.0 * norm_reg
        loss.backward()
        out[tag + "_masks"] = masks.numpy().astype(np.uint8)
        out[tag + "_feat"] = feat.detach().numpy()
        out[tag + "_sampled_pixel"] = sampled_pixel.numpy()
        out[tag + "_sampled_mask"] = sampled_mask.numpy()
        out[tag + "_C"] = C.numpy()
        out[tag + "_weights"] = weights.numpy()
        out[tag + "_C_F"] = C_F.detach().numpy()
        out[tag + "_vals"] = np.array([float(pos), float(neg), float(norm_reg), float(loss)], dtype=np.float64)
        out[tag + "_grad"] = feat.grad.numpy()
    g = torch.Generator().manual_seed(7)
    f1 = torch.randn(64, 90, generator=g, requires_grad=True)
    f2 = torch.randn(64, 120, generator=g)
    l = loss_nnfm_style(f1, f2)
    l.backward()
    out["nnfm_f1"], out["nnfm_f2"] = f1.detach().numpy(), f2.numpy()
    out["nnfm_val"] = np.array([float(l)], dtype=np.float64)
    out["nnfm_grad"] = f1.grad.numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})
    print({k: out[k] for k in out if k.endswith("_vals")})


if __name__ == "__main__":
    main()
