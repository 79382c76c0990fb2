"""Fused L1+SSIM (trase_b200/loss_utils.py) against the reference's torch formulation (oracle/loss_oracle.py
run on the GPU) at the config-2 image size: ms per forward+backward, CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trase_b200 import loss_utils
from oracle import loss_oracle

C, H, W = 3, 1014, 1352
g = torch.Generator().manual_seed(0)
gt = torch.rand(C, H, W, generator=g).cuda()
img = (gt + 0.1 * torch.randn(C, H, W, generator=g).cuda()).clamp(0, 1).requires_grad_(True)


def timeit(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def ours():
    img.grad = None
    loss_utils.l1_ssim_loss(img, gt, 0.2).backward()


def ref():
    img.grad = None
    loss_oracle.l1_ssim_ref(img, gt, 0.2).backward()


for tf32 in (True, False):
    torch.backends.cudnn.allow_tf32 = tf32
    print(f"reference torch formulation (cudnn.allow_tf32={tf32}): {timeit(ref):.3f} ms")
print(f"fused: {timeit(ours):.3f} ms")
