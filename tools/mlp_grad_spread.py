"""Run-to-run spread of the native DeformNetwork's parameter gradients in the multi-GPU parity test's set-up (one GPU):
the same two views summed by the same code, twice; also with the arena's direct accumulation and with fp32 side tensors."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import test_gpu_multi as T
from trase_b200 import _deform_native as dn
from trase_b200.distributed import GradArena

dev = torch.device("cuda:0")


def run(arena=False):
    sc, pc, net = T._setup(dev, "native")
    if arena:
        params = list(pc.parameters()) + list(net.parameters())
        a = GradArena(params, early=list(pc.parameters()))
        a.zero()
    for v in range(2):
        T._backward_one_view(pc, net, v, dev)
    torch.cuda.synchronize()
    return {n: p.grad.detach().clone() for n, p in net.named_parameters()}


for bf16 in (True, False):
    dn.BF16_WGRAD = bf16
    a, b, c = run(), run(), run(arena=True)
    print("BF16_WGRAD =", bf16, " (max-norm relative difference: plain vs plain | plain vs arena)")
    for n in a:
        s = float(a[n].abs().max())
        print("  %-28s %.2e   %.2e" % (n, float((a[n] - b[n]).abs().max()) / s, float((a[n] - c[n]).abs().max()) / s))
