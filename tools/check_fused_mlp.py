"""Fused single-kernel MLP forward against the layer-by-layer GEMM chain (same TF32 arithmetic)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trase_b200 import _deform_native as dn
from trase_b200.deform import DeformNetwork

torch.manual_seed(0)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
net = DeformNetwork(backend="native").cuda()
x = (torch.rand(M, 3, device="cuda") - 0.5) * 3
t = torch.full((M, 1), 0.37, device="cuda")
outs, grads = {}, {}
for fused in (False, True):
    dn.FUSED_FORWARD = fused
    net.zero_grad()
    o = net(x, t)
    torch.cuda.synchronize()
    g = torch.Generator(device="cuda").manual_seed(1)
    go = [torch.randn(o_.shape, device="cuda", generator=g) for o_ in o]
    torch.autograd.backward(list(o), go)
    torch.cuda.synchronize()
    outs[fused] = [a.detach().clone() for a in o]
    grads[fused] = [p.grad.detach().clone() for p in net.parameters()]
for a, b in zip(outs[False], outs[True]):
    print("out max abs diff", (a - b).abs().max().item(), "scale", a.abs().max().item())
md = max(((a - b).abs().max() / (a.abs().max() + 1e-12)).item() for a, b in zip(grads[False], grads[True]))
print("max rel grad diff", md)
for fused in (False, True):
    dn.FUSED_FORWARD = fused
    with torch.no_grad():
        for _ in range(3):
            net(x, t)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            net(x, t)
        e1.record()
        torch.cuda.synchronize()
    print("fused" if fused else "chain", f"forward {e0.elapsed_time(e1) / 10:.3f} ms")

import ctypes
from trase_b200 import _lib
lib = _lib.load()
buf = torch.zeros(148 * 8, dtype=torch.int64, device="cuda")
lib.trase_b200_mlp_fused_set_prof.argtypes = [ctypes.c_void_p]
lib.trase_b200_mlp_fused_set_prof(buf.data_ptr())
dn.FUSED_FORWARD = True
with torch.no_grad():
    net(x, t)
torch.cuda.synchronize()
lib.trase_b200_mlp_fused_set_prof(None)
for name, sl in (("leader", slice(0, 148, 2)), ("follower", slice(1, 148, 2))):
    b = buf.view(148, 8)[sl].double().mean(0).tolist()
    print(name, "cycles/CTA: total %.0f | warp-1 thread waits: tmem_empty %.0f a_ready %.0f w_full %.0f store %.0f | epilogue warp4 waits: tmem_full %.0f h_free %.0f" % tuple(b[:7]))
