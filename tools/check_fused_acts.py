import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trase_b200 import _deform_native as dn
from trase_b200.deform import DeformNetwork
M = int(sys.argv[1])
torch.manual_seed(3)
net = DeformNetwork(backend="native").cuda()
g = torch.Generator(device="cuda").manual_seed(M)
x = (torch.rand(M, 3, device="cuda", generator=g) - 0.5) * 3
t = torch.full((M, 1), 0.41, device="cuda")
sv = {}
for fused in (False, True):
    dn.FUSED_FORWARD = fused
    outs = net(x, t)
    fn = outs[0].grad_fn
    while fn is not None and "DeformMLP" not in type(fn).__name__:
        fn = fn.next_functions[0][0]
    sv[fused] = [s.clone() for s in fn.saved_tensors]
    torch.cuda.synchronize()
a, b = sv[False], sv[True]
print("masks equal frac", (a[2] == b[2]).float().mean().item())
for i in range(8):
    d = (a[3 + i] - b[3 + i]).abs()
    rows = (d.max(dim=1).values > 1e-2).nonzero().flatten()
    print("layer", i, "max diff", d.max().item(), "bad rows", rows.numel(), rows[:8].tolist(), rows[-4:].tolist())
