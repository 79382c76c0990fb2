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
go = [torch.randn(M, k, device="cuda", generator=g) for k in (3, 4, 3)]
res = {}
import time
for mode in ("torch", "chain", "fused", "fused2"):
    net.backend = "torch" if mode == "torch" else "native"
    dn.FUSED_FORWARD = mode.startswith("fused")
    dn.FUSED_BACKWARD = mode.startswith("fused")
    net.zero_grad()
    outs = net(x, t)
    torch.autograd.backward(list(outs), go)
    res[mode] = [p.grad.detach().clone() for p in net.parameters()]
names = [n for n, _ in net.named_parameters()]
for i, n in enumerate(names):
    ref = res["torch"][i]
    sc = ref.abs().max().item() + 1e-12
    print(f"{n:28s} chain-vs-fp32 {(res['chain'][i]-ref).abs().max().item()/sc:.2e}  fused-vs-fp32 {(res['fused'][i]-ref).abs().max().item()/sc:.2e}  fused-vs-chain {(res['fused'][i]-res['chain'][i]).abs().max().item()/sc:.2e}  fused-rerun {(res['fused'][i]-res['fused2'][i]).abs().max().item()/sc:.2e}")

for mode in ("chain", "fused"):
    dn.FUSED_FORWARD = dn.FUSED_BACKWARD = mode == "fused"
    for _ in range(3):
        net.zero_grad(); outs = net(x, t); torch.autograd.backward(list(outs), go)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        net.zero_grad(); outs = net(x, t); torch.autograd.backward(list(outs), go)
    e1.record(); torch.cuda.synchronize()
    print(mode, f"fwd+bwd {e0.elapsed_time(e1) / 10:.3f} ms")
