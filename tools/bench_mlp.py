"""DeformNetwork fwd+bwd at P rows: native vs torch (CUDA events)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from trase_b200 import _lib
from trase_b200.deform import DeformNetwork
dev = torch.device("cuda:0")
P = int(sys.argv[1]) if len(sys.argv) > 1 else 300000
for kv in sys.argv[2:]:
    k, v = kv.split("="); _lib.set_tuning(k, int(v))
torch.manual_seed(0)
net = DeformNetwork(backend="native").to(dev)
x = torch.randn(P, 3, device=dev); t = torch.full((P, 1), 0.3, device=dev)
go = [torch.randn(P, k, device=dev) for k in (3, 4, 3)]
def step():
    net.zero_grad(set_to_none=True)
    outs = net(x, t)
    torch.autograd.backward(list(outs), go)
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1, e2 = torch.cuda.Event(True), torch.cuda.Event(True), torch.cuda.Event(True)
n = 10
tf = tb = 0.0
for _ in range(n):
    net.zero_grad(set_to_none=True)
    e0.record(); outs = net(x, t); e1.record(); torch.autograd.backward(list(outs), go); e2.record()
    torch.cuda.synchronize(); tf += e0.elapsed_time(e1); tb += e1.elapsed_time(e2)
print("native MLP P=%d: fwd %.3f ms  bwd %.3f ms  total %.3f ms" % (P, tf / n, tb / n, (tf + tb) / n))
