"""Soak test of the CTA-pair MLP kernels: many forward+backward passes over varied row counts (tile counts that are
odd / even / below the SM count / ragged), checking finiteness and run-to-run determinism of the forward."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trase_b200.deform import DeformNetwork
torch.manual_seed(0)
net = DeformNetwork(backend="native").cuda()
sizes = [1, 2, 127, 128, 129, 255, 257, 1000, 4096, 18944, 18945, 37888, 37889, 50000, 100003, 300000, 299999]
t0 = time.time()
n = 0
g = torch.Generator(device="cuda").manual_seed(1)
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 12):
    for M in sizes:
        x = (torch.rand(M, 3, device="cuda", generator=g) - 0.5) * 3
        t = torch.full((M, 1), 0.1 * (rep % 10), device="cuda")
        net.zero_grad()
        o = net(x, t)
        o2 = [a.detach().clone() for a in o]
        torch.autograd.backward(list(o), [torch.ones_like(a) for a in o])
        with torch.no_grad():
            o3 = net(x, t)
        assert all(torch.isfinite(a).all() for a in o2)
        assert all(torch.equal(a, b) for a, b in zip(o2, o3)), ("forward not deterministic", M, rep)
        assert all(torch.isfinite(p.grad).all() for p in net.parameters())
        n += 1
torch.cuda.synchronize()
print(f"{n} forward+backward passes ok in {time.time() - t0:.1f} s")
