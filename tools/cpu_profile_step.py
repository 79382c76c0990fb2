"""cProfile of the host side of e2e steps (which Python/ctypes calls cost the most between kernels)."""
import os, sys, cProfile, pstats, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
w = bench.Workload("neu3d", dev, 0, "native")
for _ in range(5):
    w.step_e2e()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    w.step_e2e()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28)
print(s.getvalue()[:6000])
