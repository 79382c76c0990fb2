"""How close the host is to being the bottleneck: wall time to LAUNCH a step (no synchronisation inside the loop other
than the forward's own readback wait) against the device time of the step, plus a cProfile of the resident step."""
import cProfile
import io
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "neu3d"
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
w = bench.Workload(name, dev, 0, "native")
for _ in range(8):
    w.step_resident()
torch.cuda.synchronize()
K = 30
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(K):
    w.step_resident()
e1.record()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("%s: host loop %.3f ms/step (returns while the GPU still has %.3f ms queued), device %.3f ms/step"
      % (name, (t1 - t0) / K * 1e3, (t2 - t1) * 1e3, e0.elapsed_time(e1) / K))
pr = cProfile.Profile()
pr.enable()
for _ in range(K):
    w.step_resident()
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(40)
print(s.getvalue()[:9000])
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumtime").print_stats(45)
print(s.getvalue()[:9000])
