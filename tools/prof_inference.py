"""Kernel-time table of the inference workload (configs[3] shape: 2 M Gaussians, 1080p, forward only)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
w = bench.Workload("inference2m", dev, 0, "native")
def step():
    with torch.no_grad():
        return w.render(w.view, w.time)
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(4): step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=70))
