"""Kernel-time table of the end-to-end step (torch profiler), for finding what the e2e leg adds."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from torch.profiler import profile, ProfilerActivity

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
w = bench.Workload("neu3d", dev, 0, "native")
mode = sys.argv[1] if len(sys.argv) > 1 else "e2e"
fn = w.step_e2e if mode == "e2e" else w.step_resident
for _ in range(5):
    fn()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=60))
