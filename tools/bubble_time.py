"""CPU time between the end of the forward's host wait and the launch of the backward kernels (GPU idles meanwhile)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from trase_b200 import _raster, _lib
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
w = bench.Workload("neu3d", dev, 0, "native")
T = {}
of, ob = _raster.rasterize_forward, _raster.rasterize_backward
lib = _lib.load()
orig_bwd = lib.trase_b200_rasterize_backward
def f(*a, **k):
    r = of(*a, **k); T["fwd_done"] = time.perf_counter(); return r
def b(*a, **k):
    T["bwd_enter"] = time.perf_counter(); r = ob(*a, **k); T["bwd_exit"] = time.perf_counter(); return r
_raster.rasterize_forward, _raster.rasterize_backward = f, b
acc = [0.0, 0.0]
for i in range(25):
    w.step_resident()
    if i >= 5:
        acc[0] += T["bwd_enter"] - T["fwd_done"]; acc[1] += T["bwd_exit"] - T["bwd_enter"]
torch.cuda.synchronize()
print(f"forward wait done -> backward() entered: {acc[0] / 20 * 1e6:.1f} us;  inside rasterize_backward (python + launches): {acc[1] / 20 * 1e6:.1f} us")
