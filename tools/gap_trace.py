"""Where the GPU idles inside a step: kernel timeline of a few resident steps, gaps > 3 us listed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from torch.profiler import profile, ProfilerActivity

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
w = bench.Workload("neu3d", dev, 0, "native")
mode = sys.argv[1] if len(sys.argv) > 1 else "resident"
fn = w.step_e2e if mode == "e2e" else w.step_resident
for _ in range(5):
    fn()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
busy = sum(e.time_range.end - e.time_range.start for e in ev)
span = ev[-1].time_range.end - t0
print(f"{len(ev)} kernels, span {span/4:.1f} us/step, busy {busy/4:.1f} us/step, idle {(span-busy)/4:.1f} us/step")
gaps = []
for a, b in zip(ev[:-1], ev[1:]):
    g = b.time_range.start - a.time_range.end
    if g > 3:
        gaps.append((g, a.name[:50], b.name[:50], a.time_range.end - t0))
small = sum(max(0, b.time_range.start - a.time_range.end) for a, b in zip(ev[:-1], ev[1:]) if b.time_range.start - a.time_range.end <= 3)
print(f"gaps <= 3 us: {small/4:.1f} us/step in total; gaps > 3 us: {sum(g[0] for g in gaps)/4:.1f} us/step")
for g in sorted(gaps, key=lambda g: -g[0])[:40]:
    print(f"{g[0]:8.1f} us at t={g[3]:9.1f}  after {g[1]:50s} before {g[2]}")

if len(sys.argv) > 2:
    # kernel sequence of the third step
    n = len(ev) // 4
    for e in ev[2 * n:3 * n]:
        print(f"{e.time_range.start - t0:10.1f} {e.time_range.end - e.time_range.start:8.1f} us  {e.name[:90]}")
