"""Where the time of one N-rank step goes: torch.profiler (CUPTI) on rank 0 over three neu3d steps at N ranks.
Prints every NCCL kernel with its duration and what ran on the compute stream at the same time, the exposed part of
the collective (time in which no other kernel of this rank overlapped it), and the per-kernel totals.
    python -m torch.distributed.run --nproc-per-node N tools/profile_step_multi.py [--workload neu3d] > profiles/...
(Timing under a profiler is never a bench value; this is for attribution only.)"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="neu3d")
ap.add_argument("--full-arena", action="store_true")
ap.add_argument("--timeline", action="store_true", help="also list every kernel of the middle step in launch order with the idle gap before it")
args = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
bench.Workload.full_arena = args.full_arena
wl = bench.Workload(args.workload, dev, rank, mlp_backend="native")
for _ in range(8):
    wl.step_resident()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        wl.step_resident()
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range is not None]
    ker = sorted([(e.time_range.start, e.time_range.end, e.name) for e in evs if "memcpy" not in e.name.lower()
                  and "memset" not in e.name.lower()], key=lambda x: x[0])
    t0 = ker[0][0]
    print("# %d ranks, workload %s, arena %d floats; times in us relative to the first kernel of the 3 profiled steps"
          % (world, args.workload, wl.arena.numel()))
    nccl = [k for k in ker if "nccl" in k[2].lower()]
    for s, e, n in nccl:
        others = [(max(s, a), min(e, b), m) for a, b, m in ker if "nccl" not in m.lower() and b > s and a < e]
        covered = 0.0
        cur = s
        for a, b, m in sorted(others):
            if b > cur:
                covered += b - max(a, cur)
                cur = max(cur, b)
        names = sorted({m.split("(")[0][:60] for _, _, m in others})
        print("NCCL %-60s start %9.1f dur %8.1f us  overlapped %8.1f us  exposed %8.1f us  with: %s"
              % (n[:60], s - t0, e - s, covered, (e - s) - covered, ", ".join(names)[:300]))
    print("\n# step span (3 steps): %.1f us" % (ker[-1][1] - t0))
    tot = {}
    for s, e, n in ker:
        k = n.split("(")[0][:70]
        c = tot.setdefault(k, [0.0, 0])
        c[0] += e - s
        c[1] += 1
    for k, (t, c) in sorted(tot.items(), key=lambda kv: -kv[1][0])[:40]:
        print("%10.1f us  %4d x  %s" % (t, c, k))
    if args.timeline:
        n3 = len(ker) // 3
        mid = ker[n3:2 * n3]
        print("\n# middle step, in order: start (us), duration, idle gap since the end of everything before it, name")
        busy_end = ker[n3 - 1][1]
        idle = 0.0
        for s, e, n in mid:
            gap = max(0.0, s - busy_end)
            idle += gap
            print("%9.1f %8.1f %6.1f  %s" % (s - mid[0][0], e - s, gap, n.split("(")[0][:90]))
            busy_end = max(busy_end, e)
        print("# idle inside the step: %.1f us of %.1f" % (idle, mid[-1][1] - mid[0][0]))
if world > 1:
    dist.destroy_process_group()
