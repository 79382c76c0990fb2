#!/usr/bin/env bash
# neu3d bench with the compositing kernels' pixels-per-thread switches, and the kernel timeline of one step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2_w_base.json 2> gpurun_out/r2_w_base.err || tail -5 gpurun_out/r2_w_base.err
TRASE_B200_BWD_PPT=1 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2_w_bwd1.json 2>/dev/null
TRASE_B200_FWD_PPT=1 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2_w_fwd1.json 2>/dev/null
python tools/profile_step_multi.py --timeline > gpurun_out/r2_w_timeline_neu3d.txt 2>&1
python - <<'P'
import json
for t in ("base","bwd1","fwd1"):
    try:
        d=json.loads(open("gpurun_out/r2_w_%s.json"%t).read().strip().splitlines()[-1])
        print(t, d["value"], d["ms_per_step"], d["e2e"]["value"], {k:v["ms"] for k,v in d["stages"].items()})
    except Exception as e: print(t, "failed", e)
P
