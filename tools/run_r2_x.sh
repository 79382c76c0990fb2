#!/usr/bin/env bash
# MLP + glue tests, then the neu3d bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_parity.py tests/test_gpu_training_loop.py -x -q 2>&1 | tail -4
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2_x_neu3d.json 2> gpurun_out/r2_x.err || tail -5 gpurun_out/r2_x.err
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2_x_neu3d.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], {k:v["ms"] for k,v in d["stages"].items()})
P
python tools/profile_step_multi.py --timeline 2>&1 | grep -A20 "middle step" | head -12
