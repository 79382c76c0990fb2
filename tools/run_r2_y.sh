#!/usr/bin/env bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "smooth" 2>&1 | tail -3
python bench.py --workload hypernerf8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_y_hypernerf8.json 2>gpurun_out/r2_y.err || tail -5 gpurun_out/r2_y.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_y_hypernerf8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'])"
python tools/profile_step_multi.py --workload hypernerf8 2>&1 | sed -n 3,30p | cut -c1-110
