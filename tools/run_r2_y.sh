#!/usr/bin/env bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "contrastive or smooth or feature" 2>&1 | tail -3
python bench.py --workload hypernerf8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_y_hypernerf8.json 2>gpurun_out/r2_y.err || tail -5 gpurun_out/r2_y.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_y_hypernerf8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e'])"
python tools/profile_step_multi.py --workload hypernerf8 --timeline > gpurun_out/r2_z_timeline_hypernerf8.txt 2>&1; tail -1 gpurun_out/r2_z_timeline_hypernerf8.txt
sed -n 3,16p gpurun_out/r2_z_timeline_hypernerf8.txt | cut -c1-100
bash tools/run_r2_aa.sh
python tools/profile_step_multi.py --workload nnfm --timeline > gpurun_out/r2_z_timeline_nnfm.txt 2>&1; tail -1 gpurun_out/r2_z_timeline_nnfm.txt
