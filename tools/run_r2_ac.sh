#!/usr/bin/env bash
# checkpoint: full GPU suite, then the three training workloads and the host/device balance of hypernerf8
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2_ac_gpu_tests.log
for w in neu3d hypernerf8 nnfm; do
  python bench.py --workload $w --steps $([ $w = hypernerf8 ] && echo 10 || echo 30) --warmup 5 --no-cpu-baseline > gpurun_out/r2_ac_$w.json 2>gpurun_out/r2_ac_$w.err || tail -3 gpurun_out/r2_ac_$w.err
  python -c "
import json
d=json.loads(open('gpurun_out/r2_ac_$w.json').read().strip().splitlines()[-1])
print('$w', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {k:v['ms'] for k,v in d['stages'].items()})"
done
python tools/host_vs_device.py hypernerf8 2>&1 | head -1
