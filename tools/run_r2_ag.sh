#!/usr/bin/env bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_training_loop.py tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2
for w in neu3d hypernerf8 inference2m; do
  python bench.py --workload $w --steps $([ $w = neu3d ] && echo 30 || echo 10) --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$w', d['value'], d['ms_per_step'], d.get('e2e',{}).get('value'))"
done
python tools/profile_step_multi.py --timeline 2>&1 | grep "deform_embed" | head -2
