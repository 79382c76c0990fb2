#!/usr/bin/env bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_training_loop.py -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do
python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('neu3d', d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])"
done
