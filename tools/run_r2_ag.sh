#!/usr/bin/env bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_training_loop.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for w in neu3d nnfm; do
python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$w', d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])"
done
python tools/profile_step_multi.py --timeline 2>&1 | grep "assemble_sh_bwd\|assemble_bwd_kernel" | head -4
