#!/usr/bin/env bash
cd "$(dirname "$0")/.."
timeout 300 python tools/fused_mlp_stalls.py 300000 1 2>&1 | tail -5 | cut -c1-420
timeout 600 python -m pytest tests/test_gpu_mlp.py -x -q 2>&1 | tail -1
python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items() if 'mlp' in k})"
