#!/usr/bin/env bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mlp.py -x -q 2>&1 | tail -2
for w in hypernerf8 nnfm; do
  python bench.py --workload $w --steps $([ $w = hypernerf8 ] && echo 10 || echo 30) --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$w', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
done
python tools/host_vs_device.py hypernerf8 2>&1 | head -1
