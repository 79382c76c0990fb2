#!/usr/bin/env bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py tests/test_zz_no_alloc.py -m gpu -x -q -s > gpurun_out/r2_af_multi_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_af_multi_tests.log
grep -A36 "all-reduced arena" gpurun_out/r2_af_multi_tests.log | head -40; tail -3 gpurun_out/r2_af_multi_tests.log
timeout 600 python -m pytest tests/test_gpu_mlp.py -x -q -s 2>&1 | tail -30
python tools/mlp_grad_spread.py 2>&1 | sed -n 1,24p
python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('neu3d', d['value'], d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()})"
