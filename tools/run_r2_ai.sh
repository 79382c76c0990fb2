#!/usr/bin/env bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp trase_b200/libtrase_b200.so /tmp/lib_base.so
cp trase_b200/libtrase_b200_vEARLY.so trase_b200/libtrase_b200.so
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee -a gpurun_out/r2_ai_pdl_early.txt
for w in neu3d hypernerf8; do
 for v in base EARLY base EARLY; do
  [ $v = base ] && cp /tmp/lib_base.so trase_b200/libtrase_b200.so || cp trase_b200/libtrase_b200_v$v.so trase_b200/libtrase_b200.so
  python bench.py --workload $w --steps $([ $w = hypernerf8 ] && echo 10 || echo 30) --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v $w', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {k:v['ms'] for k,v in d['stages'].items()})" | tee -a gpurun_out/r2_ai_pdl_early.txt
 done
done
cp /tmp/lib_base.so trase_b200/libtrase_b200.so
