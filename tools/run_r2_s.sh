#!/usr/bin/env bash
cd "$(dirname "$0")/.."
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/r2_s_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_s_tests.log
tail -15 $O/r2_s_tests.log
for v in 0 3 4; do
  TRASE_B200_FWD_MMA=$v timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/r2_s_mma$v.json 2>$O/r2_s_mma$v.err
  python -c "
import json;d=json.load(open('$O/r2_s_mma$v.json'));print('fwd_mma=$v', d['value'], d['e2e']['value'], d['stages']['render_fwd']['ms'])"
done
