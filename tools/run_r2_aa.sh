#!/usr/bin/env bash
# render_bwd at higher occupancy: -DTB_BWD_G_MINB / TB_BWD_G_BATCH (GAUSSIAN-state variant, neu3d bench) and
# -DTB_BWD_F_MINB (FEATURE-state variant, hypernerf8 bench) variants of the library
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp trase_b200/libtrase_b200.so /tmp/lib_base.so
for v in base b4_256_f3 b4_128_f3 b5_128_f3 b6_64_f3; do
  [ $v = base ] && cp /tmp/lib_base.so trase_b200/libtrase_b200.so || cp trase_b200/libtrase_b200_v$v.so trase_b200/libtrase_b200.so
  python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v neu3d', d['value'], d['ms_per_step'], 'render_bwd', d['stages']['render_bwd']['ms'])" | tee -a gpurun_out/r2_aa_bwd_occupancy.txt
done
for v in base b4_256_f4; do
  [ $v = base ] && cp /tmp/lib_base.so trase_b200/libtrase_b200.so || cp trase_b200/libtrase_b200_v$v.so trase_b200/libtrase_b200.so
  python bench.py --workload hypernerf8 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v hypernerf8', d['value'], d['ms_per_step'], 'render_bwd', d['stages']['render_bwd']['ms'])" | tee -a gpurun_out/r2_aa_bwd_occupancy.txt
done
cp trase_b200/libtrase_b200_vb4_256_f4.so trase_b200/libtrase_b200.so
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -2
cp /tmp/lib_base.so trase_b200/libtrase_b200.so
