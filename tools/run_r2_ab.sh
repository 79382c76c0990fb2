#!/usr/bin/env bash
# occupancy variants of the compositing kernels (library variants built with -D switches / edited launch bounds)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp trase_b200/libtrase_b200.so /tmp/lib_base.so
run() {  # variant workload steps [env]
  [ $1 = base ] && cp /tmp/lib_base.so trase_b200/libtrase_b200.so || cp trase_b200/libtrase_b200_v$1.so trase_b200/libtrase_b200.so
  env $4 python bench.py --workload $2 --steps $3 --warmup 3 --no-cpu-baseline 2>gpurun_out/r2_ab_$1.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1 $4 $2', d['value'], d['ms_per_step'], 'render_fwd', d['stages']['render_fwd']['ms'], 'render_bwd', d['stages']['render_bwd']['ms'])" | tee -a gpurun_out/r2_aa_bwd_occupancy.txt
}
run base neu3d 30 A=1; run P1_3 neu3d 30 TRASE_B200_BWD_PPT=1; run P1_4 neu3d 30 TRASE_B200_BWD_PPT=1
run base hypernerf8 10 A=1; run P1_3 hypernerf8 10 TRASE_B200_BWD_PPT=1; run P1_4 hypernerf8 10 TRASE_B200_BWD_PPT=1
cp /tmp/lib_base.so trase_b200/libtrase_b200.so
python tools/host_vs_device.py hypernerf8 2>&1 | head -40 > gpurun_out/r2_ab_host_vs_device_hypernerf8.txt; head -3 gpurun_out/r2_ab_host_vs_device_hypernerf8.txt
