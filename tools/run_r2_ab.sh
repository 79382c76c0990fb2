#!/usr/bin/env bash
# more occupancy variants of the compositing kernels (library variants built with -D switches, see csrc/render.cu)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp trase_b200/libtrase_b200.so /tmp/lib_base.so
run() {  # variant workload steps
  [ $1 = base ] && cp /tmp/lib_base.so trase_b200/libtrase_b200.so || cp trase_b200/libtrase_b200_v$1.so trase_b200/libtrase_b200.so
  python bench.py --workload $2 --steps $3 --warmup 3 --no-cpu-baseline 2>gpurun_out/r2_ab_$1.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1 $2', d['value'], d['ms_per_step'], 'render_fwd', d['stages']['render_fwd']['ms'], 'render_bwd', d['stages']['render_bwd']['ms'])" | tee -a gpurun_out/r2_aa_bwd_occupancy.txt
}
run base neu3d 30; run W5 neu3d 30
run base hypernerf8 10; run F5 hypernerf8 10; run W5 hypernerf8 10
run base nnfm 30; run C7 nnfm 30; run C8 nnfm 30
cp /tmp/lib_base.so trase_b200/libtrase_b200.so
