#!/usr/bin/env bash
# Variants of the whole-network MLP kernels built with -D switches (see the top of csrc/mlp_fused.cu): stall attribution +
# stage times for each library trase_b200/libtrase_b200_v<tag>.so named on the command line, then the MLP tests on it.
#   bash tools/run_r2_u.sh base p1f0 p0f1 p0f0
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp trase_b200/libtrase_b200.so /tmp/lib_base.so
for v in "$@"; do
  [ $v = base ] && cp /tmp/lib_base.so trase_b200/libtrase_b200.so || cp trase_b200/libtrase_b200_v$v.so trase_b200/libtrase_b200.so
  echo "== variant $v" | tee -a gpurun_out/r2_u_variants.txt
  timeout 180 python tools/fused_mlp_stalls.py 300000 1 2>&1 | tail -5 | tee -a gpurun_out/r2_u_variants.txt
  [ $v = base ] || timeout 600 python -m pytest tests/test_gpu_mlp.py -x -q 2>&1 | tail -1 | tee -a gpurun_out/r2_u_variants.txt
done
cp /tmp/lib_base.so trase_b200/libtrase_b200.so
