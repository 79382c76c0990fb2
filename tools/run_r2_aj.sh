#!/usr/bin/env bash
# soak of the bit-exact parity file and the no-alloc sort with programmatic dependent launch: repeated runs must all pass
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/r2_aj_soak.txt
for i in 1 2 3 4 5 6; do
  timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_gpu_mlp.py -x -q -p no:randomly 2>&1 | tail -1 | tee -a gpurun_out/r2_aj_soak.txt
done
python tools/soak_mlp.py 2>&1 | tail -3 | tee -a gpurun_out/r2_aj_soak.txt
