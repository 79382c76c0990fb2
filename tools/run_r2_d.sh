#!/usr/bin/env bash
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r2_d_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_d_tests.log
tail -15 $O/r2_d_tests.log
for wl in neu3d hypernerf8 nnfm; do
  python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline > $O/r2_d_bench_$wl.json 2> $O/r2_d_bench_$wl.err; echo "$wl rc=$?"
done
python bench.py --workload neu3d --steps 20 --warmup 5 --no-cpu-baseline --full-arena > $O/r2_d_bench_neu3d_full.json 2> $O/r2_d_bench_neu3d_full.err
bash tools/sanitize.sh $O
