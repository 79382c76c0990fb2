#!/usr/bin/env bash
# one GPU: full gpu test-suite, then the three training workloads on both arms
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests -m gpu -x -q -s > $O/r2_b_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_b_tests.log
tail -4 $O/r2_b_tests.log
for wl in neu3d hypernerf8 nnfm; do
  python bench.py --workload $wl --steps 20 --warmup 5 > $O/r2_b_bench_$wl.json 2> $O/r2_b_bench_$wl.err; echo "$wl ours rc=$?"
  python bench.py --workload $wl --steps 10 --warmup 3 --impl reference > $O/r2_b_ref_$wl.json 2> $O/r2_b_ref_$wl.err; echo "$wl ref rc=$?"
done
python bench.py --workload neu3d --steps 10 --warmup 3 --impl reference --ref-mlp tf32 > $O/r2_b_ref_neu3d_tf32.json 2> $O/r2_b_ref_neu3d_tf32.err
for f in $O/r2_b_bench_*.json $O/r2_b_ref_*.json; do echo "== $f"; head -c 600 $f; echo; done
tail -3 $O/r2_b_*.err
