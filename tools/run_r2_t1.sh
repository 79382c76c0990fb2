#!/usr/bin/env bash
# end of round 2, ncu evidence part 1: launch list of the bench step (kernel shares)
cd "$(dirname "$0")/.."
O=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > $O/r2_t_plain.log 2> $O/r2_t_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 192 --csv --log-file $O/r2_t_launches_bench_steps2.csv $CMD > $O/r2_t_ncu1.log 2>&1
echo "launch list rc=$?"
python tools/launch_summary.py $O/r2_t_launches_bench_steps2.csv > $O/r2_t_launch_summary.txt 2>&1; head -12 $O/r2_t_launch_summary.txt
