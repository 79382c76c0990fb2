#!/usr/bin/env bash
# ncu evidence for round 2: launch list of the bench step and a full-set capture of the path's top kernels
cd "$(dirname "$0")/.."
O=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > $O/r2_k_plain.log 2> $O/r2_k_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400 --csv --log-file $O/r2_k_launches_bench_steps2.csv $CMD > $O/r2_k_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"render_bwd|render_fwd|mlp_fused_kernel|wgrad_bf16" -s 42 -c 14 -o $O/r2_k_full $CMD > $O/r2_k_ncu2.log 2>&1
echo "full rc=$?"
ncu -i $O/r2_k_full.ncu-rep --page raw --csv > $O/r2_k_full_raw.csv 2> $O/r2_k_export.err
ls -la $O/r2_k_*; tail -3 $O/r2_k_ncu2.log
rm -f $O/r2_k_full.ncu-rep
