#!/usr/bin/env bash
# end of round 2, ncu evidence part 2: full-set capture of one step's top kernels
cd "$(dirname "$0")/.."
O=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > $O/r2_t_plain2.log 2> $O/r2_t_plain2.err && \
ncu --set full --clock-control none --import-source on -k regex:"render_bwd|render_fwd|mlp_fused_kernel|wgrad_bf16" -s 15 -c 6 -o $O/r2_t_full $CMD > $O/r2_t_ncu2.log 2>&1
echo "full rc=$?"
ncu -i $O/r2_t_full.ncu-rep --page raw --csv > $O/r2_t_full_raw.csv 2> $O/r2_t_export.err
python tools/ncu_raw_summary.py $O/r2_t_full_raw.csv > $O/r2_t_ncu_summary.txt 2>&1; head -40 $O/r2_t_ncu_summary.txt
ls -la $O/r2_t_*; tail -3 $O/r2_t_ncu2.log
rm -f $O/r2_t_full.ncu-rep
