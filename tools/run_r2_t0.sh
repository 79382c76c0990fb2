#!/usr/bin/env bash
# end of round 2: full GPU suite, bench lines of every workload, reference arm at config 2
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $O/r2_t_gpu_tests.log
python bench.py --steps 30 --warmup 5 > $O/r2_t_bench_neu3d.json 2> $O/r2_t_bench_neu3d.err || tail -3 $O/r2_t_bench_neu3d.err
for w in hypernerf8 nnfm inference2m; do
  python bench.py --workload $w --steps $([ $w = neu3d -o $w = nnfm ] && echo 30 || echo 10) --warmup 5 --no-cpu-baseline > $O/r2_t_bench_$w.json 2> $O/r2_t_bench_$w.err || tail -3 $O/r2_t_bench_$w.err
done
python bench.py --impl reference --steps 10 --warmup 3 > $O/r2_t_ref_neu3d.json 2> $O/r2_t_ref_neu3d.err || tail -3 $O/r2_t_ref_neu3d.err
python - <<'P'
import json
for f in ("bench_neu3d","bench_hypernerf8","bench_nnfm","bench_inference2m","ref_neu3d"):
    try:
        d=json.loads(open("gpurun_out/r2_t_%s.json"%f).read().strip().splitlines()[-1])
        print(f, d["value"], d.get("unit"), d["ms_per_step"], "e2e", d["e2e"]["value"], "launches", d.get("gpu_launches"), "roofline", d.get("roofline",{}).get("frac"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e: print(f, "ERR", e)
P
