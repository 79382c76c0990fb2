#!/usr/bin/env bash
cd "$(dirname "$0")/.."
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_j_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_j_tests.log
tail -12 $O/r2_j_tests.log
timeout 300 python bench.py --workload hypernerf8 --steps 20 --warmup 5 --no-cpu-baseline > $O/r2_j_bench_hypernerf8.json 2> $O/r2_j_bench_hypernerf8.err; echo "hyper rc=$?"
timeout 300 python tools/profile_step_multi.py --workload hypernerf8 > $O/r2_j_profile_hyper.txt 2> $O/r2_j_profile_hyper.err
python - <<'PY'
import json
for wl in ("hypernerf8",):
    try:
        d=json.load(open("gpurun_out/r2_j_bench_%s.json"%wl))
        print(wl, d["value"], d["e2e"]["value"], d["ms_per_step"], {k:v["ms"] for k,v in d["stages"].items()})
    except Exception as e: print(wl, "ERR", e)
PY
head -45 $O/r2_j_profile_hyper.txt
