#!/usr/bin/env bash
cd "$(dirname "$0")/.."
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_h_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_h_tests.log
tail -25 $O/r2_h_tests.log
for wl in neu3d hypernerf8; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline > $O/r2_h_bench_$wl.json 2> $O/r2_h_bench_$wl.err; echo "$wl rc=$?"
done
timeout 300 python bench.py --workload hypernerf8 --steps 10 --warmup 3 --no-cpu-baseline --torch-feature-loss > $O/r2_h_bench_hypernerf8_torchloss.json 2> $O/r2_h_bench_hypernerf8_torchloss.err
python - <<'PY'
import json
for wl in ("neu3d","hypernerf8","hypernerf8_torchloss"):
    try:
        d=json.load(open("gpurun_out/r2_h_bench_%s.json"%wl))
        print(wl, d["value"], d["e2e"]["value"], d["ms_per_step"], {k:v["ms"] for k,v in d["stages"].items()})
    except Exception as e: print(wl, "ERR", e)
PY
