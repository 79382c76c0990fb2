#!/usr/bin/env bash
cd "$(dirname "$0")/.."
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_mlp.py -m gpu -x -q -s > $O/r2_e_mlp_tests.log 2>&1; echo "pytest mlp rc=$?" >> $O/r2_e_mlp_tests.log
tail -45 $O/r2_e_mlp_tests.log
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_mlp.py > $O/r2_e_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_e_tests.log
tail -5 $O/r2_e_tests.log
timeout 300 python bench.py --workload neu3d --steps 20 --warmup 5 --no-cpu-baseline > $O/r2_e_bench_neu3d.json 2> $O/r2_e_bench_neu3d.err; echo "neu3d rc=$?"
timeout 300 python bench.py --workload nnfm --steps 20 --warmup 5 --no-cpu-baseline > $O/r2_e_bench_nnfm.json 2> $O/r2_e_bench_nnfm.err; echo "nnfm rc=$?"
python - <<'PY'
import json
for wl in ("neu3d","nnfm"):
    try:
        d=json.load(open("gpurun_out/r2_e_bench_%s.json"%wl))
        print(wl, d["value"], d["e2e"]["value"], {k:v["ms"] for k,v in d["stages"].items()})
    except Exception as e: print(wl, "ERR", e)
PY
