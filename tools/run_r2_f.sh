#!/usr/bin/env bash
# eight GPUs: the N=8 scaling line of config 2 and of config 3, and the N=8 step attribution (kept short: the call is
# charged eight-fold)
cd "$(dirname "$0")/.."
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
timeout 300 $TR --master-port 29518 bench.py --gpus 8 --steps 30 --warmup 5 2> $O/r2_f_bench_n8.err | grep '^{' > $O/r2_f_bench_n8.json; echo "n8 rc=$?"
timeout 300 $TR --master-port 29525 tools/profile_step_multi.py > $O/r2_f_profile_n8.txt 2> $O/r2_f_profile_n8.err; echo "profile rc=$?"
timeout 300 $TR --master-port 29522 bench.py --gpus 8 --steps 10 --warmup 3 --workload hypernerf8 2> $O/r2_f_bench_n8_hyper.err | grep '^{' > $O/r2_f_bench_n8_hyper.json; echo "hyper rc=$?"
for f in $O/r2_f_bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1], d["n_gpus"], d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"))
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
head -24 $O/r2_f_profile_n8.txt
