#!/usr/bin/env bash
# eight GPUs: scaling lines of every workload, and the N=8 step attribution
cd "$(dirname "$0")/.."
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4; do
  timeout 300 $TR --nproc-per-node $n --master-port 2951$n bench.py --gpus $n --steps 30 --warmup 5 2> $O/r2_f_bench_n$n.err | grep '^{' > $O/r2_f_bench_n$n.json; echo "n$n rc=$?"
done
timeout 300 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 30 --warmup 5 --full-arena 2> $O/r2_f_bench_n8_full.err | grep '^{' > $O/r2_f_bench_n8_full.json
timeout 300 $TR --nproc-per-node 8 --master-port 29522 bench.py --gpus 8 --steps 10 --warmup 3 --workload hypernerf8 2> $O/r2_f_bench_n8_hyper.err | grep '^{' > $O/r2_f_bench_n8_hyper.json
timeout 300 $TR --nproc-per-node 8 --master-port 29523 bench.py --gpus 8 --steps 20 --warmup 3 --workload nnfm 2> $O/r2_f_bench_n8_nnfm.err | grep '^{' > $O/r2_f_bench_n8_nnfm.json
timeout 300 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --steps 10 --warmup 3 --workload inference2m 2> $O/r2_f_bench_n8_inf.err | grep '^{' > $O/r2_f_bench_n8_inf.json
timeout 300 python bench.py --steps 10 --warmup 3 --workload inference2m 2> $O/r2_f_bench_n1_inf.err | grep '^{' > $O/r2_f_bench_n1_inf.json
timeout 300 $TR --nproc-per-node 8 --master-port 29525 tools/profile_step_multi.py > $O/r2_f_profile_n8.txt 2> $O/r2_f_profile_n8.err
timeout 300 python tools/profile_step_multi.py > $O/r2_f_profile_n1.txt 2> $O/r2_f_profile_n1.err
for f in $O/r2_f_bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1], d["n_gpus"], d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"))
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
head -30 $O/r2_f_profile_n8.txt
