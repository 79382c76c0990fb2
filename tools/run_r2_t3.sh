#!/usr/bin/env bash
# end of round 2, eight GPUs: config 2 and config 3 at N = 8 (defaults: 8 SMs left to the all-reduce under the MLP backward,
# NCCL held to 8 CTAs in the GAUSSIAN state)
cd "$(dirname "$0")/.."
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
timeout 300 $TR --master-port 29541 bench.py --gpus 8 --steps 30 --warmup 5 2> $O/r2_t_bench_n8.err | grep '^{' > $O/r2_t_bench_n8.json
timeout 300 $TR --master-port 29542 bench.py --gpus 8 --steps 10 --warmup 3 --workload hypernerf8 2> $O/r2_t_bench_n8_hypernerf8.err | grep '^{' > $O/r2_t_bench_n8_hypernerf8.json
python - <<'P'
import json
for f in ("n8","n8_hypernerf8"):
    try:
        d=json.loads(open("gpurun_out/r2_t_bench_%s.json"%f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"].get("reserve_sms"))
    except Exception as e: print(f, "ERR", e)
P
tail -2 $O/r2_t_bench_n8.err
