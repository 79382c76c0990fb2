#!/usr/bin/env bash
# eight GPUs: SMs left to the overlapped all-reduce / NCCL CTA count sweep at N=8
cd "$(dirname "$0")/.."
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
run() { # name reserve ctas
  if [ -n "$3" ]; then export NCCL_MAX_CTAS=$3; else unset NCCL_MAX_CTAS; fi
  timeout 300 $TR --master-port $4 bench.py --gpus 8 --steps 30 --warmup 5 --reserve-sms $2 2> $O/r2_l_$1.err | grep '^{' > $O/r2_l_$1.json
  python - "$O/r2_l_$1.json" "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[2], d["value"], d["ms_per_step"], d["e2e"]["value"], {k:v["ms"] for k,v in d["stages"].items() if k.startswith("mlp")})
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
run res16 16 "" 29531
run res8_cta8 8 8 29532
run res16_cta16 16 16 29533
run res0_cta8 0 8 29534
