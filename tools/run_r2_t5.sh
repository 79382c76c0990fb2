#!/usr/bin/env bash
# end of round 2, four GPUs: config 2
cd "$(dirname "$0")/.."
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 4"
timeout 300 $TR --master-port 29551 bench.py --gpus 4 --steps 30 --warmup 5 2> $O/r2_t_bench_n4.err | grep '^{' > $O/r2_t_bench_n4.json
python -c "
import json
d=json.loads(open('gpurun_out/r2_t_bench_n4.json').read().strip().splitlines()[-1]); print('n4', d['value'], d['ms_per_step'], d['e2e']['value'])"
