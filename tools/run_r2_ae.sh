#!/usr/bin/env bash
# two GPUs, end of round 2: NCCL parity test, the no-alloc sort on a second GPU of one process, N=2 bench lines
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests/test_gpu_multi.py tests/test_zz_no_alloc.py -m gpu -x -q -s > $O/r2_ae_multi_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_ae_multi_tests.log
tail -22 $O/r2_ae_multi_tests.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus 2 --steps 30 --warmup 5 > $O/r2_ae_bench_n2.json 2> $O/r2_ae_bench_n2.err; echo "n2 rc=$?"
$TR bench.py --gpus 2 --steps 10 --warmup 3 --workload hypernerf8 > $O/r2_ae_bench_n2_hyper.json 2> $O/r2_ae_bench_n2_hyper.err; echo "n2 hyper rc=$?"
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/r2_ae_bench_n1.json 2>/dev/null
python - <<'P'
import json
for f in ("n1","n2","n2_hyper"):
    try:
        d=json.loads(open("gpurun_out/r2_ae_bench_%s.json"%f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"].get("reserve_sms"))
    except Exception as e: print(f, "ERR", e)
P
