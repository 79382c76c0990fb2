#!/usr/bin/env bash
# two GPUs: NCCL parity test, the no-alloc sort on a second GPU of one process, N=2 bench lines
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests/test_gpu_multi.py tests/test_zz_no_alloc.py -m gpu -x -q -s > $O/r2_c_multi_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_c_multi_tests.log
tail -30 $O/r2_c_multi_tests.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus 2 --steps 30 --warmup 5 > $O/r2_c_bench_n2.json 2> $O/r2_c_bench_n2.err; echo "n2 rc=$?"
$TR bench.py --gpus 2 --steps 30 --warmup 5 --full-arena > $O/r2_c_bench_n2_full.json 2> $O/r2_c_bench_n2_full.err; echo "n2 full rc=$?"
$TR bench.py --gpus 2 --steps 10 --warmup 3 --workload hypernerf8 > $O/r2_c_bench_n2_hyper.json 2> $O/r2_c_bench_n2_hyper.err; echo "n2 hyper rc=$?"
$TR bench.py --gpus 2 --steps 20 --warmup 3 --workload nnfm > $O/r2_c_bench_n2_nnfm.json 2> $O/r2_c_bench_n2_nnfm.err; echo "n2 nnfm rc=$?"
for f in $O/r2_c_bench_*.json; do echo "== $f"; head -c 400 $f; echo; done
