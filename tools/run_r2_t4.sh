#!/usr/bin/env bash
# end of round 2: reference arm on the two other training workloads
cd "$(dirname "$0")/.."
O=gpurun_out
for w in hypernerf8 nnfm; do
  python bench.py --impl reference --workload $w --steps 6 --warmup 3 > $O/r2_t_ref_$w.json 2> $O/r2_t_ref_$w.err || tail -3 $O/r2_t_ref_$w.err
  python -c "
import json
d=json.loads(open('gpurun_out/r2_t_ref_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
