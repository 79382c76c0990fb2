#!/usr/bin/env bash
# compute-sanitizer over tools/sanitize_target.py: memcheck, racecheck (shared-memory hazards: the hand-rolled
# mbarrier / CTA-pair protocol of csrc/mlp_fused.cu, the aliased staging of radix_downsweep_staged_kernel) and
# synccheck.  Logs go to gpurun_out/; copy the summaries into profiles/.
cd "$(dirname "$0")/.."
O=${1:-gpurun_out}
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_target.py > $O/sanitize_$tool.log 2>&1
  echo "$tool rc=$?" >> $O/sanitize_$tool.log
  tail -4 $O/sanitize_$tool.log
done
