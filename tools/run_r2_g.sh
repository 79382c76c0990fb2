#!/usr/bin/env bash
cd "$(dirname "$0")/.."
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_g_tests.log 2>&1; echo "pytest rc=$?" >> $O/r2_g_tests.log
tail -25 $O/r2_g_tests.log
timeout 300 python tools/profile_step_multi.py > $O/r2_g_profile_n1.txt 2> $O/r2_g_profile_n1.err; echo "profile rc=$?"
head -50 $O/r2_g_profile_n1.txt
