#!/usr/bin/env bash
cd "$(dirname "$0")/.."
cp trase_b200/libtrase_b200.so /tmp/lib8.so
cp trase_b200/libtrase_b200_epi12.so trase_b200/libtrase_b200.so
python tools/fused_mlp_stalls.py 300000 0 2>&1 | tail -8
cp /tmp/lib8.so trase_b200/libtrase_b200.so
