for pf in 0 1 2 3 4 6; do GEMM_PF=$pf python tools/bench_gemm.py 300000 20 2>&1 | grep -E "gemm fwd|gemm dgrad" | tr '\n' ' '; echo " pf=$pf"; done
