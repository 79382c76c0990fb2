for res in 0 1; do GEMM_RES=$res python tools/bench_gemm.py 300000 20 2>&1 | grep -E "gemm fwd|gemm dgrad" | tr '\n' ' '; echo " resident=$res"; done
