for d in 0 2 4 6 8 16; do echo "DIRECT=$d"; PENROZ_GEMM_DIRECT=$d timeout 200 python tools/probe_gemm.py 2>&1 | sed -n 3,3p; done
