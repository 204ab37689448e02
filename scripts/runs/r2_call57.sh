#!/bin/bash
# round 2, call 57: timeline of two slow mid-batch shapes (long K, few channel tiles)
set -x
O=gpurun_out
B200_W4A16_SPLITCAP=0 timeout 200 python scripts/trace_gemm.py --m 256 --k 8192 --n 2560 --ctas 0,1,2,3,4,73 > $O/r4e_trace_m256_8192x2560.txt 2>&1
B200_W4A16_SPLITCAP=0 timeout 200 python scripts/trace_gemm.py --m 64 --k 8192 --n 1280 --ctas 0,1,7,14,15,73 > $O/r4e_trace_m64_8192x1280.txt 2>&1
