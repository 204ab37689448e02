#!/bin/bash
# round 2, call 53: where a small mid-batch launch spends its time (tcgen05 kernel, 4096 -> 4096 at 64 / 256 tokens)
set -x
O=gpurun_out
timeout 200 python scripts/trace_gemm.py --m 64 --k 4096 --n 4096 > $O/r4a_trace_m64_4096x4096.txt 2>&1
timeout 200 python scripts/trace_gemm.py --m 256 --k 4096 --n 4096 > $O/r4a_trace_m256_4096x4096.txt 2>&1
timeout 200 python scripts/trace_gemm.py --m 512 --k 5120 --n 5120 > $O/r4a_trace_m512_5120x5120.txt 2>&1
