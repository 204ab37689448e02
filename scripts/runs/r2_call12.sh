#!/bin/bash
# round 2, call 12: why is the split-ring kernel slower at 64 tokens on big shapes? traces
set -x
O=gpurun_out
B200_W4A16_PAIR=1 timeout 120 python scripts/trace_gemm.py --kernel tcgen05 --m 64 --k 8192 --n 57344 --ctas 0,73 > $O/r2l_trace_m64_pair1.log 2>&1
B200_W4A16_PAIR=2 timeout 120 python scripts/trace_gemm.py --kernel tcgen05 --m 64 --k 8192 --n 57344 --ctas 0,73 > $O/r2l_trace_m64_pair2.log 2>&1
B200_W4A16_PAIR=2 timeout 120 python scripts/trace_gemm.py --kernel tcgen05 --m 32 --k 8192 --n 57344 --ctas 0 > $O/r2l_trace_m32_pair2.log 2>&1
B200_W4A16_FLAGS=1 B200_W4A16_PAIR=2 timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 64,2048 --shapes 5120x15360,8192x57344 --iters 5 --min-mb 100 > $O/r2l_kb_noevict.jsonl 2>&1
