#!/bin/bash
# round 2, call 33: decode kernel inner loop: HMMA halves interleaved across token blocks (+ compiler-
# schedulable mma asm), 16 instead of 12 ring stages at M = 16
set -x
O=gpurun_out
for rep in 1 2; do
for v in diagbase tune diags16; do
timeout 300 python scripts/kernel_bench.py --lib libb200_w4a16_$v.so --kernel decode --m 8,16 --shapes 4096x4096,4096x22016,11008x4096,3584x8192,8192x57344,28672x8192 --iters 10 --min-mb 300 >> $O/r3g_kb_decode_$v.jsonl 2>&1
done
done
timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_activation.py tests/test_gpu_allreduce.py tests/test_moe.py -m gpu -q -x --timeout 300 > $O/r3g_gputests.log 2>&1
tail -2 $O/r3g_gputests.log
