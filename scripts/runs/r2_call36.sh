#!/bin/bash
# round 2, call 36: super-blocked unit order at 8192 / 16384 tokens (tuning build, B200_W4A16_SB)
set -x
O=gpurun_out
B200_W4A16_SB=8 timeout 300 python scripts/check_gemm_variant.py > $O/r3j_check_sb8.log 2>&1
tail -3 $O/r3j_check_sb8.log
for rep in 1 2; do
for sb in 0 8 16 32; do
B200_W4A16_SB=$sb timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 4096,8192,16384 --shapes 5120x15360,5120x27648,13824x5120 --iters 5 --min-mb 300 >> $O/r3j_kb_sb$sb.jsonl 2>&1
done
done
