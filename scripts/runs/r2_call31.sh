#!/bin/bash
# round 2, call 31: which grouping at 128-token tiles? (PAIR, TWO) in {1,2} x {0,1}, by token count
set -x
O=gpurun_out
for rep in 1 2; do
for cfg in "2 0" "2 1" "1 1" "1 0"; do
set -- $cfg
B200_W4A16_PAIR=$1 B200_W4A16_TWO=$2 timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 128,256,512,4096 --shapes 4096x28672,14336x4096,5120x15360,5120x5120 --iters 10 --min-mb 300 >> $O/r3e_kb_p$1_t$2.jsonl 2>&1
done
done
