#!/bin/bash
# round 2, call 24: careful A/B of CTA pairs (both with one commit per tile), interleaved repeats
set -x
O=gpurun_out
for rep in 1 2 3; do
for t in 0 1; do
B200_W4A16_TWO=$t timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 1024,2048,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 10 --min-mb 300 >> $O/r2x_kb_ab_two$t.jsonl 2>&1
done
done
