#!/bin/bash
# round 2, call 34: token-tile pairs (one dequantized weight tile feeds two 128-token MMAs) on CTA
# pairs: parity, then A/B against the product policy
set -x
O=gpurun_out
B200_W4A16_TP=2 timeout 300 python scripts/check_gemm_variant.py > $O/r3h_check_tp2.log 2>&1
tail -4 $O/r3h_check_tp2.log
for rep in 1 2; do
timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 10 --min-mb 300 >> $O/r3h_kb_policy.jsonl 2>&1
B200_W4A16_TP=2 timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 10 --min-mb 300 >> $O/r3h_kb_tp2.jsonl 2>&1
done
B200_W4A16_TP=2 timeout 120 python scripts/trace_gemm.py --kernel tcgen05 --m 2048 --k 5120 --n 15360 --ctas 0 > $O/r3h_trace_tp2.log 2>&1
