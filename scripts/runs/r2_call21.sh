#!/bin/bash
# round 2, call 21: CTA-pair (cta_group::2) variant of the tcgen05 kernel: parity, then timing
set -x
O=gpurun_out
B200_W4A16_TWO=1 timeout 300 python scripts/check_gemm_variant.py > $O/r2u_check_two.log 2>&1
tail -4 $O/r2u_check_two.log
B200_W4A16_TWO=0 timeout 300 python scripts/check_gemm_variant.py > $O/r2u_check_one.log 2>&1
tail -2 $O/r2u_check_one.log
for t in 0 1; do
B200_W4A16_TWO=$t timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 512,1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 5 --min-mb 100 > $O/r2u_kb_prefill_two$t.jsonl 2>&1
done
B200_W4A16_TWO=1 timeout 120 python scripts/trace_gemm.py --kernel tcgen05 --m 2048 --k 5120 --n 15360 --ctas 0,1 > $O/r2u_trace_prefill_two.log 2>&1
