#!/bin/bash
# round 2, call 29: CTA pairs with one channel tile per CTA (double-buffered accumulators) vs two
set -x
O=gpurun_out
B200_W4A16_PAIR=1 B200_W4A16_TWO=1 timeout 300 python scripts/check_gemm_variant.py > $O/r3c_check_pair1_two.log 2>&1
tail -3 $O/r3c_check_pair1_two.log
for rep in 1 2; do
B200_W4A16_TWO=1 timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 512,1024,2048,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 10 --min-mb 300 >> $O/r3c_kb_pair2_two.jsonl 2>&1
B200_W4A16_PAIR=1 B200_W4A16_TWO=1 timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 512,1024,2048,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 10 --min-mb 300 >> $O/r3c_kb_pair1_two.jsonl 2>&1
done
timeout 300 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x --timeout 300 -k "cta_pair or channel_tile" > $O/r3c_gputests.log 2>&1
tail -2 $O/r3c_gputests.log
