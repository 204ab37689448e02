#!/bin/bash
# round 2, call 17: ring-depth variants of the tcgen05 kernel at 32 / 64 tokens; MoE tests
set -x
O=gpurun_out
for v in tune vara varb varc; do
for pr in 1 2; do
B200_W4A16_PAIR=$pr timeout 200 python scripts/kernel_bench.py --lib libb200_w4a16_$v.so --kernel tcgen05 --m 32,64 --shapes 8192x57344,4096x22016,8192x7168 --iters 5 --min-mb 300 > $O/r2q_kb_${v}_pair$pr.jsonl 2>&1
done
done
timeout 600 python -m pytest tests/test_moe.py -m gpu -q --timeout 300 > $O/r2q_gputests_moe.log 2>&1
tail -3 $O/r2q_gputests_moe.log
