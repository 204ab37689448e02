#!/bin/bash
# round 2, call 38: grouped expert launch (one decode-kernel launch over the local experts with
# tokens): MoE tests, block bench grouped vs one gated launch per expert vs dense loop
set -x
O=gpurun_out
python -m pytest tests/test_moe.py -m gpu -x -q > $O/r3l_moe_tests.log 2>&1; tail -5 $O/r3l_moe_tests.log
for epr in 8 4 2; do
  timeout 300 python scripts/moe_bench.py --tokens 1,4,8,16 --layers 2 --experts-per-rank $epr > $O/r3l_moe_bench_epr$epr.jsonl 2>&1
  cat $O/r3l_moe_bench_epr$epr.jsonl
done
python -m pytest tests -m gpu -x -q > $O/r3l_gpu_tests.log 2>&1; tail -3 $O/r3l_gpu_tests.log
