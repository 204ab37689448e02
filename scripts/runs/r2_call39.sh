#!/bin/bash
# round 2, call 39: router kernel (gate GEMM + softmax + top-k + renormalise + counts in one
# launch): MoE tests, block bench at 8 / 4 / 2 / 1 local experts
set -x
O=gpurun_out
python -m pytest tests/test_moe.py -m gpu -x -q > $O/r3m_moe_tests.log 2>&1; tail -5 $O/r3m_moe_tests.log
for epr in 8 4 2 1; do
  timeout 300 python scripts/moe_bench.py --tokens 1,4,8,16 --layers 2 --experts-per-rank $epr > $O/r3m_moe_bench_epr$epr.jsonl 2>&1
  cat $O/r3m_moe_bench_epr$epr.jsonl
done
timeout 300 python scripts/moe_bench.py --prefill --tokens 64,256,1024 --layers 2 --experts-per-rank 1 > $O/r3m_moe_prefill_epr1.jsonl 2>&1; cat $O/r3m_moe_prefill_epr1.jsonl
