#!/bin/bash
# round 2, call 41: programmatic launches for the MoE helper kernels, router without read-back,
# grouped path also for one local expert: tests + block bench 8 / 2 / 1 local experts
set -x
O=gpurun_out
python -m pytest tests/test_moe.py -m gpu -x -q > $O/r3o_moe_tests.log 2>&1; tail -5 $O/r3o_moe_tests.log
for epr in 8 2 1; do
  timeout 300 python scripts/moe_bench.py --tokens 1,4,8,16 --layers 2 --experts-per-rank $epr > $O/r3o_moe_bench_epr$epr.jsonl 2>&1
  cat $O/r3o_moe_bench_epr$epr.jsonl
done
timeout 300 python scripts/moe_bench.py --prefill --tokens 64,256,1024 --layers 2 --experts-per-rank 1 > $O/r3o_moe_prefill_epr1.jsonl 2>&1; cat $O/r3o_moe_prefill_epr1.jsonl
python -m pytest tests -m gpu -x -q > $O/r3o_gpu_tests.log 2>&1; tail -3 $O/r3o_gpu_tests.log
