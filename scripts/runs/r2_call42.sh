#!/bin/bash
# round 2, call 42: GEMM with the row count read on the device (DYN instantiations of the tcgen05
# kernel), routed MoE path without host synchronisation: tests, graph-mode block bench at
# 32 .. 1024 tokens against the dense loop, prefill bench sanity (static variants untouched)
set -x
O=gpurun_out
python -m pytest tests/test_moe.py -m gpu -x -q > $O/r3p_moe_tests.log 2>&1; tail -5 $O/r3p_moe_tests.log
for epr in 1 8; do
  timeout 400 python scripts/moe_bench.py --tokens 32,64,128,256,512,1024 --layers 2 --experts-per-rank $epr --routed-on-device-min 17 > $O/r3p_moe_mid_epr$epr.jsonl 2>&1
  cat $O/r3p_moe_mid_epr$epr.jsonl
done
python bench.py --workload prefill --no-cpu-baseline --no-extras > $O/r3p_bench_prefill.json 2> $O/r3p_bench_prefill.err; cut -c1-400 $O/r3p_bench_prefill.json
