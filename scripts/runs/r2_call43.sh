#!/bin/bash
# round 2, call 43: dynamic row counts with fewer work items than CTAs (fix), MoE tests, full suite
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_moe.py -m gpu -x -q > $O/r3q_moe_tests.log 2>&1; tail -5 $O/r3q_moe_tests.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/r3q_gpu_tests.log 2>&1; tail -3 $O/r3q_gpu_tests.log
timeout 300 python scripts/moe_bench.py --tokens 96,128,192 --layers 2 --experts-per-rank 1 --routed-on-device-min 17 > $O/r3q_moe_mid_epr1.jsonl 2>&1; cat $O/r3q_moe_mid_epr1.jsonl
