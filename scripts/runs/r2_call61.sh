#!/bin/bash
# round 2, call 61: worker cap from 32-token tiles up; the aligned policy inside the device-counted (DYN) GEMM:
# MoE + GEMM tests, mid-batch MoE block bench, kernel bench of the few-unit shapes at 24-32 tokens
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_moe.py tests/test_gpu_gemm.py tests/test_zz_no_alloc.py -m gpu -x -q > $O/r4i_tests.log 2>&1; tail -2 $O/r4i_tests.log
python scripts/kernel_bench.py --m 24,32 --shapes 8192x1280,8192x1024,14336x512,8192x2560 > $O/r4i_kb.jsonl 2>&1; cut -c1-120 $O/r4i_kb.jsonl
for epr in 1 8; do
  timeout 400 python scripts/moe_bench.py --tokens 64,128,256,512,1024 --layers 2 --experts-per-rank $epr > $O/r4i_moe_mid_epr$epr.jsonl 2>&1
  cut -c1-260 $O/r4i_moe_mid_epr$epr.jsonl
done
