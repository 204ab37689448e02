#!/bin/bash
# round 2, call 65: grouped expert launch on all SMs (the single-expert plan may be a unit-aligned, smaller grid)
set -x
O=gpurun_out
timeout 300 python -m pytest tests/test_moe.py -m gpu -x -q > $O/r4m_moe_tests.log 2>&1; tail -2 $O/r4m_moe_tests.log
timeout 200 python scripts/moe_bench.py --tokens 1,8,16 --layers 2 --experts-per-rank 8 > $O/r4m_moe_bench_epr8.jsonl 2>&1; cut -c1-300 $O/r4m_moe_bench_epr8.jsonl
