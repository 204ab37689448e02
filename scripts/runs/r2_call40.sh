#!/bin/bash
# round 2, call 40: router kernel as a CTA per token; where the MoE block's time goes (ncu launch list)
set -x
O=gpurun_out
python -m pytest tests/test_moe.py -m gpu -x -q > $O/r3n_moe_tests.log 2>&1; tail -5 $O/r3n_moe_tests.log
for epr in 8 1; do
  timeout 300 python scripts/moe_bench.py --tokens 1,8,16 --layers 2 --experts-per-rank $epr > $O/r3n_moe_bench_epr$epr.jsonl 2>&1
  cat $O/r3n_moe_bench_epr$epr.jsonl
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r3n_launches_moe_t1.csv python scripts/moe_bench.py --tokens 1 --layers 2 --experts-per-rank 8 --iters 1 > $O/r3n_ncu_moe.log 2>&1
