#!/bin/bash
# round 2, call 51: eager prefill MoE block: host-routed vs routed-on-device vs dense loop (policy for >= 512 tokens)
set -x
O=gpurun_out
for epr in 1 8; do
  timeout 400 python scripts/moe_bench.py --prefill --tokens 256,512,1024,2048,4096,8192 --layers 2 --experts-per-rank $epr --iters 10 > $O/r3y_moe_prefill_epr$epr.jsonl 2>&1
  cat $O/r3y_moe_prefill_epr$epr.jsonl
done
