#!/bin/bash
# round 2, call 47 (2 GPUs): the MoE block across real ranks (expert split, NVLink all-reduce at the end)
set -x
O=gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  scripts/moe_bench.py --tokens 1,8,64,256 --layers 2 > $O/r3u_moe_tp2.jsonl 2> $O/r3u_moe_tp2.err
cat $O/r3u_moe_tp2.jsonl; tail -5 $O/r3u_moe_tp2.err
