#!/bin/bash
# round 2, call 48 (8 GPUs): BASELINE config 5 as stated: Mixtral-8x7B expert block, TP = 8, mixed batch sizes
set -x
O=gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
  scripts/moe_bench.py --tokens 1,4,8,16,64,128,256,1024 --layers 2 > $O/r3v_moe_tp8.jsonl 2> $O/r3v_moe_tp8.err
cat $O/r3v_moe_tp8.jsonl; tail -5 $O/r3v_moe_tp8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 \
  scripts/moe_bench.py --tokens 1,8,64,256 --layers 2 > $O/r3v_moe_tp4.jsonl 2> $O/r3v_moe_tp4.err
cat $O/r3v_moe_tp4.jsonl
