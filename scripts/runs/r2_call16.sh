#!/bin/bash
# round 2, call 16 (2 GPUs): multi-rank tests incl. the two-shot all-reduce, all-reduce size sweep,
# MoE tests, TP2 bench
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_allreduce.py tests/test_gpu_norm.py tests/test_moe.py -m gpu -q -x --timeout 300 > $O/r2p_gputests_2gpu.log 2>&1
tail -5 $O/r2p_gputests_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 scripts/allreduce_bench.py > $O/r2p_allreduce_bench_tp2.jsonl 2> $O/r2p_allreduce_bench_tp2.err
tail -3 $O/r2p_allreduce_bench_tp2.jsonl
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 > $O/r2p_bench_tp2.json 2> $O/r2p_bench_tp2.err
tail -c 1500 $O/r2p_bench_tp2.json
