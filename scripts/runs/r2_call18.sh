#!/bin/bash
# round 2, call 18 (8 GPUs): multi-rank tests at 8 ranks, all-reduce sweep at TP 8 / 4, Llama-2-70B
# decode bench at TP 8 (default = fused GEMM + all-reduce; packets) and TP 4
set -x
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_allreduce.py tests/test_gpu_norm.py -m gpu -q -x --timeout 300 -k "multi_gpu" > $O/r2r_gputests_8gpu.log 2>&1
tail -3 $O/r2r_gputests_8gpu.log
timeout 600 $TR --nproc-per-node 8 --master-port 29621 bench.py --gpus 8 --steps 10 --warmup 3 > $O/r2r_bench_tp8.json 2> $O/r2r_bench_tp8.err
tail -c 600 $O/r2r_bench_tp8.json
timeout 600 $TR --nproc-per-node 8 --master-port 29622 bench.py --gpus 8 --steps 10 --warmup 3 --packet-allreduce --no-cpu-baseline > $O/r2r_bench_tp8_packets.json 2> $O/r2r_bench_tp8_packets.err
timeout 600 $TR --nproc-per-node 8 --master-port 29623 bench.py --gpus 8 --steps 10 --warmup 3 --nccl-allreduce --no-cpu-baseline > $O/r2r_bench_tp8_nccl.json 2> $O/r2r_bench_tp8_nccl.err
timeout 600 $TR --nproc-per-node 4 --master-port 29624 bench.py --gpus 4 --steps 10 --warmup 3 > $O/r2r_bench_tp4.json 2> $O/r2r_bench_tp4.err
timeout 600 $TR --nproc-per-node 4 --master-port 29625 bench.py --gpus 4 --steps 10 --warmup 3 --packet-allreduce --no-cpu-baseline > $O/r2r_bench_tp4_packets.json 2> $O/r2r_bench_tp4_packets.err
timeout 600 $TR --nproc-per-node 8 --master-port 29626 scripts/allreduce_bench.py > $O/r2r_allreduce_bench_tp8.jsonl 2> $O/r2r_allreduce_bench_tp8.err
timeout 600 $TR --nproc-per-node 4 --master-port 29627 scripts/allreduce_bench.py > $O/r2r_allreduce_bench_tp4.jsonl 2> $O/r2r_allreduce_bench_tp4.err
for f in $O/r2r_bench_*.json; do python -c "
import json,sys
for l in open('$f'):
    if l.startswith('{'):
        d=json.loads(l); print('$f', d['ms_per_step'], d['e2e']['ms_per_step'], d['allreduce'], d['parity_check']['ok'])
"; done
