#!/bin/bash
# round 2, call 27 (8 GPUs): final kernels at TP 8 / 4 (default and --layer-ops), 8-rank tests
set -x
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29641 bench.py --gpus 8 --steps 10 --warmup 3 > $O/r3a_bench_tp8.json 2> $O/r3a_bench_tp8.err
timeout 600 $TR --nproc-per-node 4 --master-port 29642 bench.py --gpus 4 --steps 10 --warmup 3 > $O/r3a_bench_tp4.json 2> $O/r3a_bench_tp4.err
timeout 600 $TR --nproc-per-node 8 --master-port 29643 bench.py --gpus 8 --steps 10 --warmup 3 --layer-ops --no-cpu-baseline > $O/r3a_bench_layerops_tp8.json 2> $O/r3a_bench_layerops_tp8.err
timeout 600 $TR --nproc-per-node 8 --master-port 29644 bench.py --gpus 8 --steps 10 --warmup 3 --batch 8 --no-cpu-baseline > $O/r3a_bench_tp8_m8.json 2> $O/r3a_bench_tp8_m8.err
timeout 600 python -m pytest tests/test_gpu_allreduce.py tests/test_gpu_norm.py -m gpu -q -x --timeout 300 -k "multi_gpu" > $O/r3a_gputests_8gpu.log 2>&1
tail -2 $O/r3a_gputests_8gpu.log
for f in $O/r3a_bench_*.json; do python -c "
import json
for l in open('$f'):
    if l.startswith('{'):
        d=json.loads(l); print('$f', d['ms_per_step'], d['e2e']['ms_per_step'], d['allreduce'], d['parity_check']['ok'], d['parity_check']['max_rel_all_ranks'])
"; done
