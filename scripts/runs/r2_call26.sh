#!/bin/bash
# round 2, call 26 (2 GPUs): --layer-ops bench mode at TP 1 / 2, TP2 default with the final kernels,
# multi-rank tests once more
set -x
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python bench.py --steps 10 --warmup 3 --layer-ops --no-extras > $O/r2z_bench_layerops_tp1.json 2> $O/r2z_bench_layerops_tp1.err
tail -c 400 $O/r2z_bench_layerops_tp1.json; tail -3 $O/r2z_bench_layerops_tp1.err
timeout 600 $TR --nproc-per-node 2 --master-port 29631 bench.py --gpus 2 --steps 10 --warmup 3 --layer-ops > $O/r2z_bench_layerops_tp2.json 2> $O/r2z_bench_layerops_tp2.err
tail -c 400 $O/r2z_bench_layerops_tp2.json; tail -3 $O/r2z_bench_layerops_tp2.err
timeout 600 $TR --nproc-per-node 2 --master-port 29632 bench.py --gpus 2 --steps 10 --warmup 3 > $O/r2z_bench_tp2.json 2> $O/r2z_bench_tp2.err
timeout 600 python -m pytest tests/test_gpu_allreduce.py tests/test_gpu_norm.py -m gpu -q -x --timeout 300 -k "multi_gpu" > $O/r2z_gputests_2gpu.log 2>&1
tail -2 $O/r2z_gputests_2gpu.log
