#!/bin/bash
# round 2, call 1: goldens for norm/activation from the reference kernels + baseline numbers
set -x
mkdir -p gpurun_out/golden
python tests/golden/make_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1
python bench.py --model llama2-70b --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_bench_70b_tp1_base.json 2> gpurun_out/r2a_bench_70b_tp1_base.err
python scripts/kernel_bench.py --m 1,8,16,32,64 --shapes 8192x1280,1024x8192,8192x7168,3584x8192 --iters 10 > gpurun_out/r2a_kb_tp8_base.jsonl 2>&1
python scripts/kernel_bench.py --m 16,32,64 --shapes 8192x10240,8192x8192,8192x57344,28672x8192 --iters 5 --min-mb 300 > gpurun_out/r2a_kb_tp1_base.jsonl 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
