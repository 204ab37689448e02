#!/bin/bash
set -x
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 240 > gpurun_out/r2i_gputests.log 2>&1
tail -3 gpurun_out/r2i_gputests.log
for k in tcdecode tcgen05; do
timeout 300 python scripts/kernel_bench.py --kernel $k --m 24,32,48,64 --shapes 4096x4096,4096x22016,11008x4096,1024x8192,8192x7168,8192x57344 --iters 10 >> gpurun_out/r2i_kb_$k.jsonl 2>&1
done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2i_bench_default.log 2>&1
tail -c 3000 gpurun_out/r2i_bench_default.log
timeout 300 python bench.py --steps 20 --warmup 3 --workload gemm1 --no-extras > gpurun_out/r2i_bench_gemm1.log 2>&1
tail -c 1500 gpurun_out/r2i_bench_gemm1.log
timeout 300 python bench.py --steps 5 --warmup 3 --workload prefill --no-extras > gpurun_out/r2i_bench_prefill.log 2>&1
tail -c 1500 gpurun_out/r2i_bench_prefill.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2i_bench_ref.log 2>&1
tail -c 1500 gpurun_out/r2i_bench_ref.log
