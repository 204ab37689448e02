#!/bin/bash
set -x
timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_allreduce.py tests/test_gpu_activation.py tests/test_moe.py -m gpu -q -x --timeout 180 -k "not llama13b_prefill" > gpurun_out/r2d_gputests.log 2>&1
tail -3 gpurun_out/r2d_gputests.log
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 > gpurun_out/r2d_trace_decode_4096.log 2>&1
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 8192 --n 57344 --ctas 0,73 > gpurun_out/r2d_trace_tcdec_57344.log 2>&1
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 4096 --n 4096 --ctas 0,73 > gpurun_out/r2d_trace_tcdec_4096.log 2>&1
for k in decode tcdecode; do
timeout 300 python scripts/kernel_bench.py --kernel $k --m 8,16 --shapes 4096x4096,4096x22016,11008x4096,8192x1280,1024x8192,8192x57344 --iters 10 >> gpurun_out/r2d_kb_$k.jsonl 2>&1
done
timeout 300 python scripts/kernel_bench.py --m 32,64 --shapes 4096x4096,4096x22016,1024x8192,8192x57344 --iters 10 >> gpurun_out/r2d_kb_tcdecode.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 512,2048 --shapes 5120x15360,13824x5120 --iters 5 --min-mb 100 >> gpurun_out/r2d_kb_prefill.jsonl 2>&1
