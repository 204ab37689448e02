#!/bin/bash
set -x
timeout 900 python -m pytest tests -m gpu -q -x --timeout 180 -k "not llama13b_prefill" > gpurun_out/r2g_gputests.log 2>&1
tail -3 gpurun_out/r2g_gputests.log
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 8192 --n 57344 --ctas 0,73 > gpurun_out/r2g_trace_tcdec_57344.log 2>&1
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 4096 --n 4096 --ctas 0,73 > gpurun_out/r2g_trace_tcdec_4096.log 2>&1
timeout 120 python scripts/trace_gemm.py --m 2048 --k 5120 --n 15360 --ctas 0,73 > gpurun_out/r2g_trace_prefill.log 2>&1
timeout 300 python scripts/kernel_bench.py --kernel tcdecode --m 8,16 --shapes 4096x4096,4096x22016,11008x4096,8192x1280,1024x8192,8192x57344 --iters 10 >> gpurun_out/r2g_kb_tcdecode.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 32,64 --shapes 4096x4096,4096x22016,1024x8192,8192x57344 --iters 10 >> gpurun_out/r2g_kb_tcdecode.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 128,256,512,2048,8192 --shapes 5120x15360,13824x5120 --iters 5 --min-mb 100 >> gpurun_out/r2g_kb_prefill.jsonl 2>&1
