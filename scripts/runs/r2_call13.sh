#!/bin/bash
# round 2, call 13: decode kernel with the collector warp (bulk-copy fetch of stream-K partials)
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_activation.py tests/test_gpu_allreduce.py tests/test_moe.py -m gpu -q -x --timeout 240 > $O/r2m_gputests.log 2>&1
tail -5 $O/r2m_gputests.log
timeout 300 python scripts/kernel_bench.py --m 1,8,16 --shapes 4096x4096,4096x12288,4096x22016,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192 --iters 10 >> $O/r2m_kb_decode.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 8,16 --shapes 8192x10240,8192x8192,8192x57344,28672x8192 --iters 5 --min-mb 300 >> $O/r2m_kb_decode.jsonl 2>&1
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 > $O/r2m_trace_decode_4096.log 2>&1
timeout 120 python scripts/trace_decode.py --m 16 --k 8192 --n 1280 > $O/r2m_trace_decode_8192x1280.log 2>&1
