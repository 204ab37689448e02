#!/bin/bash
# round 2, call 14: where do the 2.2 us of the stream-K fix-up go? stamps inside it; GPU-scope polls
set -x
O=gpurun_out
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 > $O/r2n_trace_decode_4096.log 2>&1
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 --flags 4 > $O/r2n_trace_decode_4096_gpuscope.log 2>&1
B200_W4A16_FLAGS=4 timeout 300 python scripts/kernel_bench.py --kernel decode --m 8,16 --shapes 4096x4096,8192x1280,1024x8192,4096x12288 --iters 10 >> $O/r2n_kb_decode_gpuscope.jsonl 2>&1
timeout 600 python -m pytest tests/test_moe.py tests/test_gpu_activation.py -m gpu -q -x --timeout 240 > $O/r2n_gputests_moe.log 2>&1
tail -3 $O/r2n_gputests_moe.log
