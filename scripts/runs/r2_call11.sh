#!/bin/bash
# round 2, call 11: new tcgen05 kernel (split rings, channel-tile pairs, cooperative fix-up),
# decode kernel with front-loaded fix-up polls
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_activation.py tests/test_moe.py tests/test_gpu_checkpoint.py tests/test_gpu_aux.py tests/test_golden.py -m gpu -q -x --timeout 240 > $O/r2k_gputests.log 2>&1
tail -5 $O/r2k_gputests.log
timeout 300 python scripts/kernel_bench.py --m 1,8,16 --shapes 4096x4096,4096x12288,4096x22016,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192 --iters 10 >> $O/r2k_kb_decode.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 24,32,48,64 --shapes 4096x4096,4096x22016,11008x4096,1024x8192,8192x7168,8192x57344 --iters 10 >> $O/r2k_kb_mid.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 128,256,512,1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 5 --min-mb 100 >> $O/r2k_kb_prefill.jsonl 2>&1
B200_W4A16_PAIR=1 timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 64,512,2048,8192 --shapes 5120x15360,13824x5120,8192x57344 --iters 5 --min-mb 100 >> $O/r2k_kb_pair1.jsonl 2>&1
timeout 120 python scripts/trace_gemm.py --m 2048 --k 5120 --n 15360 --ctas 0,73 > $O/r2k_trace_prefill.log 2>&1
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 > $O/r2k_trace_decode_4096.log 2>&1
ls -la $O
