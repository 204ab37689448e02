#!/bin/bash
# round 2, call 2: compact goldens + GPU test suite on the LL fix-up / packet / tcgen05-decode kernels
set -x
mkdir -p gpurun_out/golden
python tests/golden/make_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q --timeout 180 -k "not llama13b_prefill" > gpurun_out/r2b_gputests.log 2>&1
tail -5 gpurun_out/r2b_gputests.log
for k in decode tcdecode; do
timeout 300 python scripts/kernel_bench.py --kernel $k --m 8,16 --shapes 4096x4096,4096x12288,4096x22016,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192,8192x57344 --iters 10 >> gpurun_out/r2b_kb_$k.jsonl 2>&1
done
timeout 300 python scripts/kernel_bench.py --m 32,64 --shapes 4096x4096,4096x12288,4096x22016,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192,8192x57344 --iters 10 >> gpurun_out/r2b_kb_tcdecode.jsonl 2>&1
