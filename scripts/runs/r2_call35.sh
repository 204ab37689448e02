#!/bin/bash
# round 2, call 35: final policy (CTA pairs: one tile per CTA for short K, two otherwise): full suite,
# final prefill table (10 iterations, 300 MiB of weight sets), contract bench lines
set -x
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 300 > $O/r3i_gputests.log 2>&1
tail -3 $O/r3i_gputests.log
timeout 600 python scripts/kernel_bench.py --m 128,256,512,1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 10 --min-mb 300 > $O/r3i_kb_prefill.jsonl 2>&1
timeout 300 python bench.py --steps 5 --warmup 3 --workload prefill --no-extras > $O/r3i_bench_prefill.json 2> $O/r3i_bench_prefill.err
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r3i_bench_default.json 2> $O/r3i_bench_default.err
timeout 300 python bench.py --steps 20 --warmup 3 --workload gemm1 --no-extras > $O/r3i_bench_gemm1.json 2> $O/r3i_bench_gemm1.err
