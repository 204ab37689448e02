#!/bin/bash
# round 2, call 25: CTA pairs + one commit per tile as the product default: full suite, final
# prefill table, bench lines, ncu of the pair kernel
set -x
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 300 > $O/r2y_gputests.log 2>&1
tail -3 $O/r2y_gputests.log
timeout 400 python scripts/kernel_bench.py --m 128,256,512,1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 5 --min-mb 100 > $O/r2y_kb_prefill.jsonl 2>&1
timeout 300 python bench.py --steps 5 --warmup 3 --workload prefill --no-extras > $O/r2y_bench_prefill.json 2> $O/r2y_bench_prefill.err
timeout 300 python bench.py --steps 3 --warmup 3 --workload prefill --batch 8192 --no-extras --no-cpu-baseline > $O/r2y_bench_prefill_m8192.json 2> $O/r2y_bench_prefill_m8192.err
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2y_bench_default.json 2> $O/r2y_bench_default.err
CMD="python scripts/kernel_bench.py --once --m 8192 --shapes 5120x15360"
$CMD > $O/r2y_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:w4a16_gemm_kernel -c 1 -o $O/r2y_ncu_prefill_pair_m8192_5120x15360 $CMD > $O/r2y_ncu_full.log 2>&1
