#!/bin/bash
# round 2, call 20: final single-GPU tables -- decode kernel with four consumer groups, tcgen05
# kernel with weights kept in L2 across token tiles; full suite; contract bench lines; launch list
set -x
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 300 > $O/r2t_gputests.log 2>&1
tail -3 $O/r2t_gputests.log
timeout 300 python scripts/kernel_bench.py --m 1,8,16 --shapes 4096x4096,4096x12288,4096x22016,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192 --iters 10 > $O/r2t_kb_decode.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 8,16 --shapes 8192x10240,8192x8192,8192x57344,28672x8192 --iters 5 --min-mb 300 >> $O/r2t_kb_decode.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 24,32,48,64 --shapes 4096x4096,4096x22016,11008x4096,1024x8192,8192x7168,8192x57344 --iters 10 > $O/r2t_kb_mid.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 128,256,512,1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 5 --min-mb 100 > $O/r2t_kb_prefill.jsonl 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2t_bench_default.json 2> $O/r2t_bench_default.err
tail -c 300 $O/r2t_bench_default.json
timeout 300 python bench.py --steps 20 --warmup 3 --workload gemm1 --no-extras > $O/r2t_bench_gemm1.json 2> $O/r2t_bench_gemm1.err
timeout 300 python bench.py --steps 5 --warmup 3 --workload prefill --no-extras > $O/r2t_bench_prefill.json 2> $O/r2t_bench_prefill.err
CMD="python bench.py --steps 2 --warmup 3 --layers 2 --no-cpu-baseline --no-extras"
$CMD > $O/r2t_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2t_launches_bench_short.csv $CMD > $O/r2t_ncu_launch.log 2>&1
