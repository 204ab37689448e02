#!/bin/bash
# round 2, call 19: full GPU suite; decode kernel with 4 / 2 consumer groups; contract bench lines;
# MoE benches at Mixtral shapes; ncu of the prefill kernel
set -x
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 300 > $O/r2s_gputests.log 2>&1
tail -3 $O/r2s_gputests.log
for v in tune diag0g4 diag0g2; do
timeout 300 python scripts/kernel_bench.py --lib libb200_w4a16_$v.so --kernel decode --m 8,16 --shapes 4096x4096,4096x22016,11008x4096,8192x1280,3584x8192,8192x57344 --iters 10 > $O/r2s_kb_decode_$v.jsonl 2>&1
done
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2s_bench_default.json 2> $O/r2s_bench_default.err
tail -c 400 $O/r2s_bench_default.json
timeout 300 python bench.py --steps 20 --warmup 3 --workload gemm1 --no-extras > $O/r2s_bench_gemm1.json 2> $O/r2s_bench_gemm1.err
timeout 300 python bench.py --steps 5 --warmup 3 --workload prefill --no-extras > $O/r2s_bench_prefill.json 2> $O/r2s_bench_prefill.err
timeout 300 python bench.py --steps 10 --warmup 3 --model llama2-7b --no-extras > $O/r2s_bench_7b.json 2> $O/r2s_bench_7b.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2s_bench_ref.json 2> $O/r2s_bench_ref.err
timeout 300 python scripts/moe_bench.py --tokens 1,8,16,64,256 --experts-per-rank 1 --layers 16 > $O/r2s_moe_bench_tp8.jsonl 2>&1
timeout 300 python scripts/moe_bench.py --tokens 8,64 --experts-per-rank 8 --layers 3 > $O/r2s_moe_bench_tp1.jsonl 2>&1
timeout 300 python scripts/moe_bench.py --prefill --tokens 256,1024,4096 --experts-per-rank 1 --layers 4 --iters 5 > $O/r2s_moe_prefill_tp8.jsonl 2>&1
# ncu: one 13B prefill GEMM (M = 8192, 5120 -> 15360), plain run first
CMD="python scripts/kernel_bench.py --once --m 8192 --shapes 5120x15360"
$CMD > $O/r2s_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:w4a16_gemm_kernel -c 1 -o $O/r2s_ncu_prefill_m8192_5120x15360 $CMD > $O/r2s_ncu_full.log 2>&1
ls -la $O | tail -20
