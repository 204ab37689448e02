#!/bin/bash
# round 2, call 44: driver-equivalent run on the committed state: smoke, bench default (extras now
# with the MoE block of config 5), gemm1, prefill, reference arm
set -x
O=gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > $O/r3r_smoke.log 2>&1; tail -1 $O/r3r_smoke.log
python bench.py > $O/r3r_bench_default.json 2> $O/r3r_bench_default.err; tail -c 1500 $O/r3r_bench_default.json
python bench.py --workload gemm1 --no-extras > $O/r3r_bench_gemm1.json 2>/dev/null; cut -c1-300 $O/r3r_bench_gemm1.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/r3r_bench_reference.json 2>/dev/null; cut -c1-400 $O/r3r_bench_reference.json
