#!/bin/bash
# round 2, call 60: unit-aligned grids as the product default: full GPU suite, bench lines
set -x
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r4h_gpu_tests.log 2>&1; tail -3 $O/r4h_gpu_tests.log
python bench.py > $O/r4h_bench_default.json 2>/dev/null
python bench.py --workload gemm1 --no-extras --no-cpu-baseline > $O/r4h_bench_gemm1.json 2>/dev/null
python bench.py --workload prefill --no-extras --no-cpu-baseline > $O/r4h_bench_prefill.json 2>/dev/null
python bench.py --model llama2-7b --no-extras --no-cpu-baseline > $O/r4h_bench_7b.json 2>/dev/null
python - <<PY
import json
for f in ("default","gemm1","prefill","7b"):
    d=json.loads([l for l in open("$O/r4h_bench_%s.json"%f) if l.startswith("{")][-1])
    print(f, d["ms_per_step"], d["roofline"]["frac"], d["parity_check"]["ok"], d["clocks"])
    if "extra" in d:
        for k,v in d["extra"].items(): print("   ",k, round(v.get("us_per_launch", v.get("us_per_block",0)),2), v.get("us_per_block_dense_reference_loop",""))
PY
