#!/bin/bash
# round 2, call 45: are the config-5 GEMM extras (M = 64 / 256) box-dependent?  two more default runs
set -x
O=gpurun_out
nvidia-smi --query-gpu=power.limit,power.draw,clocks.sm,temperature.gpu --format=csv
for i in 1 2; do
  python bench.py --no-cpu-baseline > $O/r3s_bench_default_$i.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("$O/r3s_bench_default_$i.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["clocks"])
for k,v in d["extra"].items(): print("   ",k, round(v.get("us_per_launch", v.get("us_per_block",0)),2))
PY
done
nvidia-smi --query-gpu=power.limit,power.draw,clocks.sm,temperature.gpu --format=csv
