#!/bin/bash
# round 2, call 68: the default bench line with its extras on the final commit
python bench.py > gpurun_out/r4o_bench_default.json 2>/dev/null
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r4o_bench_default.json") if l.startswith("{")][-1])
print(d["ms_per_step"], d["roofline"]["frac"], d["parity_check"]["ok"], d["clocks"], d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
for k,v in d["extra"].items(): print("   ",k, round(v.get("us_per_launch", v.get("us_per_block",0)),2), v.get("us_per_block_dense_reference_loop",""))
PY
