#!/bin/bash
# round 2, call 46: A/B of the Mixtral expert GEMM shapes between the r3i library (worktree at
# 542f0ba under build/old_tree) and the current one, alternating, same box
set -x
O=gpurun_out
for i in 1 2; do
  for t in old new; do
    if [ $t = old ]; then S=build/old_tree/scripts/kernel_bench.py; else S=scripts/kernel_bench.py; fi
    python $S --m 8,64,256 --shapes 4096x28672,14336x4096 > $O/r3t_kb_${t}_$i.jsonl 2>&1
    python - <<PY
import json
for l in open("$O/r3t_kb_${t}_$i.jsonl"):
    if l.startswith("{"):
        r=json.loads(l); print("$t $i", r["M"], r["K"], r["N"], r["us"], r.get("tok"), r.get("grid"), r.get("kernel"))
PY
  done
done
cp build/old_tree/MEASURED_PEAKS.json /dev/null 2>&1
python build/old_tree/bench.py --no-cpu-baseline > $O/r3t_bench_old.json 2>/dev/null
python - <<PY
import json
d=json.loads(open("$O/r3t_bench_old.json").read().strip().splitlines()[-1])
print("old bench", d["ms_per_step"])
for k,v in d["extra"].items(): print("   ",k, round(v.get("us_per_launch", 0),2))
PY
