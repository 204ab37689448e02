#!/bin/bash
# round 2, call 55: A/B of the division-free range arithmetic (old = HEAD library in a worktree under build/)
set -x
O=gpurun_out
SH=4096x4096,4096x12288,11008x4096,1024x8192,8192x1280,8192x57344
for i in 1 2; do
  for t in old new; do
    if [ $t = old ]; then S=build/old_tree/scripts/kernel_bench.py; else S=scripts/kernel_bench.py; fi
    python $S --m 1,16,64,256 --shapes $SH > $O/r4c_kb_${t}_$i.jsonl 2>&1
  done
done
python - <<PY
import json
def load(f):
    return {(r["M"],r["K"],r["N"]): r["us"] for r in (json.loads(l) for l in open(f) if l.startswith("{"))}
o1,o2,n1,n2=(load("$O/r4c_kb_%s.jsonl"%s) for s in ("old_1","old_2","new_1","new_2"))
for k in sorted(o1, key=lambda k:(k[1],k[2],k[0])):
    print(k, "old", o1[k], o2[k], "new", n1[k], n2[k], "ratio %.3f" % (min(n1[k],n2[k])/min(o1[k],o2[k])))
PY
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -x -q > $O/r4c_gemm_tests.log 2>&1; tail -2 $O/r4c_gemm_tests.log
