#!/bin/bash
# round 2, call 56: cap on the CTAs sharing one output unit at 64- / 128-token tiles (few units, long K)
set -x
O=gpurun_out
SH=8192x1280,8192x1024,8192x2560,14336x512,4096x4096
for cap in 0 4 6 8 12 16; do
  B200_W4A16_SPLITCAP=$cap python scripts/kernel_bench.py --lib libb200_w4a16_tune.so --m 48,64,128,256,512 --shapes $SH > $O/r4d_kb_cap$cap.jsonl 2>&1
done
python - <<PY
import json
def load(f):
    return {(r["M"],r["K"],r["N"]): (r["us"], r.get("grid")) for r in (json.loads(l) for l in open(f) if l.startswith("{"))}
caps=(0,4,6,8,12,16)
d={c: load("$O/r4d_kb_cap%d.jsonl"%c) for c in caps}
for k in sorted(d[0], key=lambda k:(k[1],k[2],k[0])):
    print(k, " ".join("cap%d %.1f(g%s)"%(c, d[c][k][0], d[c][k][1]) for c in caps))
PY
