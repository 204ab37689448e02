#!/bin/bash
# round 2, call 63 (8 GPUs): the final state at TP 8 and TP 4 (aligned grids on the TP shards, fused kernel)
set -x
O=gpurun_out
for n in 8 4; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2954$n \
  bench.py --gpus $n --steps 20 --warmup 5 > $O/r4k_bench_tp$n.json 2> $O/r4k_bench_tp$n.err
python - <<PY
import json
d=json.loads([l for l in open("$O/r4k_bench_tp$n.json") if l.startswith("{")][-1])
print($n, d["ms_per_step"], d["value"], d["parity_check"]["ok"], d["e2e"]["value"], d["clocks"])
PY
done
