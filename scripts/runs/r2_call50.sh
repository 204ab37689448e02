#!/bin/bash
# round 2, call 50 (2 GPUs): final confirmation of the multi-rank tests and the TP 2 bench line on the
# committed state (decode kernel now a <MT, GROUPED> template; the non-grouped path must be unchanged)
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_allreduce.py tests/test_gpu_norm.py -m gpu -x -q > $O/r3x_mgpu_tests.log 2>&1; tail -3 $O/r3x_mgpu_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
  bench.py --gpus 2 --steps 20 --warmup 5 > $O/r3x_bench_tp2.json 2> $O/r3x_bench_tp2.err
python - <<PY
import json
d=json.loads([l for l in open("$O/r3x_bench_tp2.json") if l.startswith("{")][-1])
print(d["ms_per_step"], d["value"], d["parity_check"]["ok"], d["e2e"]["value"], d["clocks"])
PY
