#!/bin/bash
# round 2, call 62 (2 GPUs): multi-rank tests and the TP 2 bench line with the aligned grids (the fused decode
# kernel's grid changes on the TP shards)
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_allreduce.py tests/test_gpu_norm.py -m gpu -x -q > $O/r4j_mgpu_tests.log 2>&1; tail -2 $O/r4j_mgpu_tests.log
for extra in "" "--fused-gemm-allreduce"; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
  bench.py --gpus 2 --steps 20 --warmup 5 $extra > $O/r4j_bench_tp2$extra.json 2> $O/r4j_bench_tp2.err
python - <<PY
import json
d=json.loads([l for l in open("$O/r4j_bench_tp2$extra.json") if l.startswith("{")][-1])
print("$extra", d["ms_per_step"], d["value"], d["parity_check"]["ok"], d["e2e"]["value"], d["clocks"])
PY
done
