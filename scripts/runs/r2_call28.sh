#!/bin/bash
# round 2, call 28: CTA pairs for 32- / 64-token tiles (tuning build, forced): parity + timing
set -x
O=gpurun_out
cat > /tmp/check_mid.py <<'PY'
import importlib, os, sys
import numpy as np, torch
ROOT = os.environ.get("GRAFT_REPO_ROOT", "/root/repo")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("vllm-quick_b200")
pkg._lib.use_library("libb200_w4a16_tune.so")
from gpu_util import sampled_case
from oracle import oracle
oracle.build()
worst = 0.0
for (M, K, N, G) in [(64, 8192, 7168, 128), (33, 4096, 22016, 128), (48, 8192, 57344, 128), (20, 4096, 11 * 128 + 64, 128), (64, 2048, 4096, 64), (30, 8192, 2560, 32)]:
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, G, seed=M + N, n_words=8, edge_words=4)
    r = pkg.ops.quick_repack(qw)
    y = pkg.ops.quick_gemm(x, r, s, qz, 8); y2 = pkg.ops.quick_gemm(x, r, s, qz, 8)
    torch.cuda.synchronize()
    rel = float((np.abs(y[:, cols].float().cpu().numpy() - Yf) / np.maximum(np.abs(Yf), 1.0)).max())
    worst = max(worst, rel)
    print((M, K, N, G), pkg._lib.gemm_plan(M, K, N, G), "max_rel", round(rel, 5), "det", bool(torch.equal(y, y2)), flush=True)
print("worst", worst, "OK" if worst <= 1e-2 else "FAIL")
PY
B200_W4A16_TWO=1 timeout 300 python /tmp/check_mid.py > $O/r3b_check_mid_two.log 2>&1
tail -3 $O/r3b_check_mid_two.log
for rep in 1 2; do
for t in 0 1; do
B200_W4A16_TWO=$t timeout 300 python scripts/kernel_bench.py --kernel tcgen05 --m 32,64 --shapes 8192x57344,4096x22016,8192x7168,28672x8192,11008x4096 --iters 10 --min-mb 300 >> $O/r3b_kb_mid_two$t.jsonl 2>&1
done
done
