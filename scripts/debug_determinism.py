#!/usr/bin/env python
"""Debug aid: repeat one GEMM back to back and report where the outputs differ."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")
if len(sys.argv) > 4:
    os.environ["B200_W4A16_KERNEL"] = sys.argv[4]
    pkg._lib.use_library("libb200_w4a16_tune.so")
M, K, N = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (16, 4096, 4096)
g = torch.Generator(device="cuda").manual_seed(0)
qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device="cuda", generator=g)
qz = torch.randint(-2**31, 2**31 - 1, (K // 128, N // 8), dtype=torch.int32, device="cuda", generator=g)
s = (torch.rand((K // 128, N), device="cuda", generator=g) * 0.019 + 0.001).half()
x = torch.randn((M, K), device="cuda", generator=g).half()
r = pkg.ops.quick_repack(qw)
ref = pkg.ops.quick_gemm(x, r, s, qz, 8)
torch.cuda.synchronize()
for rep in range(4):
    ys = [pkg.ops.quick_gemm(x, r, s, qz, 8) for _ in range(6)]
    torch.cuda.synchronize()
    for i, y in enumerate(ys):
        d = (y.float() - ref.float()).abs()
        bad = (d > 0).nonzero()
        if bad.numel():
            cols = bad[:, 1].unique().cpu().numpy()
            rows = bad[:, 0].unique().cpu().numpy()
            print(f"rep {rep} launch {i}: {bad.shape[0]} diffs, max {float(d.max()):.4f}, rows {rows[:16]}, "
                  f"tiles {np.unique(cols // 128)[:24]}, cols-in-tile {np.unique(cols % 128)[:40]}")
    ws = list(pkg.ops._workspaces.values())[0]
    print("rep", rep, "workspace idle:", bool((ws == 0xFF).all()), "plan", pkg._lib.gemm_plan(M, K, N))
