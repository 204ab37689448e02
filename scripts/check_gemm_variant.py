#!/usr/bin/env python
"""Tuning aid: parity of the tcgen05 kernel variant selected by the B200_W4A16_* switches of the
tuning build (PAIR, TWO) against the CPU oracle on sampled channels, plus determinism.
  B200_W4A16_TWO=1 python scripts/check_gemm_variant.py"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("vllm-quick_b200")
pkg._lib.use_library("libb200_w4a16_tune.so")
from gpu_util import sampled_case  # noqa: E402
from oracle import oracle  # noqa: E402

oracle.build()
SHAPES = [(256, 1024, 1024, 128), (2048, 5120, 15360, 128), (512, 5120, 5120, 128), (1024, 13824, 5120, 128),
          (140, 8192, 19 * 128, 128), (300, 5120, 37 * 128, 128), (130, 16384, 11 * 128 + 64, 128),
          (520, 2048, 4096, 64), (200, 8192, 2560, 32), (8192, 5120, 27648, 128)]
worst = 0.0
for (M, K, N, G) in SHAPES:
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, G, seed=M + N, n_words=8, edge_words=4)
    r = pkg.ops.quick_repack(qw)
    y = pkg.ops.quick_gemm(x, r, s, qz, 8)
    y2 = pkg.ops.quick_gemm(x, r, s, qz, 8)
    torch.cuda.synchronize()
    d = (y[:, cols].float().cpu().numpy() - Yf)
    import numpy as np
    rel = float((np.abs(d) / np.maximum(np.abs(Yf), 1.0)).max())
    worst = max(worst, rel)
    print((M, K, N, G), pkg._lib.gemm_plan(M, K, N, G), "max_rel", round(rel, 5), "deterministic", bool(torch.equal(y, y2)),
          "finite", bool(torch.isfinite(y).all()), flush=True)
print("worst", worst, "OK" if worst <= 1e-2 else "FAIL")
