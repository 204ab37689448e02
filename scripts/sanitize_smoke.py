#!/usr/bin/env python
"""Small launches of every kernel family, each checked against the oracle: a quick end-to-end
exercise after a kernel change (and the input for compute-sanitizer where that tool is available --
it is closed on the round-2 GPU pool).  Decode kernel (split
tiles, fused activation, gated), tcgen05 kernel (single tiles, channel-tile pairs, CTA pairs --
forced through the tuning build's switches), dequantize / repack, norm, activation, MoE gather /
scatter, all-reduce kernels in one-GPU emulation.  Results are checked against the oracle where
cheap.  Usage: python scripts/sanitize_smoke.py   (B200_W4A16_TWO=1 B200_W4A16_PAIR=2 forces CTA
pairs on the small shapes through the tuning build)"""
import ctypes
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("vllm-quick_b200")
if os.environ.get("B200_W4A16_TWO") or os.environ.get("B200_W4A16_PAIR"):
    pkg._lib.use_library("libb200_w4a16_tune.so")
from gpu_util import sampled_case  # noqa: E402
from oracle import oracle  # noqa: E402

oracle.build()
L = pkg._lib.lib()


def gemm(M, K, N, G=128):
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, G, seed=M + N, n_words=4, edge_words=2)
    y = pkg.ops.quick_gemm(x, pkg.ops.quick_repack(qw), s, qz, 8)
    torch.cuda.synchronize()
    rel = float((np.abs(y[:, cols].float().cpu().numpy() - Yf) / np.maximum(np.abs(Yf), 1.0)).max())
    print("gemm", (M, K, N, G), pkg._lib.gemm_plan(M, K, N, G), "max_rel", round(rel, 5), flush=True)
    assert rel <= 1e-2


for shape in [(16, 1024, 512), (3, 512, 192, 64), (40, 1024, 1024), (128, 1024, 8192), (130, 512, 640, 32)]:
    gemm(*shape)
# fused activation + gate
x, qw, s, qz, _, _ = sampled_case(oracle, 8, 512, 1024, 128, seed=1, n_words=2)
method = pkg.QuickConfig(4, 128, True).get_linear_method()
w = {"qweight": torch.nn.Parameter(qw, requires_grad=False), "qzeros": torch.nn.Parameter(qz, requires_grad=False),
     "scales": torch.nn.Parameter(s, requires_grad=False), "qweight_layout": "awq"}
flag = torch.zeros(1, dtype=torch.int32, device="cuda")
with pkg.ops.expert_gate(flag):
    z = method.apply_weights_silu_and_mul(w, x)
flag.fill_(3)
with pkg.ops.expert_gate(flag):
    a = method.apply_weights_silu_and_mul(w, x)
torch.cuda.synchronize()
assert not z.any() and a.shape == (8, 512)
# dequantize, norm, activation, MoE plumbing
W = pkg.ops.awq_dequantize(qw, s, qz)
h = torch.randn(5, 1024, device="cuda").half()
r = torch.randn(5, 1024, device="cuda").half()
pkg.ops.fused_add_rms_norm(h, r, torch.ones(1024, device="cuda").half(), 1e-6)
o = torch.empty(5, 512, device="cuda").half()
pkg.ops.silu_and_mul(o, h)
idx = torch.tensor([4, 0, 2], dtype=torch.int32, device="cuda")
g = pkg.ops.moe_gather(h, idx)
pkg.ops.moe_scatter_add(h, g, idx, torch.ones(3, device="cuda"))
# all-reduce kernels, one-GPU emulation (no cross-GPU waits)
world, n = 4, 4096
max_bytes = 2 * n
regions = [torch.zeros(16384 + 2 * max_bytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
ptrs = (ctypes.c_void_p * world)(*[t.data_ptr() for t in regions])
parts = [torch.randn(n, device="cuda").half() for _ in range(world)]
outs = [torch.empty_like(p) for p in parts]
st = torch.cuda.current_stream().cuda_stream
for q in range(world):
    regions[q][16384 + max_bytes: 16384 + max_bytes + 2 * n] = parts[q].view(torch.uint8)
for q in range(world):
    assert L.b200_w4a16_allreduce_twoshot(ptrs, q, world, parts[q].data_ptr(), outs[q].data_ptr(), n, max_bytes, 2, st) == 0
for q in range(world):
    assert L.b200_w4a16_allreduce_twoshot(ptrs, q, world, parts[q].data_ptr(), outs[q].data_ptr(), n, max_bytes, 3, st) == 0
torch.cuda.synchronize()
acc = sum(p.float() for p in parts).half()
assert torch.equal(outs[0], acc) and torch.equal(outs[3], acc)
print("sanitize smoke ok")
