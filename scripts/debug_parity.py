#!/usr/bin/env python
"""Locate GEMM parity failures: compares quick_gemm with dequantize + fp32 matmul on the GPU
and prints which (channel tile, token tile) blocks are wrong.  Tuning / debugging aid."""
import argparse
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")


def check(M, K, N, G=128, seed=0, reps=3):
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(seed)
    qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=g)
    qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device=dev, generator=g)
    s = (torch.rand((K // G, N), device=dev, generator=g) * 0.019 + 0.001).half()
    x = torch.randn((M, K), device=dev, generator=g).half()
    r = pkg.ops.quick_repack(qw)
    W = pkg.ops.quick_dequantize(r, s, qz)
    ref = torch.empty((M, N), dtype=torch.float32, device=dev)
    for n0 in range(0, N, 4096):
        ref[:, n0:n0 + 4096] = x.float() @ W[:, n0:n0 + 4096].float()
    plan = pkg._lib.gemm_plan(M, K, N)
    tok = plan["tok_tile"]
    bad_total = 0
    for rep in range(reps):
        y = pkg.ops.quick_gemm(x, r, s, qz, 8)
        torch.cuda.synchronize()
        err = (y.float() - ref).abs() / ref.abs().clamp(min=1.0)
        bad = err > 1e-2
        nb = int(bad.sum().item())
        bad_total += nb
        if nb:
            mt = (M + tok - 1) // tok
            nt = (N + 127) // 128
            pad = torch.zeros((mt * tok, nt * 128), dtype=torch.bool, device=dev)
            pad[:M, :N] = bad
            blk = pad.view(mt, tok, nt, 128).permute(0, 2, 1, 3).reshape(mt, nt, -1).sum(-1)
            idx = torch.nonzero(blk)
            print(f"  rep {rep}: {nb} bad elements in {idx.shape[0]} blocks; max err {err.max().item():.3f}")
            KB = K // 128
            T = nt * mt * KB
            P = plan["grid"]
            for (a, b) in idx[:12].tolist():
                o = b * mt + a
                cnt = int(blk[a, b].item())
                # which CTAs own this output tile's K range
                lo, hi = o * KB, o * KB + KB - 1
                p0 = ((lo + 1) * P - 1) // T
                p1 = ((hi + 1) * P - 1) // T
                print(f"    block mt={a} nt={b} o={o} bad={cnt}/{tok * 128} owners {p0}..{p1} "
                      f"range_begin(p0)={p0 * T // P} lo={lo}")
    print(f"M={M} K={K} N={N} plan={plan} -> {'OK' if bad_total == 0 else 'BAD'}")
    return bad_total == 0


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="512x5120x27648,512x5120x15360,128x5120x27648,512x4096x4096,"
                                        "1024x4096x4096,2048x5120x5120,256x5120x27648,64x5120x27648")
    args = ap.parse_args()
    ok = True
    for sh in args.shapes.split(","):
        M, K, N = (int(v) for v in sh.split("x"))
        ok &= check(M, K, N)
    sys.exit(0 if ok else 1)
