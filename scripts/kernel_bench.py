#!/usr/bin/env python
"""Per-shape timing of the W4A16 GEMM (CUDA events around a CUDA graph of back-to-back
launches that rotate through > 2x L2 of distinct weights).  Not the contract bench (that is
bench.py); this is the tuning harness.  Usage:
  python scripts/kernel_bench.py [--m 16] [--shapes 4096x4096,4096x22016] [--iters 20] [--once]
"""
import argparse
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")

G = 128


def bytes_alg(M, K, N):
    return K * N // 2 + (K // G) * N * 2 + (K // G) * N // 2 + M * K * 2 + M * N * 2


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", default="16")
    ap.add_argument("--shapes", default="4096x4096,4096x12288,4096x22016,11008x4096")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--once", action="store_true", help="single eager launch per shape (for ncu)")
    ap.add_argument("--min-mb", type=int, default=512)
    ap.add_argument("--kernel", default=None, choices=["decode", "tcgen05"],
                    help="pin the kernel variant (loads the tuning build: make -C vllm-quick_b200/csrc tune)")
    ap.add_argument("--lib", default=None, help="load lib/<file> instead (tuning builds: make variant / diag)")
    args = ap.parse_args()
    if args.lib:
        if args.kernel:
            os.environ["B200_W4A16_KERNEL"] = args.kernel
        pkg._lib.use_library(args.lib)
    elif args.kernel:
        os.environ["B200_W4A16_KERNEL"] = args.kernel
        pkg._lib.use_library("libb200_w4a16_tune.so")
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    peak = 6541.5
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    for shape in args.shapes.split(","):
        K, N = (int(v) for v in shape.split("x"))
        for M in (int(v) for v in args.m.split(",")):
            wbytes = K * N // 2
            nsets = 1 if args.once else max(2, (args.min_mb << 20) // wbytes + 1)
            sets = []
            for _ in range(nsets):
                qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=g)
                qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device=dev, generator=g)
                s = (torch.rand((K // G, N), device=dev, generator=g) * 0.019 + 0.001).half()
                sets.append((qw, s, qz))
            x = torch.randn((M, K), device=dev, generator=g).half()
            if args.once:
                y = pkg.ops.quick_gemm(x, *sets[0], 8)
                torch.cuda.synchronize()
                print(f"once M={M} K={K} N={N} ok {float(y.float().abs().mean()):.4f}")
                continue
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for (qw, s, qz) in sets[:2]:
                    pkg.ops.quick_gemm(x, qw, s, qz, 8)
                st.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=st):
                    for (qw, s, qz) in sets:
                        pkg.ops.quick_gemm(x, qw, s, qz, 8)
                for _ in range(3):
                    graph.replay()
                st.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(args.iters):
                    graph.replay()
                e1.record(st)
                st.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (args.iters * nsets)
            gbs = bytes_alg(M, K, N) / us / 1e3
            tf = 2.0 * M * K * N / us / 1e6
            plan = pkg._lib.gemm_plan(M, K, N)
            print(json.dumps({"M": M, "K": K, "N": N, "us": round(us, 3), "GBps": round(gbs, 1),
                              "hbm_frac": round(gbs / peak, 4), "TFLOPs": round(tf, 2),
                              "grid": plan["grid"], "tok": plan["tok_tile"], "sets": nsets,
                              "kernel": args.kernel or plan["variant"]}))
            del sets, graph


if __name__ == "__main__":
    main()
