#!/usr/bin/env python
"""Time the REFERENCE's own kernels (oracle/_ref, compiled unmodified for sm_100a) on the same
shapes and protocol as scripts/kernel_bench.py: awq_gemm (split_k = 8, kernel + sum(0)) and, for
M >= 256, awq_dequantize + torch.matmul (the reference's own branch, awq.py:154-159).
Same-box "kernel to beat" numbers for profiles/ -- not part of the product."""
import argparse
import glob
import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = 128


def bytes_alg(M, K, N):
    return K * N // 2 + (K // G) * N * 2 + (K // G) * N // 2 + M * K * 2 + M * N * 2


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", default="1,16,64,256")
    ap.add_argument("--shapes", default="4096x4096,4096x12288,4096x22016,11008x4096")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--min-mb", type=int, default=512)
    args = ap.parse_args()
    path = glob.glob(os.path.join(ROOT, "oracle", "_ref", "ref_awq*.so"))[0]
    spec = importlib.util.spec_from_file_location("ref_awq", path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    for shape in args.shapes.split(","):
        K, N = (int(v) for v in shape.split("x"))
        nsets = max(2, (args.min_mb << 20) // (K * N // 2) + 1)
        sets = []
        for _ in range(nsets):
            sets.append((torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=g),
                         (torch.rand((K // G, N), device=dev, generator=g) * 0.019 + 0.001).half(),
                         torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device=dev, generator=g)))
        for M in (int(v) for v in args.m.split(",")):
            x = torch.randn((M, K), device=dev, generator=g).half()

            def run(qw, s, qz):
                if M >= 256:
                    return torch.matmul(x, ref.awq_dequantize(qw, s, qz, 0, 0, 0))
                return ref.awq_gemm(x, qw, s, qz, 8)

            for t in sets[:2]:
                run(*t)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                for t in sets:
                    run(*t)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (args.iters * nsets)
            print(json.dumps({"impl": "reference kernel (sm_100a rebuild)", "M": M, "K": K, "N": N,
                              "us": round(us, 3), "GBps": round(bytes_alg(M, K, N) / us / 1e3, 1),
                              "TFLOPs": round(2.0 * M * K * N / us / 1e6, 2),
                              "branch": "dequantize+matmul" if M >= 256 else "awq_gemm split_k=8"}))
        del sets


if __name__ == "__main__":
    main()
