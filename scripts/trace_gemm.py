#!/usr/bin/env python
"""Pipeline timeline of one GEMM launch (needs `make -C vllm-quick_b200/csrc trace`).  Prints, for a few CTAs, per-tile clock offsets of every pipeline role.
Usage: python scripts/trace_gemm.py --m 16 --k 4096 --n 22016"""
import argparse
import ctypes
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")
pkg._lib.use_library("libb200_w4a16_trace.so")      # make -C vllm-quick_b200/csrc trace


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=16)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--n", type=int, default=22016)
    ap.add_argument("--ctas", default="0,73,147")
    ap.add_argument("--warm", type=int, default=4)
    ap.add_argument("--kernel", default=None, choices=["decode", "tcgen05"])
    args = ap.parse_args()
    if args.kernel:
        os.environ["B200_W4A16_KERNEL"] = args.kernel      # honoured by tuning / trace builds only
    M, K, N, G = args.m, args.k, args.n, 128
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    sets = []
    for _ in range(args.warm + 1):
        qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=g)
        qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device=dev, generator=g)
        s = (torch.rand((K // G, N), device=dev, generator=g) * 0.019 + 0.001).half()
        sets.append((qw, s, qz))
    x = torch.randn((M, K), device=dev, generator=g).half()
    L = pkg._lib.lib()
    L.b200_w4a16_debug_set_trace.argtypes = [ctypes.c_void_p]
    L.b200_w4a16_debug_set_trace.restype = None
    nl = len(sets)
    traces = torch.zeros((nl, 148, 8, 64), dtype=torch.int32, device=dev)
    for (qw, s, qz) in sets[:2]:
        pkg.ops.quick_gemm(x, qw, s, qz, 8)
    torch.cuda.synchronize()
    # flush L2
    junk = torch.empty(512 << 20, dtype=torch.uint8, device=dev); junk.zero_(); torch.cuda.synchronize()
    # back-to-back launches, one trace buffer each (the pointer is a launch-time parameter)
    for i, (qw, s, qz) in enumerate(sets):
        L.b200_w4a16_debug_set_trace(traces[i].data_ptr())
        pkg.ops.quick_gemm(x, qw, s, qz, 8)
    torch.cuda.synchronize()
    L.b200_w4a16_debug_set_trace(None)
    ta = traces.cpu().numpy().astype(np.uint32)
    t = ta[-1]
    grid = pkg._lib.gemm_plan(M, K, N)["grid"]
    # absolute timeline (globaltimer ns): kernel entry / setup done / epilogue done per launch
    g0 = np.int64(ta[0, :grid, 7, 2].min())
    for i in range(nl):
        ent = ta[i, :grid, 7, 2].astype(np.int64) - g0
        setup = ta[i, :grid, 7, 3].astype(np.int64) - g0
        done = ta[i, :grid, 7, 4].astype(np.int64) - g0
        cyc = (ta[i, :grid, 7, 6].astype(np.int64) - ta[i, :grid, 7, 5].astype(np.int64))
        ns = done - ent
        print(f"launch {i}: entry ns min {ent.min()} med {int(np.median(ent))} max {ent.max()} | "
              f"setup-entry med {int(np.median(setup - ent))} | done min {done.min()} med "
              f"{int(np.median(done))} max {done.max()} | CTA life ns med {int(np.median(ns))} "
              f"cycles med {int(np.median(cyc))} => {np.median(cyc) / max(np.median(ns), 1):.3f} GHz")
    plan = pkg._lib.gemm_plan(M, K, N)
    print("plan", plan)
    names = ["issue", "landed", "dq_start", "dq_done", "mma_start", "mma_issued"]
    for cta in (int(v) for v in args.ctas.split(",")):
        if cta >= plan["grid"]:
            continue
        base = t[cta, 7, 0]
        rel = lambda v: int(np.int64(v) - np.int64(base)) if v else -1
        print(f"== CTA {cta}: exit at {rel(t[cta,7,1])} cycles after setup; prologue (cycles before setup): entry "
              f"{rel(t[cta,7,5])} walk {rel(t[cta,7,8])} barriers {rel(t[cta,7,9])} cursor {rel(t[cta,7,10])} "
              f"ring issued {rel(t[cta,7,11])} tmem {rel(t[cta,7,12])} | dependency resolved {rel(t[cta,7,13])}")
        print("  it " + " ".join(f"{n:>10s}" for n in names))
        for it in range(64):
            if not t[cta, 0, it]:
                break
            print(f"  {it:2d} " + " ".join(f"{rel(t[cta, e, it]):10d}" for e in range(6)))
        segs = [rel(v) for v in t[cta, 6, :12]]
        print("  epilogue (dfull seen, seg done) pairs:", segs)
    # summary over all CTAs
    ex = np.array([np.int64(t[c, 7, 1]) - np.int64(t[c, 7, 0]) for c in range(plan["grid"])])
    print("exit cycles: min %d median %d max %d" % (ex.min(), np.median(ex), ex.max()))


if __name__ == "__main__":
    main()
