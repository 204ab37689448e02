#!/usr/bin/env python
"""Timeline of back-to-back decode-kernel launches (needs `make -C vllm-quick_b200/csrc trace`).
Prints, per launch, when each phase of the CTAs happens (globaltimer ns relative to the first
CTA entry of the first traced launch).  Usage: python scripts/trace_decode.py --m 16 --k 4096 --n 4096"""
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

NAMES = ["entry", "w_filled", "prod_dep", "cons_dep", "tile0", "seg0_done", "seg0_out", "exit",
         "lastseg_done"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=16)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--n", type=int, default=4096)
    ap.add_argument("--launches", type=int, default=6)
    ap.add_argument("--kernel", default=None, choices=["decode", "tcgen05"])
    ap.add_argument("--flags", default=None, help="B200_W4A16_FLAGS for the trace build")
    args = ap.parse_args()
    if args.flags:
        os.environ["B200_W4A16_FLAGS"] = args.flags
    if args.kernel:
        os.environ["B200_W4A16_KERNEL"] = args.kernel      # honoured by tuning / trace builds only
    M, K, N, G = args.m, args.k, args.n, 128
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    sets = []
    for _ in range(args.launches):
        qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=g)
        qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device=dev, generator=g)
        s = (torch.rand((K // G, N), device=dev, generator=g) * 0.019 + 0.001).half()
        sets.append((qw, s, qz))
    x = torch.randn((M, K), device=dev, generator=g).half()
    L = pkg._lib.lib()
    L.b200_w4a16_debug_set_trace.argtypes = [ctypes.c_void_p]
    L.b200_w4a16_debug_set_trace.restype = None
    grid = pkg._lib.gemm_plan(M, K, N)["grid"]
    nl = len(sets)
    traces = torch.zeros((nl, grid, 128), dtype=torch.int32, device=dev)
    for (qw, s, qz) in sets[:2]:
        pkg.ops.quick_gemm(x, qw, s, qz, 8)
    torch.cuda.synchronize()
    junk = torch.empty(512 << 20, dtype=torch.uint8, device=dev); junk.zero_(); torch.cuda.synchronize()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        pkg.ops.quick_gemm(x, *sets[0], 8)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            for i, (qw, s, qz) in enumerate(sets):
                L.b200_w4a16_debug_set_trace(traces[i].data_ptr())
                pkg.ops.quick_gemm(x, qw, s, qz, 8)
        L.b200_w4a16_debug_set_trace(None)
        graph.replay(); st.synchronize()
        junk.zero_(); st.synchronize()
        graph.replay(); st.synchronize()
    ta = traces.cpu().numpy().astype(np.uint32).astype(np.int64)
    g0 = ta[0, :, 0].min()
    print("plan", pkg._lib.gemm_plan(M, K, N), "tiles/CTA", ta[0, :, 9].min(), ta[0, :, 9].max(),
          "segs", ta[0, :, 10].min(), ta[0, :, 10].max())
    for i in range(nl):
        row = []
        for e, nm in enumerate(NAMES):
            v = ta[i, :, e] - g0
            row.append(f"{nm} {v.min()}/{int(np.median(v))}/{v.max()}")
        print(f"launch {i}: " + " | ".join(row))
        cyc = (ta[i, :, 12] - ta[i, :, 11]) % (1 << 32)
        ns = ta[i, :, 7] - ta[i, :, 3]
        mhz = 1e3 * cyc / np.maximum(ns, 1)
        print(f"          SM clock inside the kernel (clock64 / globaltimer): {mhz.min():.0f}/{np.median(mhz):.0f}/{mhz.max():.0f} MHz")
    # the slowest CTAs of the last launch: who makes the dependents wait?
    last = ta[nl - 1]
    rel0 = int(np.median(last[:, 3]))
    order = np.argsort(-last[:, 7])
    print(f"== launch {nl - 1}: CTAs by exit time (ns after the median dependency release)")
    for cta in list(order[:10]) + list(order[-3:]):
        r = last[cta]
        fix = f" fix_begin {r[13] - rel0:6d} fix_summed {r[14] - rel0:6d}" if r[13] else ""
        print(f"   cta {cta:3d} tiles {r[9]:3d} segs {r[10]} entry {r[0] - rel0:6d} dep {r[3] - rel0:6d} tile0 {r[4] - rel0:6d} "
              f"seg0_done {r[5] - rel0:6d} seg0_out {r[6] - rel0:6d} lastseg_done {r[8] - rel0:6d}{fix} exit {r[7] - rel0:6d}")
    # per-tile stamps of group 0 / warp 0 in the last launch, three CTAs
    for cta in (0, grid // 2, grid - 1):
        rel = ta[nl - 1, cta, 3]  # consumer past griddepcontrol.wait
        print(f"== launch {nl - 1} CTA {cta}: per tile (wait begins, landed, math done) ns after dependency release")
        for k in range(28):
            a, b, c, d = (ta[nl - 1, cta, 16 + 4 * k + e] for e in range(4))
            if a == 0:
                break
            print(f"   tile {k:2d}: issued {d - rel:6d} | {a - rel:6d} {b - rel:6d} {c - rel:6d}   "
                  f"wait {b - a:5d} math {c - b:5d} issue->landed {b - d:6d}")


if __name__ == "__main__":
    main()
