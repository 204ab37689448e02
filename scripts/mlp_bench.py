"""Llama MLP block (gate_up_proj -> SiluAndMul -> down_proj, llama.py:48-80), decode batches:
activation in the gate_up GEMM's epilogue (LlamaMLP(fuse_activation=True)) vs the reference's
three steps (GEMM, silu_and_mul kernel, GEMM).  CUDA graph of `layers` dependent blocks with
distinct weights (> L2), µs per block, median of 5 replays.

    python scripts/mlp_bench.py [--model llama2-7b|llama2-70b-tp8] [--layers 16]
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

SHAPES = {"llama2-7b": (4096, 11008), "llama2-13b": (5120, 13824), "llama2-70b-tp8": (8192, 28672 // 8)}


def fill(layer, gen):
    for w in (layer.gate_up_proj.linear_weights, layer.down_proj.linear_weights):
        for name in ("qweight", "qzeros"):
            w[name].data = torch.randint(-2**31, 2**31 - 1, tuple(w[name].shape), dtype=torch.int32,
                                         device="cuda", generator=gen)
        w["scales"].data = (torch.rand(tuple(w["scales"].shape), device="cuda", generator=gen) * 0.019
                            + 0.001).half()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama2-7b")
    ap.add_argument("--layers", type=int, default=16)
    ap.add_argument("--tokens", default="1,8,16")
    args = ap.parse_args()
    hidden, inter = SHAPES[args.model]
    method = pkg.QuickConfig(4, 128, True).get_linear_method()
    gen = torch.Generator(device="cuda").manual_seed(0)
    stacks = {}
    for fuse in (True, False):
        blocks = []
        for _ in range(args.layers):
            mlp = pkg.LlamaMLP(hidden, inter, "silu", method, torch.float16, fuse_activation=fuse).cuda()
            fill(mlp, gen)
            blocks.append(mlp)
        stacks[fuse] = blocks
    for M in [int(t) for t in args.tokens.split(",")]:
        x = (torch.randn(M, hidden, device="cuda", generator=gen) * 0.1).half()
        res = {}
        for fuse, blocks in stacks.items():
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                def run():
                    h = x
                    for b in blocks:
                        h = b(h) * 0 + x          # keep magnitudes bounded, keep the dependency
                    return h
                run()
                st.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    run()
                for _ in range(3):
                    g.replay()
                ts = []
                for _ in range(5):
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(st); g.replay(); b.record(st); st.synchronize()
                    ts.append(a.elapsed_time(b) * 1e3 / len(blocks))
                ts.sort()
                res["fused_epilogue" if fuse else "three_steps"] = round(ts[2], 2)
        res["speedup"] = round(res["three_steps"] / res["fused_epilogue"], 3)
        print(json.dumps({"op": "llama_mlp_block", "model": args.model, "tokens": M, "layers": args.layers,
                          "us_per_block": res}), flush=True)


if __name__ == "__main__":
    main()
