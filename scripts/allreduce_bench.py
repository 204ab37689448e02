#!/usr/bin/env python
"""All-reduce latency by message size on N GPUs of one node: packet / one-shot / two-shot kernels
of csrc/w4a16_allreduce.cu against NCCL ("us per call at 16 KiB ... 4 MiB; bus GB/s").
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/allreduce_bench.py
Every variant is timed as a CUDA graph of `--calls` dependent all-reduces (device time, max over
ranks); results are checked against NCCL and for bit-equality across ranks."""
import argparse
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")
pu = pkg.parallel_utils


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-kib", default="16,64,256,512,1024,2048,4096,8192")
    ap.add_argument("--calls", type=int, default=20)
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pu.initialize_model_parallel(world)
    pu.custom_all_reduce.init_custom_all_reduce(max_bytes=8 << 20, ll_bytes=1 << 20)
    h = pu.custom_all_reduce._handle
    g = torch.Generator(device=dev).manual_seed(rank)

    def timed(fn, x):
        st = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(st):
            y = fn(x)
            st.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=st):
                t = x
                for _ in range(args.calls):
                    t = fn(t)
            for _ in range(3):
                graph.replay()
            st.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(args.iters):
                graph.replay()
            e1.record(st)
            st.synchronize()
        us = torch.tensor([e0.elapsed_time(e1) * 1e3 / (args.iters * args.calls)], device=dev)
        dist.all_reduce(us, op=dist.ReduceOp.MAX)
        return float(us), y

    def nccl(t):
        o = t.clone()
        dist.all_reduce(o)
        return o

    def variant(ll, oneshot_max, twoshot):
        def fn(t):
            h.use_ll, h.oneshot_max_bytes, h.use_twoshot, h.twoshot_min_bytes = ll, oneshot_max, twoshot, 0
            o = pu.custom_all_reduce.maybe_all_reduce(t)
            assert o is not None
            return o
        return fn

    for kib in (int(v) for v in args.sizes_kib.split(",")):
        n = kib * 1024 // 2
        x = (torch.randn(n, device=dev, generator=g) * 0.01).half()     # 20 chained sums stay finite
        row = {"world": world, "KiB": kib}
        us, ref = timed(nccl, x)
        row["nccl_us"] = round(us, 2)
        cands = [("oneshot", variant(False, 1 << 30, False)), ("twoshot", variant(False, 0, True))]
        if kib * 1024 <= h.ll_bytes:
            cands.insert(0, ("packets", variant(True, 0, False)))
        for name, fn in cands:
            h.ll_max_bytes = 1 << 30
            us, y = timed(fn, x)
            ok = bool(torch.allclose(y.float(), ref.float(), rtol=2e-3, atol=2e-2))
            gathered = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(gathered, y)
            same = all(torch.equal(gathered[0], t) for t in gathered)
            row[name + "_us"] = round(us, 2)
            row[name + "_ok"] = ok and same
        best = min((row[k], k) for k in row if k.endswith("_us"))
        # bus bandwidth of an all-reduce: 2 (W - 1) / W x bytes / time
        row["best"] = best[1][:-3]
        row["best_bus_GBps"] = round(2 * (world - 1) / world * kib * 1024 / best[0] / 1e3, 1)
        if rank == 0:
            print(json.dumps(row), flush=True)
    torch.cuda.synchronize(dev)
    dist.barrier()
    pu.custom_all_reduce.destroy_custom_all_reduce()
    os._exit(0)


if __name__ == "__main__":
    main()
