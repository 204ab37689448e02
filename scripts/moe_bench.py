#!/usr/bin/env python
"""BASELINE configs[4]: Mixtral-8x7B AWQ expert MLP GEMMs at mixed batch sizes, one GPU
holding `--experts-per-rank` of the 8 experts (8 = TP 1, 1 = one rank of TP 8).
Times one MoE block (router + experts; CUDA graph over `--layers` distinct blocks so the
weights exceed L2) with idle experts skipped on the device (b200_w4a16_gemm_gated) and with
the reference's dense loop (every local expert streams its weights for every batch).
Usage: python scripts/moe_bench.py [--tokens 1,4,16] [--layers 3] [--experts-per-rank 8]
Under torchrun (WORLD_SIZE > 1) it is the real thing: every rank builds its share of the experts
(`MixtralMoE`'s own split), the block ends in the tensor-parallel all-reduce (NVLink kernels), time =
max over ranks, the ranks' results are compared bit for bit:
  torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 scripts/moe_bench.py --tokens 1,8,64,256
"""
import argparse
import importlib
import json
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")
G = 128


def fill(moe, gen):
    for mlp in moe.experts:
        if mlp is None:
            continue
        for lin in (mlp.w1, mlp.w2, mlp.w3):
            lin.qweight.data.copy_(torch.randint(-2**31, 2**31 - 1, lin.qweight.shape, dtype=torch.int32,
                                                 device="cuda", generator=gen))
            lin.qzeros.data.copy_(torch.randint(-2**31, 2**31 - 1, lin.qzeros.shape, dtype=torch.int32,
                                                device="cuda", generator=gen))
            lin.scales.data.copy_((torch.rand(lin.scales.shape, device="cuda", generator=gen) * 0.004
                                   + 0.001).half())
    moe.gate.weight.data.copy_(torch.randn(moe.gate.weight.shape, device="cuda", generator=gen).half())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", default="1,4,16")
    ap.add_argument("--layers", type=int, default=3)
    ap.add_argument("--experts-per-rank", type=int, default=8)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--prefill", action="store_true",
                    help="eager launches (no CUDA graph): batches above 16 tokens take the routed "
                         "gather -> per-expert GEMM -> scatter-add path; compared with the dense loop")
    ap.add_argument("--routed-on-device-min", type=int, default=None,
                    help="override MixtralMoE.routed_on_device_min_tokens (graph mode, batches above 16 tokens)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    dist = pu = None
    if world > 1:
        import torch.distributed as dist
        local = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        pu = pkg.parallel_utils
        pu.initialize_model_parallel(world)
        pu.custom_all_reduce.init_custom_all_reduce(max_bytes=max(int(v) for v in args.tokens.split(",")) * 4096 * 2)
        args.experts_per_rank = 8 // world
    hidden, ffn, experts, top_k = 4096, 14336, 8, 2
    cfg = types.SimpleNamespace(num_local_experts=experts, num_experts_per_tok=top_k,
                                hidden_size=hidden, intermediate_size=ffn)
    gen = torch.Generator(device="cuda").manual_seed(0)
    method = pkg.QuickConfig(4, G, True).get_linear_method()
    torch.set_default_dtype(torch.float16)
    blocks = []
    with torch.device("cuda"):
        for _ in range(args.layers):
            moe = pkg.MixtralMoE(cfg, method)
            if world == 1:
                # one rank of a wider tensor-parallel group: keep the first n experts only
                moe.expert_indicies = list(range(args.experts_per_rank))
                for e in range(args.experts_per_rank, experts):
                    moe.experts[e] = None
            blocks.append(moe)
    torch.set_default_dtype(torch.float32)
    for moe in blocks:
        fill(moe, gen)
        if world > 1:                      # the router is replicated: same weights on every rank
            dist.broadcast(moe.gate.weight.data, 0)
        if args.routed_on_device_min is not None:
            moe.routed_on_device_min_tokens = args.routed_on_device_min
    bytes_expert = 3 * (hidden * ffn // 2 + (hidden // G) * ffn * 5 // 2)   # packed + scales + zeros (approx.)
    if args.prefill:
        for moe in blocks:
            moe.routed_min_tokens = 17        # show the crossover; the product threshold is 512
        for T in (int(v) for v in args.tokens.split(",")):
            x = torch.randn(1, T, hidden, device="cuda", generator=gen).half()
            res = {}
            # True: host-routed (one read of the counts); "device": routed with the counts read on
            # the device (no synchronisation); False: the reference's dense loop
            for skip in (True, "device", False):
                for moe in blocks:
                    moe.skip_idle_experts = bool(skip)
                    moe.routed_min_tokens = (1 << 30) if skip == "device" else 17
                    moe.routed_on_device_min_tokens = 17
                for moe in blocks:
                    moe(x)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.iters):
                    for moe in blocks:
                        moe(x)
                e1.record()
                torch.cuda.synchronize()
                res[skip] = e0.elapsed_time(e1) * 1e3 / (args.iters * len(blocks))
            print(json.dumps({"mode": "prefill_eager", "tokens": T, "local_experts": args.experts_per_rank,
                              "us_per_block_routed": round(res[True], 1),
                              "us_per_block_routed_on_device": round(res["device"], 1),
                              "us_per_block_dense_reference_loop": round(res[False], 1),
                              "speedup": round(res[False] / res[True], 2)}), flush=True)
        return
    for T in (int(v) for v in args.tokens.split(",")):
        x = torch.randn(1, T, hidden, device="cuda", generator=gen).half()
        if world > 1:
            dist.broadcast(x, 0)
        res = {}
        same = True
        # grouped: one launch per projection over the local experts with tokens (<= 16 tokens);
        # gated: one gated launch per expert; False: the reference's dense loop
        for skip in (True, "gated", False):
            for moe in blocks:
                moe.skip_idle_experts = bool(skip)
                moe.grouped_max_tokens = 0 if skip == "gated" else 16
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for moe in blocks:
                    moe(x)
                st.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=st):
                    for moe in blocks:
                        moe(x)
                for _ in range(3):
                    graph.replay()
                st.synchronize()
                if world > 1:
                    dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(args.iters):
                    graph.replay()
                e1.record(st)
                st.synchronize()
            res[skip] = e0.elapsed_time(e1) * 1e3 / (args.iters * len(blocks))
            if world > 1:
                t = torch.tensor([res[skip]], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                res[skip] = float(t)
                y = blocks[0](x)                             # every rank must hold the same bits
                ys = [torch.empty_like(y) for _ in range(world)]
                dist.all_gather(ys, y)
                same = same and all(torch.equal(ys[0], t_) for t_ in ys)
            del graph
        active = []
        for moe in blocks:
            _, _, counts = moe.route(x.view(-1, hidden))
            active.append(int((counts[:args.experts_per_rank] > 0).sum()))
        mean_active = sum(active) / len(active)
        if rank != 0:
            continue
        print(json.dumps({"tokens": T, "local_experts": args.experts_per_rank, "world": world,
                          "ranks_bit_identical": same,
                          "active_experts_mean": round(mean_active, 2),
                          "us_per_block_skip_idle": round(res[True], 1),
                          "us_per_block_gated_launch_per_expert": round(res["gated"], 1),
                          "us_per_block_dense_reference_loop": round(res[False], 1),
                          "speedup": round(res[False] / res[True], 2),
                          "GBps_active_weights": round(mean_active * bytes_expert / res[True] / 1e3, 1)}), flush=True)
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        pu.custom_all_reduce.destroy_custom_all_reduce()
        torch.cuda.synchronize()
        os._exit(0)


if __name__ == "__main__":
    main()
