#!/usr/bin/env python
"""Benchmark of the QUICK / AWQ W4A16 linear path on B200 (contract: see the task brief).

Workloads (`--workload`, all synthetic random-init AWQ-packed weights, G = 128):

  decode   (default) one decode step of Llama-2-70B: every quantized linear of every decoder
           layer (qkv, o, gate_up, down x 80), batch M = 16, tensor parallel over the N ranks
           the bench runs on (N = 1: the whole model on one B200) -- BASELINE.json configs[1]
           shapes sweep / configs[3]; `--model llama2-7b` gives the 7B stack.  Distinct weights
           per layer: one step streams 35.6 GB / N of packed weights (>> 126 MB L2).
  gemm1    BASELINE configs[0]: ONE GEMM, M = 16, 4096 -> 4096, rotating through > 2 x L2 of
           distinct weights.
  prefill  BASELINE configs[2]: Llama-2-13B prefill GEMMs (qkv, o, gate_up, down), M = 2048
           (`--batch`), compute bound: the roofline is the tensor pipe.

metric = algorithmic GB/s of the GEMMs (packed weights + scales + zeros +
activations + outputs, each counted once); decode rooflines against measured HBM bandwidth,
prefill against the measured dense bf16 tensor throughput.  The default run also reports the
other BASELINE configs under "extra" (short, N = 1 only).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch M] [--impl b200|reference]
                  [--workload decode|gemm1|prefill] [--model llama2-70b] [--no-extras]
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODELS = {
    # name: (hidden, intermediate, q_heads, kv_heads, head_dim, layers)
    "llama2-7b": (4096, 11008, 32, 32, 128, 32),
    "llama2-13b": (5120, 13824, 40, 40, 128, 40),
    "llama2-70b": (8192, 28672, 64, 8, 128, 80),
}
GROUP = 128


def layer_shapes(model: str, tp: int):
    """(name, K, N, row_parallel) per rank, sharded as vLLM's TP linear layers shard them
    (reference vllm/model_executor/layers/linear.py:312-437 QKV, :218-309 merged, :440-552 row)."""
    hidden, inter, qh, kvh, hd, _ = MODELS[model]
    kv_per_rank = max(1, kvh // tp)  # kv heads are replicated when tp > kv_heads
    qkv_n = (qh // tp + 2 * kv_per_rank) * hd
    return [
        ("qkv_proj", hidden, qkv_n, False),
        ("o_proj", hidden // tp, hidden, True),
        ("gate_up_proj", hidden, 2 * inter // tp, False),
        ("down_proj", inter // tp, hidden, True),
    ]


def bytes_alg(M, K, N, G=GROUP):
    return K * N // 2 + (K // G) * N * 2 + (K // G) * N // 2 + M * K * 2 + M * N * 2


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            d = json.load(open(path))
            return {"hbm": float(d["hbm_gbs"]), "tensor_burst": float(d["bf16_tflops"]),
                    "tensor_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    "source": "measured (MEASURED_PEAKS.json)"}
        except Exception:
            pass
    return {"hbm": 6650.0, "tensor_burst": 1650.0, "tensor_sustained": 1400.0,
            "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-i", str(index), "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [s.strip() for s in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                   "samples": len(sm)}
        return out


# ----------------------------------------------------------------------------- workloads

def workload_gemms(args, model, tp):
    """The GEMMs of one step: (name, M, K, N, row_parallel, count per step)."""
    M = args.batch
    if args.workload == "decode":
        return [(n, M, K, N, row, MODELS[model][5] if args.layers is None else args.layers)
                for (n, K, N, row) in layer_shapes(model, tp)]
    if args.workload == "gemm1":
        return [("gemm1", M, 4096, 4096, False, 64)]        # 64 x 8 MiB of distinct weights
    if args.workload == "prefill":
        return [(n, M, K, N, False, 2) for (n, K, N, _) in layer_shapes("llama2-13b", 1)]
    raise ValueError(args.workload)


def workload_config(args, model, tp):
    M = args.batch
    if args.workload == "decode":
        layers = MODELS[model][5] if args.layers is None else args.layers
        what = (f"{model} AWQ W4 g{GROUP} decode step: quantized linear layers (qkv,o,gate_up,down x {layers} "
                f"layers), batch M={M}, tensor-parallel={tp}")
        if getattr(args, "layer_ops", False):
            what += " + the layer's RMSNorms (fused behind the all-reduce) and SiLU-and-multiply (gate_up epilogue)"
    elif args.workload == "gemm1":
        what = f"single QUICK-AWQ W4 g{GROUP} GEMM M={M} K=4096 N=4096 fp16, 64 distinct weight sets per step"
    else:
        what = (f"llama2-13b AWQ W4 g{GROUP} prefill GEMMs (qkv,o,gate_up,down, 2 distinct weight sets each), "
                f"M={M} tokens, one GPU per rank")
    return {"workload": what, "global_batch": M, "group_size": GROUP,
            "parallelism": f"tp{tp}" if args.workload == "decode" else f"replicas{tp}",
            "l2_policy": "inputs larger than L2 (distinct packed weights per launch, rotating)"}


# ----------------------------------------------------------------------------- reference arm

def cpu_sample_tensors(gemms):
    import torch
    g = torch.Generator().manual_seed(0)
    tensors = []
    for (_, M, K, N, _, _) in gemms:
        tensors.append((torch.randn((M, K), generator=g).half(),
                        torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, generator=g),
                        (torch.rand((K // GROUP, N), generator=g) * 0.019 + 0.001).half(),
                        torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, generator=g)))
    return tensors


def cpu_sample_step(tensors):
    """One bounded sample of the workload through the CPU path (PyTorch dequantize + matmul)."""
    from oracle import torch_cpu
    t0 = time.perf_counter()
    for (x, qw, s, qz) in tensors:
        torch_cpu.awq_gemm(x, qw, s, qz)
    return time.perf_counter() - t0


def cpu_sample(args, model, tp):
    """The bounded CPU sample: ONE launch of every distinct GEMM of the step (for decode: the
    four linears of one decoder layer), so every shape of the workload is in it."""
    gemms = workload_gemms(args, model, tp)
    if args.workload == "prefill":          # 13B prefill on the CPU: minutes per GEMM -- M capped
        gemms = [(n, min(M, 256), K, N, r, c) for (n, M, K, N, r, c) in gemms]
    desc = "1 launch each of " + ", ".join(f"{n} M={M} {K}->{N}" for (n, M, K, N, _, _) in gemms)
    return gemms, desc


def run_reference(args, model, tp):
    """--impl reference: the reference's CPU definition of the path (PyTorch dequantize + matmul,
    oracle/torch_cpu.py) on all host cores (rank 0 only under torchrun)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    gemms, desc = cpu_sample(args, model, tp)
    tensors = cpu_sample_tensors(gemms)
    for _ in range(min(args.warmup, 1)):
        cpu_sample_step(tensors)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_sample_step(tensors)
    step_bytes = sum(bytes_alg(M, K, N) for (_, M, K, N, _, _) in gemms)
    value = step_bytes * args.steps / t / 1e9
    print(json.dumps({
        "impl": "reference", "metric": "quick_gemm_algorithmic_GBps", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.workload == "decode" else "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(args, model, tp),
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cores, "kind": "port",
                         "sample": desc + " per step (PyTorch CPU dequantize + matmul, oracle/torch_cpu.py)"},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------- B200 arm helpers

def measured_traffic(workload, model, tp, M):
    """DRAM bytes per launch of the dominant kernel from a committed `ncu --set full` capture of
    THIS workload (profiles/traffic.json, keyed "<workload>/<model>/tp<tp>/M<M>"), else None --
    never an estimate and never another workload's number."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return float(d[f"{workload}/{model}/tp{tp}/M{M}"]["avg_dram_bytes_per_launch"])
    except Exception:
        return None


def kernel_name(pkg, M, K, N):
    plan = pkg._lib.gemm_plan(M, K, N)
    if plan["variant"] == "decode":
        return f"w4a16_decode_kernel<{plan['tok_tile'] // 8}>"
    two = plan["cta_pair"]
    return (f"w4a16_gemm_kernel<{plan['tok_tile']}, {plan['unit_tiles'] // (2 if two else 1)}, "
            f"{'true' if two else 'false'}, {plan['token_tiles']}>")


def rand_awq(torch, dev, g, K, N):
    qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=g)
    qz = torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, device=dev, generator=g)
    s = (torch.rand((K // GROUP, N), device=dev, generator=g) * 0.019 + 0.001).half()
    return qw, s, qz


def oracle_sample_check(torch, pkg, x, qw_awq, s, qz, y, n_words=16, seed=0):
    """max |y - ref| / max(1, |ref|) of a GEMM output against the CPU oracle (fp32
    dequant-then-matmul, oracle/awq_oracle.c) on a random sample of 8 * n_words channels."""
    import numpy as np
    from oracle import oracle
    g = torch.Generator(device=x.device).manual_seed(seed)
    N = s.shape[1]
    words = torch.randperm(N // 8, device=x.device, generator=g)[:n_words].sort().values
    cols = (words[:, None] * 8 + torch.arange(8, device=x.device)[None, :]).reshape(-1)
    W = oracle.dequantize(qw_awq[:, words].contiguous().cpu().numpy(),
                          s[:, cols].contiguous().cpu().view(torch.int16).numpy().view(np.float16),
                          qz[:, words].contiguous().cpu().numpy())
    Yf, _ = oracle.gemm_f32(x.cpu().view(torch.int16).numpy().view(np.float16), W)
    d = np.abs(y[:, cols].float().cpu().numpy() - Yf)
    return float((d / np.maximum(np.abs(Yf), 1.0)).max(initial=0.0))


def time_graph(torch, stream, graph, iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(iters):
        graph.replay()
    e1.record(stream)
    stream.synchronize()
    return e0.elapsed_time(e1) / iters


def bench_gemm_list(torch, pkg, dev, gemms, iters=10, warm=3, seed=7):
    """Short device-timed run of a list of (name, M, K, N, sets) GEMMs on rotating weight sets
    through ops.quick_gemm, one CUDA graph; returns per-GEMM-list totals (for the extras)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    work = []
    for (name, M, K, N, sets) in gemms:
        x = torch.randn((M, K), device=dev, generator=g).half()
        for _ in range(sets):
            qw, s, qz = rand_awq(torch, dev, g, K, N)
            work.append((x, pkg.ops.quick_repack(qw), s, qz))
            del qw
    st = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(st):
        for (x, r, s, qz) in work[:2]:
            pkg.ops.quick_gemm(x, r, s, qz, 8)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            for (x, r, s, qz) in work:
                pkg.ops.quick_gemm(x, r, s, qz, 8)
        for _ in range(warm):
            graph.replay()
        st.synchronize()
        ms = time_graph(torch, st, graph, iters)
    del graph, work
    torch.cuda.empty_cache()
    b = sum(bytes_alg(M, K, N) * sets for (_, M, K, N, sets) in gemms)
    f = sum(2.0 * M * K * N * sets for (_, M, K, N, sets) in gemms)
    n = sum(sets for g_ in gemms for sets in [g_[4]])
    return {"us_per_launch": ms * 1e3 / n, "GBps": b / ms / 1e6, "TFLOPs": f / ms / 1e9, "launches": n}


def bench_moe_block(torch, pkg, dev, tokens, local_experts=1, blocks=2, iters=10, seed=11):
    """One Mixtral-8x7B MoE block (router + this rank's experts + accumulation) as a rank of a
    TP = 8 group holds it (config 5), CUDA-graph replay over `blocks` distinct blocks: the
    product path against the reference's dense loop (every local expert on every token)."""
    import types
    hidden, ffn, experts, top_k, G = 4096, 14336, 8, 2, 128
    cfg = types.SimpleNamespace(num_local_experts=experts, num_experts_per_tok=top_k,
                                hidden_size=hidden, intermediate_size=ffn)
    g = torch.Generator(device=dev).manual_seed(seed)
    method = pkg.QuickConfig(4, G, True).get_linear_method()
    torch.set_default_dtype(torch.float16)
    try:
        with torch.device(dev):
            moes = [pkg.MixtralMoE(cfg, method) for _ in range(blocks)]
    finally:
        torch.set_default_dtype(torch.float32)
    for moe in moes:
        moe.expert_indicies = list(range(local_experts))
        for e in range(local_experts, experts):
            moe.experts[e] = None
        for mlp in moe.experts:
            if mlp is None:
                continue
            for lin in (mlp.w1, mlp.w2, mlp.w3):
                lin.qweight.data.copy_(torch.randint(-2**31, 2**31 - 1, lin.qweight.shape, dtype=torch.int32, device=dev, generator=g))
                lin.qzeros.data.copy_(torch.randint(-2**31, 2**31 - 1, lin.qzeros.shape, dtype=torch.int32, device=dev, generator=g))
                lin.scales.data.copy_((torch.rand(lin.scales.shape, device=dev, generator=g) * 0.004 + 0.001).half())
        moe.gate.weight.data.copy_(torch.randn(moe.gate.weight.shape, device=dev, generator=g).half())
    x = torch.randn(1, tokens, hidden, device=dev, generator=g).half()
    res = {}
    for name, skip in (("us_per_block", True), ("us_per_block_dense_reference_loop", False)):
        for moe in moes:
            moe.skip_idle_experts = skip
        st = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(st):
            for moe in moes:
                moe(x)
            st.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=st):
                for moe in moes:
                    moe(x)
            for _ in range(3):
                graph.replay()
            st.synchronize()
            res[name] = time_graph(torch, st, graph, iters) * 1e3 / blocks
        del graph
    del moes
    torch.cuda.empty_cache()
    res["local_experts"] = local_experts
    res["tokens"] = tokens
    return res


def run_extras(torch, pkg, dev, peaks):
    """The other BASELINE configs, briefly (N = 1): kernel-only numbers with their roofline
    fraction, so every config of the metric has a driver-visible figure in the default run."""
    out = {}
    r = bench_gemm_list(torch, pkg, dev, [("gemm1", 16, 4096, 4096, 64)])
    out["config1_gemm_M16_4096x4096"] = {**r, "hbm_frac": r["GBps"] / peaks["hbm"]}
    for M in (1, 16, 64):
        gem = [(n, M, K, N, 8) for (n, K, N, _) in layer_shapes("llama2-7b", 1)]
        r = bench_gemm_list(torch, pkg, dev, gem)
        out[f"config2_llama2-7b_layer_M{M}"] = {**r, "hbm_frac": r["GBps"] / peaks["hbm"]}
    for M in (1, 8):        # the default workload's model at the other decode batch sizes (M = 16 is the headline)
        gem = [(n, M, K, N, 2) for (n, K, N, _) in layer_shapes("llama2-70b", 1)]
        r = bench_gemm_list(torch, pkg, dev, gem)
        out[f"config4_llama2-70b_layer_M{M}"] = {**r, "hbm_frac": r["GBps"] / peaks["hbm"]}
    for M in (8, 64, 256):
        gem = [("w1w3", M, 4096, 28672, 8), ("w2", M, 14336, 4096, 8)]
        r = bench_gemm_list(torch, pkg, dev, gem)
        out[f"config5_mixtral_expert_gemms_M{M}"] = {**r, "hbm_frac": r["GBps"] / peaks["hbm"]}
    for T in (8, 256):      # the whole MoE block as one rank of the TP = 8 group runs it
        out[f"config5_mixtral_moe_block_tp8_rank_T{T}"] = bench_moe_block(torch, pkg, dev, T)
    # the prefill GEMMs last: they drive the GPU into its power cap, and the memory-bound lines
    # measured right behind them read up to 1.6x slow on some boxes (profiles/r3t_ab_config5.md)
    for M in (512, 2048, 8192):
        gem = [(n, M, K, N, 2) for (n, K, N, _) in layer_shapes("llama2-13b", 1)]
        r = bench_gemm_list(torch, pkg, dev, gem, iters=4, warm=2)
        out[f"config3_llama2-13b_prefill_M{M}"] = {**r, "tensor_frac_burst": r["TFLOPs"] / peaks["tensor_burst"]}
    return out


# ----------------------------------------------------------------------------- B200 arm

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="decode", choices=["decode", "gemm1", "prefill"])
    ap.add_argument("--model", default="llama2-70b", help="decode workload: llama2-7b | llama2-13b | llama2-70b")
    ap.add_argument("--layers", type=int, default=None, help="override layer count (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the short runs of the other BASELINE configs")
    ap.add_argument("--fused-gemm-allreduce", action="store_true",
                    help="force the row-parallel GEMM fused with its all-reduce in one kernel")
    ap.add_argument("--packet-allreduce", action="store_true",
                    help="force GEMM, then the packet all-reduce kernel")
    ap.add_argument("--unfused-allreduce", action="store_true",
                    help="row-parallel GEMM and the one-shot barrier all-reduce as two kernels")
    ap.add_argument("--layer-ops", action="store_true",
                    help="decode workload as the whole linear path of a decoder layer: residual add + RMSNorm "
                         "before qkv / gate_up (fused behind the all-reduce at TP > 1: SURVEY 8f-1), SiLU-and-"
                         "multiply in the gate_up epilogue (8f-2), real data flowing o_proj -> norm -> gate_up "
                         "-> down_proj -> norm; same metric (algorithmic bytes of the GEMMs)")
    ap.add_argument("--nccl-allreduce", action="store_true",
                    help="row-parallel all-reduce through NCCL instead of the NVLink peer-memory kernels")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.batch is None:
        args.batch = 2048 if args.workload == "prefill" else 16

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    tp = world
    model = args.model if args.workload == "decode" else "llama2-13b" if args.workload == "prefill" else "none"

    if args.impl == "reference":
        run_reference(args, model, tp)
        return

    import torch
    import torch.distributed as dist
    pkg = importlib.import_module("vllm-quick_b200")
    pkg._lib.lib()  # fail loudly if the CUDA library is missing
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    M = args.batch
    gemms = workload_gemms(args, model, tp)
    pl = importlib.import_module("vllm-quick_b200.parallel_linear")
    pu = importlib.import_module("vllm-quick_b200.parallel_utils")
    decode = args.workload == "decode"
    if world > 1:
        pu.initialize_model_parallel(world if decode else 1)
    if decode and world > 1 and not args.nccl_allreduce:
        hidden = MODELS[model][0]
        pu.custom_all_reduce.init_custom_all_reduce(
            max_bytes=M * hidden * 2, fused_out_features=0 if args.unfused_allreduce else hidden)
        if args.unfused_allreduce:      # this mode means "the one-shot barrier kernel", whatever the policy says
            pu.custom_all_reduce._handle.oneshot_max_bytes = 1 << 30
            pu.custom_all_reduce._handle.use_ll = False
        if args.fused_gemm_allreduce:
            pu.custom_all_reduce._handle.prefer_fused_gemm = True
        if args.packet_allreduce:
            pu.custom_all_reduce._handle.prefer_fused_gemm = False
    method = pkg.QuickConfig(4, GROUP, True).get_linear_method()
    g = torch.Generator(device=dev).manual_seed(1234 + rank)

    keep_awq = {}                       # checkpoint-layout copy of the first layer's weights (parity check)

    def fill(layer, keep=None):
        w = layer.linear_weights
        for name in ("qweight", "qzeros"):
            t = w[name]
            t.data = torch.randint(-2**31, 2**31 - 1, tuple(t.shape), dtype=torch.int32, device=dev, generator=g)
        w["scales"].data = (torch.rand(tuple(w["scales"].shape), device=dev, generator=g) * 0.019 + 0.001).half()
        if keep is not None:
            keep_awq[keep] = w["qweight"].data.clone()
        method.process_weights_after_loading(w)
        return layer

    if decode:
        # ---- the four tensor-parallel quantized linears of every decoder layer, built through
        # the plugin's own create_weights (inside the layer classes) + post-load interleave
        hidden, inter, qh, kvh, hd, _ = MODELS[model]
        n_layers = gemms[0][5]
        shapes = layer_shapes(model, tp)
        kw = dict(bias=False, linear_method=method, params_dtype=torch.float16)
        layers = []
        for li in range(n_layers):
            k = (lambda i: i) if li == 0 else (lambda i: None)
            layers.append([
                fill(pl.QKVParallelLinear(hidden, hd, qh, kvh, **kw), k(0)),
                fill(pl.RowParallelLinear(hidden, hidden, input_is_parallel=True, **kw), k(1)),
                fill(pl.MergedColumnParallelLinear(hidden, [inter, inter], **kw), k(2)),
                fill(pl.RowParallelLinear(inter, hidden, input_is_parallel=True, **kw), k(3)),
            ])
        for (name, K, N, _), layer in zip(shapes, layers[0]):
            got = (layer.linear_weights["scales"].shape[0] * GROUP, layer.linear_weights["scales"].shape[1])
            assert got == (K, N), (name, got, (K, N))
        # inputs of the four GEMMs (attention and the activation between them are outside the
        # path: fixed stand-in activations of the right per-rank shape)
        xs = {K: torch.randn((M, K), device=dev, generator=g).half() for (_, K, _, _) in shapes}
        io_width = hidden

        def step(x_in):
            """All quantized linears of all layers; returns the last down_proj output."""
            out = None
            for lyr in layers:
                for (name, K, N, row), layer in zip(shapes, lyr):
                    inp = x_in if K == hidden and not row else xs[K]
                    out, _ = layer(inp)          # row-parallel layers all-reduce inside
            return out

        if args.layer_ops:
            # ---- the layer as models/llama.py:205-222 + :76-80 run it on this library: two RMSNorms
            # (weights = ones), row-parallel layers WITHOUT their own all-reduce feeding
            # RMSNorm.forward_reduced (all-reduce + residual add + norm in one kernel at TP > 1),
            # gate_up with the activation in its epilogue.  Attention stays a stand-in (fixed o_proj
            # input of the right per-rank shape); everything else is real data flow.
            norms = []
            for lyr in layers:
                lyr[1].reduce_results = False
                lyr[3].reduce_results = False
                norms.append((pkg.RMSNorm(hidden).to(dev).half(), pkg.RMSNorm(hidden).to(dev).half()))
            for lyr in layers:                  # channel-pair the gate_up weights once, before any capture
                lyr[2].forward_silu_and_mul(xs[hidden])
            residual0 = torch.randn((M, hidden), device=dev, generator=g).half()

            def step(x_in):                     # noqa: F811
                residual = residual0.clone()
                h = x_in.clone()
                for lyr, (n_in, n_post) in zip(layers, norms):
                    h, residual = n_in(h, residual)                      # input_layernorm
                    lyr[0](h)                                            # qkv_proj (attention: stand-in below)
                    o_partial, _ = lyr[1](xs[shapes[1][1]])              # o_proj on this rank's K shard
                    h, residual = n_post.forward_reduced(o_partial, residual)   # all-reduce + add + norm
                    act = lyr[2].forward_silu_and_mul(h)                 # gate_up + SiLU-and-multiply
                    h, _ = lyr[3](act)                                   # down_proj (partial sums at TP > 1)
                    if tp > 1:                                           # reduced by the NEXT layer's norm in a model;
                        h = pu.tensor_model_parallel_all_reduce(h)       # here: the packet all-reduce
                return h
    else:
        # ---- plain GEMM lists through the operator (ops.quick_gemm), rotating weight sets
        work = []
        xs = {}
        for (name, Mg, K, N, _, count) in gemms:
            xs.setdefault(K, torch.randn((Mg, K), device=dev, generator=g).half())
            for c in range(count):
                qw, s, qz = rand_awq(torch, dev, g, K, N)
                if not keep_awq.get(name) and c == 0:
                    keep_awq[name] = (qw.clone(), len(work))
                work.append((K, pkg.ops.quick_repack(qw), s, qz))
                del qw
        io_width = gemms[0][2]

        def step(x_in):
            out = None
            for i, (K, r, s, qz) in enumerate(work):
                out = pkg.ops.quick_gemm(x_in if i == 0 else xs[K], r, s, qz, 8)
            return out

    x_host = torch.randn((M, io_width)).half().pin_memory()
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        x_dev = xs[io_width].clone()
        for _ in range(2):
            y = step(x_dev)      # eager: warms NCCL and the per-stream scratch
        stream.synchronize()
        y_host = torch.empty(tuple(y.shape), dtype=torch.float16).pin_memory()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            y_static = step(x_dev)
        launches_per_step = sum(c for (*_, c) in gemms)     # GEMM launches (roofline denominator)
        fused_ar = decode and world > 1 and pu.custom_all_reduce.fused_handle(M, MODELS[model][0], GROUP) is not None
        ar_launches = 2 * gemms[0][5] if (decode and world > 1 and pu.custom_all_reduce.is_initialized()
                                          and not fused_ar) else 0
        if decode and args.layer_ops:        # two norm kernels per layer (+ the packet all-reduce at TP > 1)
            fused_ar = False
            ar_launches = (2 + (1 if world > 1 else 0)) * gemms[0][5]

        def barrier():
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)

        # ---- kernel-path throughput: inputs resident, graph replay, CUDA events
        for _ in range(args.warmup):
            graph.replay()
        barrier()
        sampler = ClockSampler(local_rank) if rank == 0 else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            graph.replay()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None

        # ---- end to end: host buffers, H2D + D2H inside the timed region, every step
        for _ in range(args.warmup):
            x_dev.copy_(x_host, non_blocking=True); graph.replay(); y_host.copy_(y_static, non_blocking=True)
            stream.synchronize()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(args.steps):
            x_dev.copy_(x_host, non_blocking=True)
            graph.replay()
            y_host.copy_(y_static, non_blocking=True)
            stream.synchronize()
        f1.record(stream)
        barrier()
        ms_e2e = f0.elapsed_time(f1)

        # ---- parity of what was just timed (after the timed regions; not part of them)
        parity = {"ok": True, "max_rel": 0.0, "checks": []}
        if decode and args.layer_ops:
            # the fused step against the same layers run op by op (separate all-reduce, norm and
            # activation kernels): the library's own unfused path, bit-identical by construction
            def unfused(x_in):
                residual = residual0.clone()
                h = x_in.clone()
                for lyr, (n_in, n_post) in zip(layers[:2], norms[:2]):
                    h, residual = n_in(h, residual)
                    o_partial, _ = lyr[1](xs[shapes[1][1]])
                    o_full = pu.tensor_model_parallel_all_reduce(o_partial) if tp > 1 else o_partial
                    h, residual = n_post(o_full.clone(), residual)
                    w = lyr[2].linear_weights
                    full = pkg.ops.quick_gemm(h, w["qweight"].data, w["scales"].data, w["qzeros"].data, 8)
                    act = torch.empty((M, full.shape[1] // 2), dtype=torch.float16, device=dev)
                    pkg.ops.silu_and_mul_paired(act, full)
                    h, _ = lyr[3](act)
                    if tp > 1:
                        h = pu.tensor_model_parallel_all_reduce(h)
                return h
            keep = layers, norms
            layers, norms = layers[:2], norms[:2]
            got2 = step(x_dev)
            layers, norms = keep
            want2 = unfused(x_dev)
            d = (got2.float() - want2.float()).abs() / want2.float().abs().clamp_min(1.0)
            parity["checks"].append({"what": "two decoder layers, fused ops (benchmarked path) vs the same layers "
                                             "op by op (separate all-reduce / norm / activation kernels)",
                                     "max_rel": float(d.max()), "finite": bool(torch.isfinite(got2).all())})
        elif decode:
            for i, ((name, K, N, row), layer) in enumerate(zip(shapes, layers[0])):
                w = layer.linear_weights
                inp = xs[K]
                local = pkg.ops.quick_gemm(inp, w["qweight"].data, w["scales"].data, w["qzeros"].data, 8)
                rel = oracle_sample_check(torch, pkg, inp, keep_awq[i], w["scales"].data, w["qzeros"].data,
                                          local, seed=i)
                parity["checks"].append({"what": f"{name} GEMM (this rank's shard) vs CPU oracle", "max_rel": rel})
                if row and world > 1:
                    got, _ = layer(inp)                       # the benchmarked path: fused / packets / NCCL
                    # reference: the ranks' fp16 GEMM outputs summed in fp32 (NCCL on an fp32 copy),
                    # rounded once -- what linear.py:541-542 computes up to NCCL's own fp16 rounding
                    # order (an fp16 ring sum of 8 partials is itself off by up to ~2e-2 of this)
                    ref = local.float()
                    dist.all_reduce(ref)
                    d = (got.float() - ref).abs() / ref.abs().clamp_min(1.0)
                    parity["checks"].append({"what": f"{name} GEMM + all-reduce (benchmarked path) vs "
                                                     "fp32 NCCL all-reduce of the unfused GEMM outputs",
                                             "max_rel": float(d.max())})
        else:
            for name, (qw_awq, idx) in keep_awq.items():
                K, r, s, qz = work[idx]
                inp = x_dev if idx == 0 else xs[K]
                rel = oracle_sample_check(torch, pkg, inp, qw_awq, s, qz, pkg.ops.quick_gemm(inp, r, s, qz, 8),
                                          n_words=4 if M > 512 else 16)
                parity["checks"].append({"what": f"{name} GEMM vs CPU oracle", "max_rel": rel})
        stream.synchronize()
        parity["max_rel"] = max(c["max_rel"] for c in parity["checks"])
        parity["ok"] = bool(parity["max_rel"] <= 1e-2)
        parity["tolerance"] = 1e-2

    t = torch.tensor([ms, ms_e2e, parity["max_rel"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, worst = t.tolist()
    parity["max_rel_all_ranks"] = worst
    parity["ok"] = bool(worst <= 1e-2)

    per_rank_bytes = sum(bytes_alg(Mg, K, N) * c for (_, Mg, K, N, _, c) in gemms)
    per_rank_flops = sum(2.0 * Mg * K * N * c for (_, Mg, K, N, _, c) in gemms)
    total_bytes = per_rank_bytes * world
    ms_step = ms / args.steps
    value = total_bytes / (ms_step * 1e-3) / 1e9
    e2e_value = total_bytes / (ms_e2e / args.steps * 1e-3) / 1e9
    peaks = load_peaks()
    # dominant kernel = the GEMM (every launch in the step); per-launch averages on one GPU
    alg_per_launch = per_rank_bytes / launches_per_step
    us_per_launch = ms_step * 1e3 / launches_per_step
    if args.workload == "prefill":
        achieved = per_rank_flops / launches_per_step / (us_per_launch * 1e-6) / 1e12
        # a multi-millisecond step of back-to-back GEMMs: the sustained figure is the peak
        roof = {"bound": "tensor", "achieved": achieved, "peak": peaks["tensor_sustained"], "unit": "TFLOP/s",
                "frac": achieved / peaks["tensor_sustained"], "frac_of_burst_peak": achieved / peaks["tensor_burst"],
                "alg_flops_per_launch": per_rank_flops / launches_per_step}
    else:
        achieved = alg_per_launch / (us_per_launch * 1e-6) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm"], "unit": "GB/s",
                "frac": achieved / peaks["hbm"], "alg_bytes_per_launch": alg_per_launch}
    roof.update({"traffic": measured_traffic(args.workload, model, tp, M), "peak_source": peaks["source"],
                 "kernel": kernel_name(pkg, gemms[0][1], gemms[0][2], gemms[0][3]),
                 "us_per_launch": us_per_launch})

    if rank == 0:
        line = {
            "metric": "quick_gemm_algorithmic_GBps", "value": value, "unit": "GB/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong" if decode else "weak",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": workload_config(args, model, tp),
            "tokens_per_s": M / (ms_step * 1e-3) * (1 if decode else world),
            "gpu_launches": (launches_per_step + ar_launches) * args.steps,
            "layer_ops": bool(decode and args.layer_ops),
            "allreduce": None if not (decode and world > 1) else (
                "o_proj: all-reduce + residual add + RMSNorm in one kernel (packets); down_proj: packet kernel"
                if args.layer_ops and pu.custom_all_reduce.is_initialized() else
                "fused into the row-parallel GEMM, packet exchange (b200_w4a16_gemm_allreduce)" if fused_ar else
                ("nvlink packet kernel after the GEMM (b200_w4a16_allreduce_ll)"
                 if pu.custom_all_reduce._handle.use_ll else
                 "nvlink one-shot barrier kernel after the GEMM (b200_w4a16_allreduce)")
                if ar_launches else "nccl"),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": x_host.numel() * 2,
                    "d2h_bytes_per_step": y_host.numel() * 2, "ms_per_step": ms_e2e / args.steps},
            "roofline": roof,
            "parity_check": parity,
        }
        if not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            cgemms, desc = cpu_sample(args, model, tp)
            tensors = cpu_sample_tensors(cgemms)
            reps, tcpu = 0, 0.0
            budget = 12.0 if world == 1 else 4.0
            while tcpu < budget and reps < 5:
                tcpu += cpu_sample_step(tensors)
                reps += 1
            cpu_bytes = sum(bytes_alg(Mg, K, N) for (_, Mg, K, N, _, _) in cgemms) * reps
            line["cpu_baseline"] = {
                "value": cpu_bytes / tcpu / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
                "sample": f"{desc} x {reps} reps, PyTorch CPU dequantize + matmul (oracle/torch_cpu.py)"}
            del tensors
        if world == 1 and decode and not args.no_extras:
            del graph
            layers = None
            torch.cuda.empty_cache()
            try:
                line["extra"] = run_extras(torch, pkg, dev, peaks)
            except Exception as e:  # the headline line must survive a failing side measurement
                line["extra"] = {"error": repr(e)}
        print(json.dumps(line))
    if world > 1:
        # Tear down without destroying communicators that captured graphs still reference
        # (an NCCL communicator with live captured collectives can block in its destructor):
        # drop the graph, drain the device, meet the other ranks, then leave the process.
        sys.stdout.flush()
        graph = None
        torch.cuda.synchronize(dev)
        dist.barrier()
        pu.custom_all_reduce.destroy_custom_all_reduce()
        torch.cuda.synchronize(dev)
        os._exit(0)


if __name__ == "__main__":
    main()
