#!/usr/bin/env python
"""Benchmark of the QUICK / AWQ W4A16 linear path on B200 (contract: see the task brief).

A "step" is one pass of the hot path over one decode batch: every quantized linear of every
decoder layer of a Llama-2 model (qkv, o, gate_up, down), synthetic random-init AWQ-packed
weights, distinct per layer, so one step streams several GB of packed weights (far more
than the 126 MB L2) -- BASELINE.json configs[1] at 1 GPU (Llama-2-7B shapes) and
configs[3] under torchrun (Llama-2-70B, tensor parallel over the ranks with an NCCL
all-reduce after the row-parallel o_proj / down_proj).

metric = algorithmic GB/s of the GEMMs (packed weights + scales + zeros +
activations + outputs, each counted once) -- the roofline for decode is HBM bandwidth.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch M] [--impl b200|reference]
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
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


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


# ----------------------------------------------------------------------------- reference arm

def cpu_reference_step(sample_shapes, M, tensors=None):
    """One bounded sample of the workload through the CPU path (PyTorch dequantize + matmul)."""
    import torch
    from oracle import torch_cpu
    if tensors is None:
        g = torch.Generator().manual_seed(0)
        tensors = []
        for (_, K, N, _) in sample_shapes:
            tensors.append((torch.randn((M, K), generator=g).half(),
                            torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, generator=g),
                            (torch.rand((K // GROUP, N), generator=g) * 0.019 + 0.001).half(),
                            torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, generator=g)))
        return tensors
    t0 = time.perf_counter()
    for (x, qw, s, qz) in tensors:
        torch_cpu.awq_gemm(x, qw, s, qz)
    return time.perf_counter() - t0


def run_reference(args, model, tp):
    """--impl reference: the reference's CPU path on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    M = args.batch
    shapes = layer_shapes(model, tp)
    sample = [shapes[1]]  # one o_proj-sized GEMM per step keeps K steps within minutes
    tensors = cpu_reference_step(sample, M)
    for _ in range(args.warmup):
        cpu_reference_step(sample, M, tensors)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_reference_step(sample, M, tensors)
    step_bytes = sum(bytes_alg(M, K, N) for (_, K, N, _) in sample)
    value = step_bytes * args.steps / t / 1e9
    desc = f"1 GEMM/step: {sample[0][0]} M={M} K={sample[0][1]} N={sample[0][2]} G={GROUP}"
    print(json.dumps({
        "impl": "reference", "metric": "quick_gemm_algorithmic_GBps", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(model, tp, M),
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cores, "kind": "port",
                         "sample": desc + " (PyTorch CPU dequantize + matmul, oracle/torch_cpu.py)"},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def measured_traffic(model, tp, M):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture
    (profiles/traffic_decode16.json: one Llama-2-7B layer at M = 16), or None for any other
    workload -- never an estimate."""
    if model != "llama2-7b" or tp != 1 or M != 16:
        return None
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "traffic_decode16.json")))[
            "avg_dram_bytes_per_launch"])
    except Exception:
        return None


def kernel_name(pkg, M, K, N):
    plan = pkg._lib.gemm_plan(M, K, N)
    if plan["variant"] == "decode":
        return f"w4a16_decode_kernel<{plan['tok_tile'] // 8}>"
    return f"w4a16_gemm_kernel<{plan['tok_tile']}>"


def workload_config(model, tp, M):
    return {"workload": f"{model} AWQ W4 g{GROUP} decode linear layers (qkv,o,gate_up,down x all layers), "
                        f"batch M={M}, tensor-parallel={tp}",
            "global_batch": M, "group_size": GROUP, "parallelism": f"tp{tp}",
            "l2_policy": "inputs larger than L2 (distinct packed weights per layer)"}


# ----------------------------------------------------------------------------- B200 arm

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default=None, help="llama2-7b | llama2-13b | llama2-70b")
    ap.add_argument("--layers", type=int, default=None, help="override layer count (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fused-gemm-allreduce", action="store_true",
                    help="force the row-parallel GEMM fused with its all-reduce in one kernel (default at TP > 2)")
    ap.add_argument("--packet-allreduce", action="store_true",
                    help="force GEMM, then the packet all-reduce kernel (default at TP = 2)")
    ap.add_argument("--unfused-allreduce", action="store_true",
                    help="row-parallel GEMM and the peer-memory all-reduce as two kernels")
    ap.add_argument("--nccl-allreduce", action="store_true",
                    help="row-parallel all-reduce through NCCL instead of the NVLink peer-memory kernel")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    tp = world
    model = args.model or ("llama2-7b" if tp == 1 else "llama2-70b")

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
    shapes = layer_shapes(model, tp)
    n_layers = args.layers or MODELS[model][5]
    hidden, inter, qh, kvh, hd, _ = MODELS[model]
    pl = importlib.import_module("vllm-quick_b200.parallel_linear")
    pu = importlib.import_module("vllm-quick_b200.parallel_utils")
    if world > 1:
        pu.initialize_model_parallel(world)
        if not args.nccl_allreduce:
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

    # ---- the four tensor-parallel quantized linears of every decoder layer, built through the
    # plugin's own create_weights (inside the layer classes) + post-load interleave
    def fill(layer):
        w = layer.linear_weights
        for name in ("qweight", "qzeros"):
            t = w[name]
            t.data = torch.randint(-2**31, 2**31 - 1, tuple(t.shape), dtype=torch.int32, device=dev, generator=g)
        w["scales"].data = (torch.rand(tuple(w["scales"].shape), device=dev, generator=g) * 0.019 + 0.001).half()
        method.process_weights_after_loading(w)
        return layer

    kw = dict(bias=False, linear_method=method, params_dtype=torch.float16)
    layers = []
    for _ in range(n_layers):
        layers.append([
            fill(pl.QKVParallelLinear(hidden, hd, qh, kvh, **kw)),
            fill(pl.RowParallelLinear(hidden, hidden, input_is_parallel=True, **kw)),
            fill(pl.MergedColumnParallelLinear(hidden, [inter, inter], **kw)),
            fill(pl.RowParallelLinear(inter, hidden, input_is_parallel=True, **kw)),
        ])
    for (name, K, N, _), layer in zip(shapes, layers[0]):
        got = (layer.linear_weights["scales"].shape[0] * GROUP, layer.linear_weights["scales"].shape[1])
        assert got == (K, N), (name, got, (K, N))
    # inputs of the four GEMMs (attention and the activation between them are outside the path:
    # fixed stand-in activations of the right per-rank shape)
    xs = {K: torch.randn((M, K), device=dev, generator=g).half() for (_, K, _, _) in shapes}
    x_host = torch.randn((M, hidden)).half().pin_memory()
    y_host = torch.empty((M, hidden), dtype=torch.float16).pin_memory()

    def step(x_in):
        """All quantized linears of all layers; returns the last down_proj output."""
        out = None
        for lyr in layers:
            for (name, K, N, row), layer in zip(shapes, lyr):
                inp = x_in if K == hidden and not row else xs[K]
                out, _ = layer(inp)          # row-parallel layers all-reduce inside
        return out

    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        x_dev = xs[hidden].clone()
        for _ in range(2):
            y = step(x_dev)      # eager: allocates the per-stream workspace, warms NCCL
        stream.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            y_static = step(x_dev)
        launches_per_step = n_layers * len(shapes)          # GEMM launches (roofline denominator)
        fused_ar = world > 1 and pu.custom_all_reduce.fused_handle(M, hidden, GROUP) is not None
        ar_launches = 2 * n_layers if (world > 1 and pu.custom_all_reduce.is_initialized()
                                       and not fused_ar) else 0

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

    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = t.tolist()

    per_rank_bytes = n_layers * sum(bytes_alg(M, K, N) for (_, K, N, _) in shapes)
    total_bytes = per_rank_bytes * world
    ms_step = ms / args.steps
    value = total_bytes / (ms_step * 1e-3) / 1e9
    e2e_value = total_bytes / (ms_e2e / args.steps * 1e-3) / 1e9
    peak, peak_src = load_peaks()
    # dominant kernel = the GEMM (every launch in the step); per-launch averages on one GPU
    alg_per_launch = per_rank_bytes / launches_per_step
    us_per_launch = ms_step * 1e3 / launches_per_step
    achieved = alg_per_launch / (us_per_launch * 1e-6) / 1e9

    if rank == 0:
        line = {
            "metric": "quick_gemm_algorithmic_GBps", "value": value, "unit": "GB/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": workload_config(model, tp, M),
            "tokens_per_s": M / (ms_step * 1e-3),
            "gpu_launches": (launches_per_step + ar_launches) * args.steps,
            "allreduce": None if world == 1 else (
                "fused into the row-parallel GEMM (b200_w4a16_gemm_allreduce)" if fused_ar else
                ("nvlink packet kernel after the GEMM (b200_w4a16_allreduce_ll)"
                 if pu.custom_all_reduce._handle.use_ll else
                 "nvlink one-shot barrier kernel after the GEMM (b200_w4a16_allreduce)")
                if ar_launches else "nccl"),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": M * hidden * 2,
                    "d2h_bytes_per_step": M * hidden * 2, "ms_per_step": ms_e2e / args.steps},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": measured_traffic(model, tp, M),
                         "peak_source": peak_src,
                         "kernel": kernel_name(pkg, M, shapes[0][1], shapes[0][2]),
                         "alg_bytes_per_launch": alg_per_launch, "us_per_launch": us_per_launch},
        }
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            tensors = cpu_reference_step(shapes, M)
            cpu_reference_step(shapes[1:2], M, tensors[1:2])  # warm-up on the small GEMM
            reps, tcpu = 0, 0.0
            while tcpu < 10.0 and reps < 5:
                tcpu += cpu_reference_step(shapes, M, tensors)
                reps += 1
            cpu_bytes = sum(bytes_alg(M, K, N) for (_, K, N, _) in shapes) * reps
            line["cpu_baseline"] = {
                "value": cpu_bytes / tcpu / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
                "sample": f"1 of {n_layers} decoder layers (4 GEMMs, M={M}) x {reps} reps, "
                          "PyTorch CPU dequantize + matmul (oracle/torch_cpu.py)"}
        print(json.dumps(line))
    if world > 1:
        # Tear down without destroying communicators that captured graphs still reference
        # (an NCCL communicator with live captured collectives can block in its destructor):
        # drop the graph, drain the device, meet the other ranks, then leave the process.
        sys.stdout.flush()
        del graph
        torch.cuda.synchronize(dev)
        dist.barrier()
        pu.custom_all_reduce.destroy_custom_all_reduce()
        torch.cuda.synchronize(dev)
        os._exit(0)


if __name__ == "__main__":
    main()
