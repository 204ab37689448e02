"""Timing of the residual add + RMSNorm kernels (csrc/w4a16_norm.cu).

Single process: GB/s of ``fused_add_rms_norm`` (algorithmic bytes = read x, residual; write
out, residual; + weight) against the measured copy bandwidth, rotating through buffers larger
than L2 for the big shapes.  Under torchrun (one rank per GPU): µs per call of
``all_reduce -> fused_add_rms_norm`` (two kernels) vs ``forward_reduced`` (one kernel) vs NCCL +
norm, each as a CUDA graph of dependent calls; max over ranks.

    python scripts/norm_bench.py
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/norm_bench.py
"""
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")


def graph_time(fn, iters, reps=5):
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn()
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for i in range(iters):
                fn(i)
        for _ in range(3):
            g.replay()
        best = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            g.replay()
            b.record(st)
            st.synchronize()
            best.append(a.elapsed_time(b) * 1e3 / iters)
    best.sort()
    return best[len(best) // 2], best[0]


def single():
    peak = 6544.7
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    shapes = [(16, 4096), (16, 8192), (256, 8192), (4096, 8192), (16384, 8192)]
    if len(sys.argv) > 1:                      # e.g. 4096x8192 (ncu captures)
        shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
    for rows, hidden in shapes:
        nbytes = rows * hidden * 2
        sets = max(2, min(64, (512 << 20) // (4 * nbytes) + 1))
        xs = [torch.randn(rows, hidden, device="cuda").half() for _ in range(sets)]
        rs = [torch.randn(rows, hidden, device="cuda").half() for _ in range(sets)]
        w = torch.ones(hidden, device="cuda").half()

        def fn(i=0):
            pkg.ops.fused_add_rms_norm(xs[i % sets], rs[i % sets], w, 1e-6)
        med, mn = graph_time(fn, 2 * sets)
        alg = 4 * nbytes + hidden * 2
        print(json.dumps({"op": "fused_add_rms_norm", "rows": rows, "hidden": hidden, "us": round(med, 2),
                          "us_min": round(mn, 2), "GBps": round(alg / med / 1e3, 1),
                          "hbm_frac": round(alg / med / 1e3 / peak, 3), "buffer_sets": sets}), flush=True)


def multi():
    import torch.distributed as dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    pu = pkg.parallel_utils
    pu.initialize_model_parallel(world)
    pu.custom_all_reduce.init_custom_all_reduce(max_bytes=256 * 8192 * 2, ll_bytes=256 * 8192 * 2)
    pu.custom_all_reduce._handle.oneshot_max_bytes = 1 << 30     # time the kernel at every size
    pu.custom_all_reduce._handle.ll_max_bytes = 1 << 30
    for rows, hidden in [(1, 8192), (16, 8192), (64, 8192), (256, 8192)]:
        norm = pkg.RMSNorm(hidden, eps=1e-5).half().to(dev)
        partial = torch.randn(rows, hidden, device=dev).half() * 0.01
        residual = torch.zeros(rows, hidden, device=dev).half()

        def two(i=0):
            norm(pu.tensor_model_parallel_all_reduce(partial), residual)

        def one(i=0):
            pu.custom_all_reduce._handle.all_reduce_add_rms_norm(partial, residual, norm.weight.data, 1e-5)

        def one_barrier(i=0):
            h = pu.custom_all_reduce._handle
            h.use_ll = False
            h.all_reduce_add_rms_norm(partial, residual, norm.weight.data, 1e-5)
            h.use_ll = True

        def nccl(i=0):
            t = partial.clone()
            dist.all_reduce(t)
            norm(t, residual)

        res = {}
        for name, fn in (("two_kernels", two), ("fused_barrier", one_barrier), ("fused_packets", one),
                         ("nccl_plus_norm", nccl)):
            dist.barrier()
            med, mn = graph_time(fn, 50)
            t = torch.tensor([med], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[name] = round(float(t.item()), 2)
            residual.zero_()
        if rank == 0:
            print(json.dumps({"op": "allreduce+add+rmsnorm", "world": world, "rows": rows, "hidden": hidden,
                              "us_per_call": res}), flush=True)
    pu.custom_all_reduce.destroy_custom_all_reduce()
    dist.destroy_process_group()


if __name__ == "__main__":
    if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        multi()
    else:
        single()
