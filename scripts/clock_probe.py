#!/usr/bin/env python
"""Tuning probe: does the SM clock hold while the decode GEMM streams?  Replays a CUDA graph of
back-to-back GEMMs for a few seconds while a thread samples NVML (SM clock, power, throttle
reasons) every 10 ms; prints the per-launch time of each 250 ms window next to the clocks.
Usage: python scripts/clock_probe.py [--m 16] [--shape 16384x22016] [--seconds 4]
"""
import argparse
import importlib
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")
G = 128


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=16)
    ap.add_argument("--shape", default="16384x22016")
    ap.add_argument("--seconds", type=float, default=4.0)
    args = ap.parse_args()
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    K, N = (int(v) for v in args.shape.split("x"))
    nsets = max(2, (512 << 20) // (K * N // 2) + 1)
    sets = []
    for _ in range(nsets):
        qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=g)
        qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device=dev, generator=g)
        s = (torch.rand((K // G, N), device=dev, generator=g) * 0.019 + 0.001).half()
        sets.append((qw, s, qz))
    x = torch.randn((args.m, K), device=dev, generator=g).half()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for (qw, s, qz) in sets[:2]:
            pkg.ops.quick_gemm(x, qw, s, qz, 8)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            for (qw, s, qz) in sets:
                pkg.ops.quick_gemm(x, qw, s, qz, 8)
        graph.replay()
        st.synchronize()
    samples = []
    stop = False

    def sampler():
        while not stop:
            try:
                samples.append((time.perf_counter(),
                                pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
                                pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_MEM),
                                pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0,
                                pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)))
            except Exception as e:  # noqa: BLE001
                samples.append((time.perf_counter(), -1, -1, -1.0, str(e)))
            time.sleep(0.01)

    th = threading.Thread(target=sampler)
    th.start()
    t_end = time.perf_counter() + args.seconds
    windows = []
    with torch.cuda.stream(st):
        while time.perf_counter() < t_end:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            reps = 0
            e0.record(st)
            while time.perf_counter() - t0 < 0.25:
                for _ in range(20):
                    graph.replay()
                reps += 20
                st.synchronize()
            e1.record(st)
            st.synchronize()
            windows.append((t0, time.perf_counter(), e0.elapsed_time(e1) * 1e3 / (reps * nsets)))
    stop = True
    th.join()
    for (t0, t1, us) in windows:
        ss = [s for s in samples if t0 <= s[0] <= t1 and s[1] > 0]
        if ss:
            clk = sorted(s[1] for s in ss)[len(ss) // 2]
            mem = sorted(s[2] for s in ss)[len(ss) // 2]
            pw = max(s[3] for s in ss)
            rs = sorted({s[4] for s in ss})
        else:
            clk = mem = pw = -1
            rs = []
        print(json.dumps({"M": args.m, "K": K, "N": N, "us_per_launch": round(us, 2), "sm_mhz": clk,
                          "mem_mhz": mem, "power_w_max": round(pw, 1), "reasons": rs,
                          "samples": len(ss)}))


if __name__ == "__main__":
    main()
