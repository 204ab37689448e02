"""-m gpu tests of the NVLink peer-memory all-reduce (csrc/w4a16_allreduce.cu).

One GPU: the two ranks' regions both live on cuda:0, rank 1's slot is pre-filled and the
kernel runs without the cross-GPU wait (``barrier=0``) -- the documented way to exercise a
multi-rank kernel without co-scheduling waiting kernels on one device.  Two or more GPUs:
real ranks under ``torch.multiprocessing`` with NCCL for the control plane, compared with
``torch.distributed.all_reduce`` and checked bit-identical across ranks and CUDA-graph replays."""
import ctypes
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu
HEADER = 16384


def _expected(parts):
    acc = torch.zeros_like(parts[0], dtype=torch.float32)
    for p in parts:                      # rank order, fp32 accumulate, one rounding
        acc += p.float()
    return acc.half()


@pytest.mark.parametrize("world,n", [(2, 16 * 8192), (4, 8 * 4096), (8, 1 * 8192), (2, 8)])
def test_single_gpu_emulation_matches_rank_order_sum(pkg, cuda_lib, world, n):
    L = cuda_lib
    max_bytes = 2 * 16 * 8192
    slot = (max_bytes + 255) // 256 * 256
    assert L.b200_w4a16_ar_region_bytes(max_bytes, world, 0) == HEADER + 2 * slot
    assert L.b200_w4a16_ar_region_bytes(max_bytes, world, 8192) == HEADER + 2 * slot + 2 * world * 8192 * 64
    regions = [torch.zeros(HEADER + 2 * slot, dtype=torch.uint8, device="cuda") for _ in range(world)]
    ptrs = (ctypes.c_void_p * world)(*[r.data_ptr() for r in regions])
    g = torch.Generator(device="cuda").manual_seed(n + world)
    for call in range(3):                # epochs 1, 2, 3 -> slots 1, 0, 1
        parts = [torch.randn(n, device="cuda", generator=g).half() for _ in range(world)]
        s = (call + 1) & 1
        for q in range(1, world):        # what the peers' kernels would have staged
            regions[q][HEADER + s * slot: HEADER + s * slot + 2 * n] = parts[q].view(torch.uint8)
        out = torch.empty_like(parts[0])
        st = L.b200_w4a16_allreduce(ptrs, 0, world, parts[0].data_ptr(), out.data_ptr(), n, max_bytes, 0,
                                    torch.cuda.current_stream().cuda_stream)
        assert st == 0
        torch.cuda.synchronize()
        assert torch.equal(out, _expected(parts)), call


@pytest.mark.parametrize("world,n", [(2, 16 * 8192), (8, 8192), (3, 8)])
def test_packet_allreduce_single_gpu_emulation(pkg, cuda_lib, world, n):
    """b200_w4a16_allreduce_ll, rank 0 of `world` on one GPU: peers' packets {data, epoch, data,
    epoch} pre-placed; result = rank-order fp32 sum, and rank 0's packets land in the peers'."""
    L = cuda_lib
    ll_bytes, ll_off = 2 * 16 * 8192, HEADER
    regions = [torch.zeros(ll_off + 4 * world * ll_bytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
    ptrs = (ctypes.c_void_p * world)(*[r.data_ptr() for r in regions])
    g = torch.Generator(device="cuda").manual_seed(n + world)
    s = torch.cuda.current_stream().cuda_stream
    for epoch in (1, 2, 3):
        parts = [torch.randn(n, device="cuda", generator=g).half() for _ in range(world)]
        base = ll_off + (epoch & 1) * world * 2 * ll_bytes
        for q in range(1, world):
            data = parts[q].view(torch.int32)
            pk = torch.stack([data, torch.full_like(data, epoch)], dim=1).reshape(-1)
            regions[0][base + q * 2 * ll_bytes: base + q * 2 * ll_bytes + 4 * n].view(torch.int32).copy_(pk)
        out = torch.empty_like(parts[0])
        assert L.b200_w4a16_allreduce_ll(ptrs, 0, world, parts[0].data_ptr(), out.data_ptr(), n, ll_off,
                                         ll_bytes, 0, s) == 0
        torch.cuda.synchronize()
        assert torch.equal(out, _expected(parts)), epoch
        mine = parts[0].view(torch.int32)
        want = torch.stack([mine, torch.full_like(mine, epoch)], dim=1).reshape(-1)
        for q in range(1, world):
            assert torch.equal(regions[q][base: base + 4 * n].view(torch.int32), want)
    assert L.b200_w4a16_allreduce_ll(ptrs, 0, world, parts[0].data_ptr(), out.data_ptr(), n, ll_off, n,
                                     0, s) == pkg._lib.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_allreduce_ll(ptrs, 0, 1, parts[0].data_ptr(), out.data_ptr(), n, ll_off, ll_bytes,
                                     0, s) == pkg._lib.B200_ERR_BAD_ARG


def _to_packets(Y, MT, epoch):
    """[16, N] fp16 -> the fused path's packet slot of one source rank: per 128-channel tile,
    2 MT x 128 packets {half2, epoch, half2, epoch}; packet e = (b MT + i) 128 + ct of thread
    ct = 32 w + 4 g + t holds tokens 8i+2t, 8i+2t+1 of channels 32 w + 16 b + g and + 8
    (push_packet in csrc/w4a16_decode.cu)."""
    N = Y.shape[1]
    data = np.zeros((N // 128, 2 * MT, 128, 2, 2), dtype=np.float16)
    for w in range(4):
        for g in range(8):
            for t in range(4):
                ct = 32 * w + 4 * g + t
                for i in range(MT):
                    for b in range(2):
                        for r in range(2):
                            ch = 32 * w + 16 * b + 8 * r + g
                            data[:, b * MT + i, ct, r, 0] = Y[8 * i + 2 * t, ch::128]
                            data[:, b * MT + i, ct, r, 1] = Y[8 * i + 2 * t + 1, ch::128]
    words = data.view(np.uint32).reshape(N // 128, 2 * MT, 128, 2)
    pk = np.empty((N // 128, 2 * MT, 128, 4), dtype=np.uint32)
    pk[..., 0] = words[..., 0]
    pk[..., 2] = words[..., 1]
    pk[..., 1] = epoch
    pk[..., 3] = epoch
    return pk


@pytest.mark.parametrize("M,K,N", [(16, 1024, 512), (5, 512, 256), (16, 4096, 4096), (16, 1024, 8192), (1, 3584, 8192)])
def test_fused_gemm_allreduce_single_gpu_emulation(pkg, oracle, cuda_lib, M, K, N):
    """Rank 0 of a world of two, both regions on cuda:0: rank 1's packets are pre-placed, so the
    kernel never waits on another launch.  Result must equal
    fp16(float(fp16(y0)) + float(y1)) with y0 the unfused GEMM of the same inputs."""
    from gpu_util import to_dev
    L = cuda_lib
    world, G, max_bytes, fused_n = 2, 128, 16 * 8192 * 2, 8192
    region_bytes = L.b200_w4a16_ar_region_bytes(max_bytes, world, fused_n)
    regions = [torch.zeros(region_bytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
    ptrs = (ctypes.c_void_p * world)(*[r.data_ptr() for r in regions])
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=N + M)
    dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(to_dev(qw))
    y0 = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    MT = 1 if M <= 8 else 2
    data_off = HEADER + 2 * ((max_bytes + 255) // 256 * 256)
    slot = fused_n * 64
    ws = torch.empty(L.b200_w4a16_workspace_bytes(M, K, N, 1), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    assert L.b200_w4a16_workspace_init(ws.data_ptr(), ws.numel(), stream) == 0
    rng = np.random.default_rng(M)
    for epoch in (1, 2, 3):
        Y1 = np.zeros((16, N), dtype=np.float16)
        Y1[:M] = rng.standard_normal((M, N)).astype(np.float16)
        par = epoch & 1
        pk = torch.from_numpy(_to_packets(Y1, MT, epoch).view(np.uint8).reshape(-1)).cuda()
        off = data_off + par * world * slot + 1 * slot
        regions[0][off: off + pk.numel()] = pk
        y = torch.empty((M, N), dtype=torch.float16, device="cuda")
        st = L.b200_w4a16_gemm_allreduce(dx.data_ptr(), r.data_ptr(), ds.data_ptr(), dqz.data_ptr(),
                                         y.data_ptr(), M, K, N, G, 0, ws.data_ptr(), ws.numel(), ptrs,
                                         0, world, max_bytes, fused_n, stream)
        assert st == 0
        torch.cuda.synchronize()
        want = (y0.float() + torch.from_numpy(Y1[:M].astype(np.float32)).cuda()).half()
        assert torch.equal(y, want), epoch
        # the packets it pushed to the "peer" (and to itself) carry its own fp16 GEMM output
        mine = _to_packets(np.pad(y0.cpu().numpy(), ((0, 16 - M), (0, 0))), MT, epoch).view(np.uint8).reshape(-1)
        for q in range(world):
            got = regions[q][data_off + par * world * slot: data_off + par * world * slot + mine.size]
            assert np.array_equal(got.cpu().numpy(), mine), (epoch, q)
        assert int(regions[0][3072:3076].view(torch.int32).item()) == epoch     # call closed
        assert bool((ws == 0xFF).all())                                         # scratch handed back idle


@pytest.mark.parametrize("world,n", [(4, 64 * 8192), (8, 512 * 8192 + 8), (3, 8), (2, 40)])
def test_twoshot_allreduce_single_gpu_emulation(pkg, cuda_lib, world, n):
    """b200_w4a16_allreduce_twoshot with every rank's region on one GPU: first all ranks run
    stage + reduce-scatter + push (mode 2, no cross-GPU waits), then all ranks run the final copy
    (mode 3).  Every rank must end with the rank-order fp32 sum, bit for bit; two calls to cover
    both slot parities."""
    L = cuda_lib
    max_bytes = (2 * n + 255) // 256 * 256
    regions = [torch.zeros(HEADER + 2 * max_bytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
    ptrs = (ctypes.c_void_p * world)(*[r.data_ptr() for r in regions])
    g = torch.Generator(device="cuda").manual_seed(n + world)
    s = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        parts = [torch.randn(n, device="cuda", generator=g).half() for _ in range(world)]
        outs = [torch.empty_like(p) for p in parts]
        # all ranks stage first (what barrier A guarantees), then reduce + push: run mode 2 twice
        # is not possible (the epoch would advance), so stage by hand for the peers of each rank
        slot = HEADER + ((call + 1) & 1) * max_bytes
        for q in range(world):
            regions[q][slot: slot + 2 * n] = parts[q].view(torch.uint8)
        for r in range(world):
            assert L.b200_w4a16_allreduce_twoshot(ptrs, r, world, parts[r].data_ptr(), outs[r].data_ptr(), n,
                                                  max_bytes, 2, s) == 0
        for r in range(world):
            assert L.b200_w4a16_allreduce_twoshot(ptrs, r, world, parts[r].data_ptr(), outs[r].data_ptr(), n,
                                                  max_bytes, 3, s) == 0
        torch.cuda.synchronize()
        want = _expected(parts)
        for r in range(world):
            assert torch.equal(outs[r], want), (call, r)
    x = torch.zeros(64, dtype=torch.float16, device="cuda")
    assert L.b200_w4a16_allreduce_twoshot(ptrs, 0, world, x.data_ptr(), x.data_ptr(), 12, max_bytes, 1, s) == pkg._lib.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_allreduce_twoshot(ptrs, 0, world, x.data_ptr(), x.data_ptr(), 64, max_bytes, 0, s) == pkg._lib.B200_ERR_BAD_ARG
    assert L.b200_w4a16_allreduce_twoshot(ptrs, 0, 1, x.data_ptr(), x.data_ptr(), 64, max_bytes, 1, s) == pkg._lib.B200_ERR_BAD_ARG


def test_allreduce_rejects_bad_arguments(pkg, cuda_lib):
    L = cuda_lib
    r = torch.zeros(HEADER + 1024, dtype=torch.uint8, device="cuda")
    ptrs = (ctypes.c_void_p * 1)(r.data_ptr())
    x = torch.zeros(64, dtype=torch.float16, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    assert L.b200_w4a16_allreduce(ptrs, 0, 1, x.data_ptr(), x.data_ptr(), 12, 512, 1, s) == pkg._lib.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_allreduce(ptrs, 0, 1, x.data_ptr(), x.data_ptr(), 512, 512, 1, s) == pkg._lib.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_allreduce(ptrs, 1, 1, x.data_ptr(), x.data_ptr(), 64, 512, 1, s) == pkg._lib.B200_ERR_BAD_ARG
    assert L.b200_w4a16_allreduce(ptrs, 0, 9, x.data_ptr(), x.data_ptr(), 64, 512, 1, s) == pkg._lib.B200_ERR_BAD_ARG
    assert L.b200_w4a16_allreduce(ptrs, 0, 1, x.data_ptr(), x.data_ptr(), 0, 512, 1, s) == 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, q):
    try:
        import torch.distributed as dist
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        pkg = importlib.import_module("vllm-quick_b200")
        pu = pkg.parallel_utils
        pu.initialize_model_parallel(world)
        pu.custom_all_reduce.init_custom_all_reduce(max_bytes=256 * 8192 * 2, fused_out_features=8192)
        assert pu.custom_all_reduce._handle.oneshot_max_bytes in (128 << 10, 512 << 10)
        pu.custom_all_reduce._handle.oneshot_max_bytes = 1 << 30     # exercise the kernel at every size
        assert pu.custom_all_reduce.fused_handle(16, 4096, 128) is not None   # the fused GEMM kernel at every TP
        pu.custom_all_reduce._handle.prefer_fused_gemm = False
        assert pu.custom_all_reduce.fused_handle(16, 4096, 128) is None       # ... unless packets are asked for
        pu.custom_all_reduce._handle.prefer_fused_gemm = True
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        for shape in [(1, 8192), (16, 8192), (256, 8192), (3, 4096)]:
            x = torch.randn(shape, device=dev, generator=g).half()
            ref = x.clone()
            dist.all_reduce(ref)
            y = pu.tensor_model_parallel_all_reduce(x)
            assert y.data_ptr() != x.data_ptr()                     # the kernel took it
            assert torch.allclose(y.float(), ref.float(), rtol=2e-3, atol=2e-2), shape
            gathered = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(gathered, y)
            assert all(torch.equal(gathered[0], t) for t in gathered)   # bit-identical on every rank
        # two-shot kernel (reduce-scatter + all-gather), forced at every world size, against the
        # one-shot kernel (same rank-order fp32 sums: same bits) and NCCL; mixed with small calls
        h = pu.custom_all_reduce._handle
        saved = (h.oneshot_max_bytes, h.use_twoshot, h.use_ll)
        h.twoshot_min_bytes = 0
        for shape in [(256, 8192), (64, 8192), (33, 1000)]:
            x = torch.randn(shape, device=dev, generator=g).half()
            h.oneshot_max_bytes, h.use_twoshot, h.use_ll = 1 << 30, False, False
            one = pu.tensor_model_parallel_all_reduce(x)
            h.oneshot_max_bytes, h.use_twoshot, h.use_ll = 0, True, False
            two = pu.tensor_model_parallel_all_reduce(x)
            small = pu.tensor_model_parallel_all_reduce(x[:1].contiguous())     # another size in between
            two_again = pu.tensor_model_parallel_all_reduce(x)
            assert two.data_ptr() != x.data_ptr() and torch.equal(one, two) and torch.equal(two, two_again), shape
            ref = x.clone()
            dist.all_reduce(ref)
            assert torch.allclose(two.float(), ref.float(), rtol=2e-3, atol=2e-2), shape
            assert torch.allclose(small.float(), ref[:1].float(), rtol=2e-3, atol=2e-2), shape
        h.oneshot_max_bytes, h.use_twoshot, h.use_ll = saved
        # row-parallel layer: the fused GEMM + all-reduce kernel vs GEMM, then all-reduce kernel
        K, N = 2048 * world, 4096
        method = pkg.QuickConfig(4, 128, True).get_linear_method()
        row = pkg.RowParallelLinear(K, N, bias=False, input_is_parallel=True, linear_method=method,
                                    params_dtype=torch.float16).to(dev)
        gw = torch.Generator(device=dev).manual_seed(7 + rank)
        w = row.linear_weights
        w["qweight"].data = torch.randint(-2**31, 2**31 - 1, tuple(w["qweight"].shape), dtype=torch.int32, device=dev, generator=gw)
        w["qzeros"].data = torch.randint(-2**31, 2**31 - 1, tuple(w["qzeros"].shape), dtype=torch.int32, device=dev, generator=gw)
        w["scales"].data = (torch.rand(tuple(w["scales"].shape), device=dev, generator=gw) * 0.019 + 0.001).half()
        for Mtok in (1, 8, 16):
            xin = torch.randn(Mtok, K // world, device=dev, generator=gw).half()
            assert pu.custom_all_reduce.fused_handle(Mtok, N, 128) is not None
            fused, _ = row(xin)
            two_step = pu.tensor_model_parallel_all_reduce(method.apply_weights(w, xin))
            assert torch.equal(fused, two_step), Mtok
            ref = method.apply_weights(w, xin).float()
            dist.all_reduce(ref)
            assert torch.allclose(fused.float(), ref, rtol=2e-3, atol=3e-2)
        xin = torch.randn(64, K // world, device=dev, generator=gw).half()       # M > 16: two kernels
        assert pu.custom_all_reduce.fused_handle(64, N, 128) is None
        out64, _ = row(xin)
        assert torch.equal(out64, pu.tensor_model_parallel_all_reduce(method.apply_weights(w, xin)))
        # fused path under CUDA-graph replay
        xin = torch.randn(16, K // world, device=dev, generator=gw).half()
        st2 = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(st2):
            eager, _ = row(xin)
            st2.synchronize()
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2, stream=st2):
                o1, _ = row(xin)
                o2, _ = row(xin)
            for _ in range(3):
                g2.replay()
            st2.synchronize()
            assert torch.equal(o1, eager) and torch.equal(o2, eager)
        big = torch.randn(2048, 8192, device=dev).half()            # above the registered size: NCCL
        assert pu.tensor_model_parallel_all_reduce(big).data_ptr() == big.data_ptr()
        # CUDA-graph capture + replay with fresh inputs
        x = torch.randn(16, 8192, device=dev, generator=g).half()
        st = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(st):
            pu.tensor_model_parallel_all_reduce(x)
            st.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=st):
                y = pu.tensor_model_parallel_all_reduce(x)
                z = pu.tensor_model_parallel_all_reduce(y)
            for it in range(3):
                x.copy_(torch.full_like(x, float(rank + it)))
                graph.replay()
                st.synchronize()
                s1 = sum(r + it for r in range(world))
                assert torch.equal(y, torch.full_like(y, float(s1)))
                assert torch.equal(z, torch.full_like(z, float(s1 * world)))
        pu.custom_all_reduce.destroy_custom_all_reduce()
        pu.destroy_model_parallel()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except BaseException as e:  # noqa: BLE001
        import traceback
        q.put((rank, "".join(traceback.format_exception(type(e), e, e.__traceback__))))


@pytest.mark.timeout(300)
def test_multi_gpu_allreduce_matches_nccl(pkg, cuda_lib):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs on the box (gpurun --gpus 2); the one-GPU emulation test above "
                    "covers the reduction arithmetic")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(30)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}: {msg}"
