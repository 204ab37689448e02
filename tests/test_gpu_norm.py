"""-m gpu parity tests of the residual add + RMSNorm kernels (csrc/w4a16_norm.cu)
through the C ABI / the ops mirror, against the oracle (oracle/oracle.py np_rms_norm_*).

Tolerances: the residual stream is exact (bit for bit); the normalised output is floating
point with a reduction, and the bar is the reference's own for this kernel, atol = rtol = 1e-2
(tests/kernels/test_layernorm.py:52-59) -- the tests below hold the kernel to 2 fp16 ulps of
the kernel-arithmetic oracle, far inside it.  Shapes and input recipe follow the reference's
test (tokens 7 / 83 / 4096, hidden 768 / 5120 / 8192, x ~ N(0,1) / (2 hidden), w ~ N(1, 0.1))."""
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


def _inputs(tokens, hidden, seed, scale=None):
    rng = np.random.default_rng(seed)
    scale = 1.0 / (2 * hidden) if scale is None else scale
    x = (rng.standard_normal((tokens, hidden)) * scale).astype(np.float16)
    r = (rng.standard_normal((tokens, hidden)) * scale).astype(np.float16)
    w = rng.normal(1.0, 0.1, hidden).astype(np.float16)
    return x, r, w


def _ulps(got: torch.Tensor, want: np.ndarray) -> float:
    g = got.detach().float().cpu().numpy()
    w = want.astype(np.float32)
    ulp = np.spacing(np.maximum(np.abs(want), np.float16(6.1e-5))).astype(np.float32)
    return float((np.abs(g - w) / ulp).max(initial=0.0))


@pytest.mark.parametrize("tokens", [1, 7, 83, 4096])
@pytest.mark.parametrize("hidden", [768, 4096, 5120, 8192, 16384])
@pytest.mark.parametrize("add_residual", [False, True])
def test_rms_norm_matches_oracle(pkg, oracle, tokens, hidden, add_residual):
    from gpu_util import bits, to_dev
    if tokens == 4096 and hidden == 16384:
        tokens = 1024
    x, r, w = _inputs(tokens, hidden, tokens * 3 + hidden)
    dx, dr, dw = to_dev(x), to_dev(r), to_dev(w)
    if add_residual:
        want, want_res = oracle.np_rms_norm_kernel(x, w, 1e-6, r)
        pkg.ops.fused_add_rms_norm(dx, dr, dw, 1e-6)                 # in place, like the reference
        assert np.array_equal(bits(dr), want_res.view(np.uint16))    # residual stream: exact
        out = dx
        fwd = oracle.np_rms_norm_forward(x, w, 1e-6, r)[0]
    else:
        want = oracle.np_rms_norm_kernel(x, w, 1e-6)
        out = torch.empty_like(dx)
        pkg.ops.rms_norm(out, dx, dw, 1e-6)
        assert np.array_equal(bits(dx), x.view(np.uint16))           # input untouched
        fwd = oracle.np_rms_norm_forward(x, w, 1e-6)
    assert _ulps(out, want) <= 2.0
    # the reference's own acceptance against its PyTorch statement (test_layernorm.py:52-59)
    assert np.allclose(out.float().cpu().numpy(), fwd.astype(np.float32), atol=1e-2, rtol=1e-2)


def test_rms_norm_layer_and_properties(pkg, oracle):
    """RMSNorm mirror; scale invariance (x -> 4x leaves the output unchanged when eps = 0, a
    power of two commutes with every rounding); determinism; unit-scale inputs."""
    from gpu_util import bits, to_dev
    x, r, w = _inputs(64, 8192, 1, scale=1.0)
    layer = pkg.RMSNorm(8192, eps=0.0).half().cuda()
    layer.weight.data.copy_(to_dev(w))
    dx = to_dev(x)
    a = layer(dx)
    b = layer(dx * 4)
    assert torch.equal(a, b)
    assert torch.equal(a, layer(dx))
    assert _ulps(a, oracle.np_rms_norm_kernel(x, w, 0.0)) <= 2.0
    out, res = layer(dx.clone(), to_dev(r))
    want, want_res = oracle.np_rms_norm_kernel(x, w, 0.0, r)
    assert np.array_equal(bits(res), want_res.view(np.uint16))
    assert _ulps(out, want) <= 2.0
    # zero rows stay zero (rsqrt(eps) * 0), empty batch is a no-op
    layer.variance_epsilon = 1e-6
    assert not layer(torch.zeros(3, 8192, dtype=torch.float16, device="cuda")).any()
    assert layer(torch.zeros(0, 8192, dtype=torch.float16, device="cuda")).shape == (0, 8192)


def test_rms_norm_rejects_bad_arguments(pkg, cuda_lib):
    L = cuda_lib
    s = torch.cuda.current_stream().cuda_stream
    x = torch.zeros(4, 64, dtype=torch.float16, device="cuda")
    w = torch.ones(64, dtype=torch.float16, device="cuda")
    e = pkg._lib
    assert L.b200_w4a16_add_rms_norm(x.data_ptr(), None, w.data_ptr(), x.data_ptr(), None, 1e-6, 4, 12, s) == e.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_add_rms_norm(x.data_ptr(), None, w.data_ptr(), x.data_ptr(), None, 1e-6, 4, 16392, s) == e.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_add_rms_norm(None, None, w.data_ptr(), x.data_ptr(), None, 1e-6, 4, 64, s) == e.B200_ERR_BAD_ARG
    assert L.b200_w4a16_add_rms_norm(x.data_ptr() + 2, None, w.data_ptr(), x.data_ptr(), None, 1e-6, 1, 8, s) == e.B200_ERR_BAD_ARG
    assert L.b200_w4a16_add_rms_norm(x.data_ptr(), None, w.data_ptr(), x.data_ptr(), None, 1e-6, 0, 64, s) == 0
    with pytest.raises(ValueError):
        pkg.ops.rms_norm(x, x, torch.ones(32, dtype=torch.float16, device="cuda"), 1e-6)
    with pytest.raises(ValueError):
        pkg.ops.rms_norm(x.float(), x.float(), w.float(), 1e-6)


def test_rms_norm_cuda_graph(pkg):
    x = torch.randn(16, 4096, device="cuda").half()
    r = torch.randn(16, 4096, device="cuda").half()
    w = torch.ones(4096, device="cuda").half()
    x0, r0 = x.clone(), r.clone()
    pkg.ops.fused_add_rms_norm(x0, r0, w, 1e-6)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        xa, ra = x.clone(), r.clone()
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            pkg.ops.fused_add_rms_norm(xa, ra, w, 1e-6)
        xa.copy_(x); ra.copy_(r)
        g.replay()
        st.synchronize()
    assert torch.equal(xa, x0) and torch.equal(ra, r0)


@pytest.mark.parametrize("world,rows,hidden", [(2, 16, 8192), (4, 5, 4096), (8, 1, 8192), (2, 40, 5120),
                                               (2, 3, 16384), (3, 100, 768)])
def test_allreduce_norm_single_gpu_emulation(pkg, oracle, cuda_lib, world, rows, hidden):
    """Rank 0 of `world`, all regions on cuda:0, peers' slots pre-filled (barrier = 0, as in
    test_gpu_allreduce.py).  Must equal, bit for bit, the all-reduce kernel followed by the
    norm kernel; the reduced residual must equal the oracle exactly."""
    from gpu_util import bits
    L = cuda_lib
    max_bytes = 2 * 128 * 8192
    slot = (max_bytes + 255) // 256 * 256
    regions = [torch.zeros(HEADER + 2 * slot, dtype=torch.uint8, device="cuda") for _ in range(world)]
    ptrs = (ctypes.c_void_p * world)(*[r.data_ptr() for r in regions])
    g = torch.Generator(device="cuda").manual_seed(rows + world)
    n = rows * hidden
    s = torch.cuda.current_stream().cuda_stream
    w = (torch.randn(hidden, device="cuda", generator=g) * 0.1 + 1).half()
    for call in range(4):                # epochs 1..4; call 2 is a plain all-reduce (mixed use)
        parts = [torch.randn(rows, hidden, device="cuda", generator=g).half() for _ in range(world)]
        res = torch.randn(rows, hidden, device="cuda", generator=g).half()
        par = (call + 1) & 1
        for q in range(1, world):
            regions[q][HEADER + par * slot: HEADER + par * slot + 2 * n] = parts[q].view(torch.uint8).view(-1)
        reduced = torch.empty_like(parts[0])
        if call == 2:
            assert L.b200_w4a16_allreduce(ptrs, 0, world, parts[0].data_ptr(), reduced.data_ptr(), n,
                                          max_bytes, 0, s) == 0
            torch.cuda.synchronize()
            want = oracle.np_allreduce_f16([p.cpu().numpy() for p in parts])
            assert np.array_equal(bits(reduced), want.view(np.uint16))
            continue
        out = torch.empty_like(parts[0])
        res_out = torch.empty_like(res)
        st = L.b200_w4a16_allreduce_add_rms_norm(ptrs, 0, world, parts[0].data_ptr(), res.data_ptr(), w.data_ptr(),
                                           out.data_ptr(), res_out.data_ptr(), 1e-5, rows, hidden,
                                           max_bytes, 0, s)
        assert st == 0
        torch.cuda.synchronize()
        # every block's epoch advanced, whatever the grid was
        epochs = regions[0][2048:2048 + 256].view(torch.int32)
        assert torch.all(epochs == call + 1), call
        red = oracle.np_allreduce_f16([p.cpu().numpy() for p in parts])
        want, want_res = oracle.np_rms_norm_kernel(red, w.cpu().numpy(), 1e-5, res.cpu().numpy())
        assert np.array_equal(bits(res_out), want_res.view(np.uint16)), call
        assert _ulps(out, want) <= 2.0
        # two-step on the device: identical bits
        two = torch.from_numpy(red.view(np.int16).copy()).view(torch.float16).cuda()
        r2 = res.clone()
        pkg.ops.fused_add_rms_norm(two, r2, w, 1e-5)
        assert torch.equal(out, two) and torch.equal(res_out, r2), call


@pytest.mark.parametrize("world,rows,hidden", [(2, 16, 8192), (8, 1, 8192), (4, 40, 5120), (3, 70, 768)])
def test_allreduce_norm_packets_single_gpu_emulation(pkg, oracle, cuda_lib, world, rows, hidden):
    """The flag-in-data variant, rank 0 of `world` on one GPU: the peers' packets
    {data, epoch, data, epoch} are pre-placed in rank 0's packet area; what rank 0 pushes lands
    in the "peers'" regions and is checked too.  Result must equal the barrier variant's."""
    from gpu_util import bits
    L = cuda_lib
    ll_bytes = 2 * 128 * 8192
    ll_off = HEADER + 4096
    region_bytes = ll_off + 4 * world * ll_bytes
    regions = [torch.zeros(region_bytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
    ptrs = (ctypes.c_void_p * world)(*[r.data_ptr() for r in regions])
    g = torch.Generator(device="cuda").manual_seed(rows + world)
    n = rows * hidden
    s = torch.cuda.current_stream().cuda_stream
    w = (torch.randn(hidden, device="cuda", generator=g) * 0.1 + 1).half()
    for call in range(3):
        epoch = call + 1
        parts = [torch.randn(rows, hidden, device="cuda", generator=g).half() for _ in range(world)]
        res = torch.randn(rows, hidden, device="cuda", generator=g).half()
        base = ll_off + (epoch & 1) * world * 2 * ll_bytes
        for q in range(1, world):
            data = parts[q].view(torch.int32).reshape(-1)                       # 2 halfs per word
            pk = torch.stack([data, torch.full_like(data, epoch)], dim=1).reshape(-1)
            off = base + q * 2 * ll_bytes
            regions[0][off: off + 4 * n].view(torch.int32).copy_(pk)
        out = torch.empty_like(parts[0])
        res_out = torch.empty_like(res)
        st = L.b200_w4a16_allreduce_add_rms_norm_ll(ptrs, 0, world, parts[0].data_ptr(), res.data_ptr(),
                                                    w.data_ptr(), out.data_ptr(), res_out.data_ptr(), 1e-5,
                                                    rows, hidden, ll_off, ll_bytes, 0, s)
        assert st == 0
        torch.cuda.synchronize()
        red = oracle.np_allreduce_f16([p.cpu().numpy() for p in parts])
        two = torch.from_numpy(red.view(np.int16).copy()).view(torch.float16).cuda()
        r2 = res.clone()
        pkg.ops.fused_add_rms_norm(two, r2, w, 1e-5)
        assert torch.equal(out, two) and torch.equal(res_out, r2), call
        # rank 0's own packets arrived in every peer's area, source slot 0
        mine = parts[0].view(torch.int32).reshape(-1)
        want_pk = torch.stack([mine, torch.full_like(mine, epoch)], dim=1).reshape(-1)
        for q in range(1, world):
            got = regions[q][base: base + 4 * n].view(torch.int32)
            assert torch.equal(got, want_pk), (call, q)
        assert torch.all(regions[0][2048:2048 + 256].view(torch.int32) == epoch)
    e = pkg._lib
    assert L.b200_w4a16_allreduce_add_rms_norm_ll(ptrs, 0, world, parts[0].data_ptr(), None, w.data_ptr(),
                                                  out.data_ptr(), None, 1e-5, rows, hidden, ll_off, hidden, 0,
                                                  s) == e.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_allreduce_add_rms_norm_ll(ptrs, 0, world, parts[0].data_ptr(), None, w.data_ptr(),
                                                  out.data_ptr(), None, 1e-5, rows, hidden, ll_off + 16, ll_bytes,
                                                  0, s) == e.B200_ERR_BAD_ARG


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
        pu.custom_all_reduce.init_custom_all_reduce(max_bytes=256 * 8192 * 2, ll_bytes=256 * 8192 * 2)
        pu.custom_all_reduce._handle.oneshot_max_bytes = 1 << 30     # exercise the kernel at every size
        pu.custom_all_reduce._handle.ll_max_bytes = 1 << 30
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        gs = torch.Generator(device=dev).manual_seed(5)                 # same on every rank
        for rows, hidden in [(1, 8192), (16, 8192), (256, 8192), (3, 4096), (16, 5120), (70, 768)]:
            norm = pkg.RMSNorm(hidden, eps=1e-5).half().to(dev)
            norm.weight.data = (torch.randn(hidden, device=dev, generator=gs) * 0.1 + 1).half()
            partial = torch.randn(rows, hidden, device=dev, generator=g).half()
            residual = torch.randn(rows, hidden, device=dev, generator=gs).half()
            r_two = residual.clone()
            two, r_two = norm(pu.tensor_model_parallel_all_reduce(partial), r_two)
            r_one = residual.clone()
            one, r_one = norm.forward_reduced(partial, r_one)        # dispatch (fused up to 32 rows)
            assert torch.equal(one, two) and torch.equal(r_one, r_two), (rows, hidden)
            for use_ll in (True, False):                             # the fused kernels themselves, any rows
                pu.custom_all_reduce._handle.use_ll = use_ll
                r_k = residual.clone()
                k = pu.custom_all_reduce._handle.all_reduce_add_rms_norm(partial, r_k, norm.weight.data, 1e-5)
                assert torch.equal(k, two) and torch.equal(r_k, r_two), (rows, hidden, use_ll)
            pu.custom_all_reduce._handle.use_ll = True
            ref = partial.float()
            dist.all_reduce(ref)
            assert torch.allclose(r_one.float(), ref + residual.float(), rtol=2e-3, atol=2e-2)
            gathered = [torch.empty_like(one) for _ in range(world)]
            dist.all_gather(gathered, one)
            assert all(torch.equal(gathered[0], t) for t in gathered)   # bit-identical on every rank
        # mixed with plain all-reduces of other sizes, under CUDA-graph replay
        norm = pkg.RMSNorm(8192, eps=1e-5).half().to(dev)
        partial = torch.randn(16, 8192, device=dev, generator=g).half()
        small = torch.randn(2, 4096, device=dev, generator=g).half()
        residual = torch.randn(16, 8192, device=dev, generator=gs).half()
        want_small = pu.tensor_model_parallel_all_reduce(small)
        want, want_res = norm.forward_reduced(partial, residual.clone())
        st = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(st):
            res_g = residual.clone()
            st.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=st):
                s1 = pu.tensor_model_parallel_all_reduce(small)
                o1, _ = norm.forward_reduced(partial, res_g)
                s2 = pu.tensor_model_parallel_all_reduce(small)
            for _ in range(3):
                res_g.copy_(residual)
                graph.replay()
                st.synchronize()
                assert torch.equal(o1, want) and torch.equal(res_g, want_res)
                assert torch.equal(s1, want_small) and torch.equal(s2, want_small)
        pu.custom_all_reduce.destroy_custom_all_reduce()
        pu.destroy_model_parallel()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except BaseException as e:  # noqa: BLE001
        import traceback
        q.put((rank, "".join(traceback.format_exception(type(e), e, e.__traceback__))))


@pytest.mark.timeout(300)
def test_multi_gpu_allreduce_norm_matches_two_step(pkg, cuda_lib):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs on the box (gpurun --gpus 2); the one-GPU emulation test above "
                    "covers the arithmetic")
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
