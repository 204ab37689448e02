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

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu
HEADER = 4096


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
    assert L.b200_w4a16_ar_region_bytes(max_bytes) == HEADER + 2 * slot
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
        pu.custom_all_reduce.init_custom_all_reduce(max_bytes=256 * 8192 * 2)
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
