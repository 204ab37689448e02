"""CPU tests of the drop-in boundary: the library loads, exports every symbol the header
declares, validates shapes like the reference host code
(csrc/quantization/awq/gemm_kernels.cu:412-419) and refuses to compute without a B200."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for fn in os.listdir(inc):
        if fn.endswith(".h"):
            src = open(os.path.join(inc, fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(b200_w4a16_\w+)\s*\(", src))
    return names


def test_library_exports_every_declared_symbol(pkg, cuda_lib):
    declared = _declared_symbols()
    assert declared == set(pkg._lib.EXPORTS)
    for name in declared:
        assert hasattr(cuda_lib, name), name


def test_version_and_strerror(cuda_lib):
    assert cuda_lib.b200_w4a16_version() == 200
    assert cuda_lib.b200_w4a16_strerror(0) == b"ok"
    assert b"shape" in cuda_lib.b200_w4a16_strerror(1)
    assert b"no sm_100" in cuda_lib.b200_w4a16_strerror(5)


@pytest.mark.parametrize("K,N,words", [(128, 128, 2048), (4096, 4096, 4096 * 512),
                                       (256, 192, 2 * 2 * 2048), (128, 64, 2048)])
def test_repacked_words(cuda_lib, K, N, words):
    assert cuda_lib.b200_w4a16_repacked_words(K, N) == words


@pytest.mark.parametrize("M,K,N,G", [
    (16, 4096, 4000, 128),   # N % 64 != 0
    (16, 4096, 4096, 48),    # G % 32 != 0
    (16, 4100, 4096, 128),   # K % G != 0
    (16, 4096 + 64, 4096, 64),  # K % 128 != 0
    (-1, 4096, 4096, 128),
    (16, 4096, 4096, 96),    # G not in {32, 64, 128k}
])
def test_gemm_rejects_bad_shapes(pkg, cuda_lib, M, K, N, G):
    fake = ctypes.c_void_p(0x1000)
    st = cuda_lib.b200_w4a16_gemm(fake, fake, fake, fake, fake, M, K, N, G, 1, 8, fake, 1 << 30,
                                  None)
    assert st == pkg._lib.B200_ERR_BAD_SHAPE
    with pytest.raises(ValueError):
        pkg._lib.check(st, "gemm")


def test_gemm_rejects_bad_args(pkg, cuda_lib):
    fake = ctypes.c_void_p(0x1000)
    assert cuda_lib.b200_w4a16_gemm(None, fake, fake, fake, fake, 16, 128, 128, 128, 1, 8, fake,
                                    1 << 30, None) == pkg._lib.B200_ERR_BAD_ARG
    assert cuda_lib.b200_w4a16_gemm(fake, fake, fake, fake, fake, 16, 128, 128, 128, 7, 8, fake,
                                    1 << 30, None) == pkg._lib.B200_ERR_BAD_ARG
    mis = ctypes.c_void_p(0x1004)
    assert cuda_lib.b200_w4a16_gemm(mis, fake, fake, fake, fake, 16, 128, 128, 128, 1, 8, fake,
                                    1 << 30, None) == pkg._lib.B200_ERR_BAD_ARG
    # M == 0 is a no-op that needs no device
    assert cuda_lib.b200_w4a16_gemm(None, None, None, None, None, 0, 128, 128, 128, 1, 8, None, 0,
                                    None) == pkg._lib.B200_OK


def test_dequant_and_repack_validation(pkg, cuda_lib):
    fake = ctypes.c_void_p(0x1000)
    L = pkg._lib
    assert cuda_lib.b200_w4a16_dequant(fake, fake, fake, fake, 128, 60, 128, 0, None) == L.B200_ERR_BAD_SHAPE
    assert cuda_lib.b200_w4a16_dequant(fake, fake, fake, fake, 100, 64, 128, 0, None) == L.B200_ERR_BAD_SHAPE
    assert cuda_lib.b200_w4a16_dequant(fake, fake, fake, fake, 64, 64, 64, 1, None) == L.B200_ERR_BAD_SHAPE
    assert cuda_lib.b200_w4a16_dequant(fake, fake, fake, fake, 128, 64, 128, 5, None) == L.B200_ERR_BAD_ARG
    assert cuda_lib.b200_w4a16_dequant(None, None, None, None, 0, 64, 128, 0, None) == L.B200_OK
    assert cuda_lib.b200_w4a16_repack(fake, fake, 100, 64, 0, None) == L.B200_ERR_BAD_SHAPE
    assert cuda_lib.b200_w4a16_repack(fake, ctypes.c_void_p(0x2000), 128, 64, 3, None) == L.B200_ERR_BAD_ARG
    assert cuda_lib.b200_w4a16_repack(fake, fake, 128, 64, 0, None) == L.B200_ERR_BAD_ARG  # overlap


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device refusal")
def test_no_device_means_no_compute(pkg, cuda_lib):
    fake = ctypes.c_void_p(0x1000)
    st = cuda_lib.b200_w4a16_gemm(fake, fake, fake, fake, fake, 16, 128, 128, 128, 1, 8, fake,
                                  1 << 30, None)
    assert st == pkg._lib.B200_ERR_NO_DEVICE
    assert cuda_lib.b200_w4a16_dequant(fake, fake, fake, fake, 128, 64, 128, 0, None) == pkg._lib.B200_ERR_NO_DEVICE
    with pytest.raises(RuntimeError):
        pkg._lib.check(st, "gemm")
    # the torch-facing ops refuse CPU tensors instead of falling back
    x = torch.zeros(2, 128, dtype=torch.float16)
    qw = torch.zeros(128, 16, dtype=torch.int32)
    s = torch.zeros(1, 128, dtype=torch.float16)
    qz = torch.zeros(1, 16, dtype=torch.int32)
    with pytest.raises(RuntimeError):
        pkg.ops.awq_gemm(x, qw, s, qz, 8)
    with pytest.raises(RuntimeError):
        pkg.ops.awq_dequantize(qw, s, qz, 0, 0, 0)


@pytest.mark.parametrize("M,K,N,tok,grid", [
    # M <= 16: mma.sync decode kernel, one persistent CTA per SM, 8- or 16-token tiles
    (1, 4096, 22016, 8, 148), (16, 4096, 12288, 16, 148), (3, 128, 64, 8, 1), (16, 256, 128, 16, 2),
    # ... on short walks with ranges that end at tile borders: 32 channel tiles x 4 CTAs
    (1, 4096, 4096, 8, 128), (16, 8192, 1280, 16, 140), (16, 8192, 8192, 16, 148),
    # above: tcgen05 stream-K kernel, one persistent CTA per SM
    (2048, 5120, 15360, 128, 148), (8192, 5120, 15360, 128, 148),
    # fewer output units than SMs: a whole number of CTAs per unit when that keeps >= 64 % of the SMs
    # busy (32 units x 4, 64 x 2, 96 x 1), at most four per unit from 32-token tiles up (10 units x 4)
    (17, 4096, 4096, 32, 128), (64, 11008 // 128 * 128, 4096, 64, 128), (256, 4096, 4096, 128, 128),
    (64, 4096, 12288, 64, 96), (64, 8192, 1280, 64, 40), (32, 8192, 1280, 32, 40),
    (512, 4096, 4096, 128, 128), (40, 4096, 22016, 64, 148),
])
def test_gemm_plan(pkg, M, K, N, tok, grid):
    plan = pkg._lib.gemm_plan(M, K, N)
    assert plan["tok_tile"] == tok
    assert plan["grid"] == grid
    assert plan["smem_bytes"] <= 227 * 1024
    ws = pkg._lib.lib().b200_w4a16_workspace_bytes(M, K, N, 1)
    assert 256 <= ws <= 148 * 2 * 2 * 128 * 128 * 4
    assert pkg._lib.lib().b200_w4a16_workspace_bytes(M, K, N, 0) >= ws + K * N // 2


def test_gemm_plan_invariants_over_the_benchmark_shapes(pkg):
    """Planner invariants the kernels rely on, over every (M, K, N) the benchmarks and BASELINE
    configs use: every CTA resident (grid <= 148), CTA pairs only with an even grid and 128-token
    tiles, shared memory within the opt-in limit, split-K scratch within the base workspace."""
    L = pkg._lib.lib()
    shapes = [(4096, 4096), (4096, 12288), (4096, 22016), (11008, 4096),            # Llama-2-7B
              (5120, 15360), (5120, 5120), (5120, 27648), (13824, 5120),             # 13B
              (8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192),             # 70B TP1
              (8192, 1280), (1024, 8192), (8192, 7168), (3584, 8192),                # 70B TP8
              (4096, 28672), (14336, 4096), (4096, 768), (512, 4096)]                # Mixtral
    base = 148 * 2 * 2 * 128 * 128 * 4
    for (K, N) in shapes:
        for M in (1, 8, 16, 17, 64, 128, 256, 512, 2048, 4096, 8192):
            plan = pkg._lib.gemm_plan(M, K, N)
            assert 1 <= plan["grid"] <= 148 and plan["smem_bytes"] <= 227 * 1024, (M, K, N, plan)
            assert plan["variant"] == ("decode" if M <= 16 else "tcgen05"), (M, K, N, plan)
            if plan["cta_pair"]:
                assert plan["grid"] % 2 == 0 and plan["tok_tile"] == 128, (M, K, N, plan)
                assert plan["unit_tiles"] in (2, 4) and plan["token_tiles"] in (1, 2), (M, K, N, plan)
            else:
                assert plan["unit_tiles"] in (1, 2) and plan["token_tiles"] == 1, (M, K, N, plan)
            if plan["tok_tile"] < 128:
                assert plan["unit_tiles"] == 1, (M, K, N, plan)      # pairs only for 128-token tiles
            assert 256 <= L.b200_w4a16_workspace_bytes(M, K, N, 1) <= base, (M, K, N)
