"""-m gpu parity tests (through the C ABI) of the dequantize and layout-conversion kernels:
bit-exact against the CPU oracle (SURVEY.md 8c acceptance rule)."""
import numpy as np
import pytest
import torch

from gpu_util import bits, to_dev

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("K,N,G", [(128, 64, 128), (128, 8, 32), (256, 192, 64), (512, 4096, 128),
                                   (4096, 4096, 128), (1024, 11008, 128), (384, 128, 384),
                                   (36, 24, 12)])
def test_awq_dequantize_bit_exact(pkg, oracle, cuda_lib, K, N, G):
    _, qw, s, qz = oracle.make_inputs(1, K, N, G, seed=K ^ N ^ G, signed_scales=True)
    W = pkg.ops.awq_dequantize(to_dev(qw), to_dev(s), to_dev(qz), 0, 0, 0)
    assert W.shape == (K, N) and W.dtype == torch.float16
    assert np.array_equal(bits(W), oracle.dequantize(qw, s, qz).view(np.uint16))


@pytest.mark.parametrize("K,N", [(128, 128), (128, 64), (384, 192), (4096, 4096), (11008, 4096),
                                 (4096, 22016)])
def test_repack_matches_oracle_and_roundtrips(pkg, oracle, cuda_lib, K, N):
    rng = np.random.default_rng(K + N)
    qw = rng.integers(-2**31, 2**31, size=(K, N // 8), dtype=np.int64).astype(np.int32)
    d = to_dev(qw)
    r = pkg.ops.quick_repack(d)
    assert r.numel() == oracle.b200_words(K, N)
    assert np.array_equal(r.cpu().numpy().reshape(-1), oracle.b200_repack(qw))
    back = pkg.ops.quick_unrepack(r, K, N)
    assert torch.equal(back, d)


@pytest.mark.parametrize("K,N,G", [(128, 64, 128), (256, 192, 64), (256, 128, 32), (4096, 4096, 128),
                                   (512, 256, 256)])
def test_quick_dequantize_bit_exact(pkg, oracle, cuda_lib, K, N, G):
    _, qw, s, qz = oracle.make_inputs(1, K, N, G, seed=7 + K + N + G, signed_scales=True)
    r = pkg.ops.quick_repack(to_dev(qw))
    W = pkg.ops.quick_dequantize(r, to_dev(s), to_dev(qz))
    assert np.array_equal(bits(W), oracle.dequantize(qw, s, qz).view(np.uint16))


def test_dequantize_every_nibble_zero_scale_combination(pkg, oracle, cuda_lib):
    # all 16 x 16 (q, z) pairs in every nibble slot, odd scale mantissas and both signs
    K, N = 256, 128
    q = (np.arange(K)[:, None] + np.arange(N)[None, :]) % 16
    z = (np.arange(N)[None, :] // 8 + np.arange(2)[:, None] * 5) % 16
    qw, qz = oracle.pack(q.astype(np.uint8)), oracle.pack(z.astype(np.uint8))
    s = (np.linspace(-0.0313, 0.0291, 2 * N).reshape(2, N)).astype(np.float16)
    W = pkg.ops.awq_dequantize(to_dev(qw), to_dev(s), to_dev(qz), 0, 0, 0)
    assert np.array_equal(bits(W), oracle.dequantize(qw, s, qz).view(np.uint16))
    Wq = pkg.ops.quick_dequantize(pkg.ops.quick_repack(to_dev(qw)), to_dev(s), to_dev(qz))
    assert torch.equal(W, Wq)


def test_large_roundtrip_property(pkg, cuda_lib):
    # BASELINE full size (70B gate_up, one rank of TP=2): repack is a bijection
    K, N = 8192, 28672
    g = torch.Generator(device="cuda").manual_seed(0)
    qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device="cuda", generator=g)
    r = pkg.ops.quick_repack(qw)
    assert not torch.equal(r.view(-1), qw.view(-1))
    assert torch.equal(pkg.ops.quick_unrepack(r, K, N), qw)
    # nibble histogram is preserved
    def hist(t):
        return torch.stack([((t >> (4 * i)) & 0xF).view(-1).bincount(minlength=16) for i in range(8)]).sum(0)
    assert torch.equal(hist(r), hist(qw))
