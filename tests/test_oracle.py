"""CPU tests of the oracle itself (oracle/ is test infrastructure): the C restatement, the
NumPy restatement and the bit-level emulation of the reference's PTX sequence
(csrc/quantization/awq/dequantize.cuh:17-84) must agree with each other."""
import numpy as np
import pytest


def test_lop3_emulation_matches_shift_unpack_all_halfwords(oracle):
    # Every 16-bit pattern in the low and in the high half covers every nibble combination
    # each lop3 extraction can see.
    for hi in (0x0000, 0xFFFF, 0x1234, 0xA5C3):
        for lo in range(0, 1 << 16, 7):  # stride 7 still hits every nibble value in every slot
            w = (hi << 16) | lo
            got = oracle.dequant_word_lop3(w)
            exp = np.array([(w >> (4 * o)) & 0xF for o in oracle.AWQ_ORDER], dtype=np.float16)
            assert np.array_equal(got, exp), hex(w)
    for lo in (0x0000, 0xFFFF, 0x89AB):
        for hi in range(0, 1 << 16, 11):
            w = (hi << 16) | lo
            got = oracle.dequant_word_lop3(w)
            exp = np.array([(w >> (4 * o)) & 0xF for o in oracle.AWQ_ORDER], dtype=np.float16)
            assert np.array_equal(got, exp), hex(w)


def test_pack_unpack_roundtrip(oracle):
    rng = np.random.default_rng(0)
    q = rng.integers(0, 16, size=(64, 256), dtype=np.uint8)
    packed = oracle.pack(q)
    assert packed.shape == (64, 32) and packed.dtype == np.int32
    assert np.array_equal(oracle.unpack(packed), q)
    assert np.array_equal(oracle.np_unpack(packed), q)
    assert np.array_equal(oracle.np_pack(q), packed)
    # nibble order: logical element j sits in nibble (0,4,1,5,2,6,3,7)[j]
    one = np.zeros((1, 8), dtype=np.uint8)
    one[0, 1] = 0xF
    assert (int(oracle.pack(one)[0, 0]) & 0xFFFFFFFF) == 0xF << 16


@pytest.mark.parametrize("K,N,G", [(128, 64, 128), (256, 128, 32), (256, 192, 64), (512, 256, 128),
                                   (512, 128, 256)])
def test_dequant_c_numpy_lop3_agree(oracle, K, N, G):
    _, qw, s, qz = oracle.make_inputs(1, K, N, G, seed=K + N + G, signed_scales=True)
    a = oracle.dequantize(qw, s, qz)
    b = oracle.np_dequantize(qw, s, qz)
    c = oracle.dequantize(qw, s, qz, lop3=True)
    assert np.array_equal(a.view(np.uint16), b.view(np.uint16))
    assert np.array_equal(a.view(np.uint16), c.view(np.uint16))


def test_dequant_zero_times_negative_scale_is_positive_zero(oracle):
    # fma(q - z, s, +0) yields +0 when q == z even for a negative scale (gemm_kernels.cu:153-160)
    q = np.full((128, 64), 5, dtype=np.uint8)
    qw, qz = oracle.pack(q), oracle.pack(np.full((1, 64), 5, dtype=np.uint8))
    s = np.full((1, 64), -0.01, dtype=np.float16)
    W = oracle.dequantize(qw, s, qz)
    assert np.all(W.view(np.uint16) == 0)


def test_dequant_extremes(oracle):
    # q - z spans [-15, 15]; the largest fp16 scale that keeps 15 * s finite
    q = np.zeros((128, 64), dtype=np.uint8)
    q[:, ::2] = 15
    z = np.zeros((1, 64), dtype=np.uint8)
    z[:, 1::2] = 15
    s = np.full((1, 64), 4096.0, dtype=np.float16)
    W = oracle.dequantize(oracle.pack(q), s, oracle.pack(z))
    assert np.all(W[:, ::2] == 15 * 4096.0) and np.all(W[:, 1::2] == -15 * 4096.0)


@pytest.mark.parametrize("K,N", [(128, 128), (256, 64), (384, 192), (128, 512)])
def test_b200_repack_roundtrip_and_numpy_agree(oracle, K, N):
    rng = np.random.default_rng(K * N)
    qw = rng.integers(-2**31, 2**31, size=(K, N // 8), dtype=np.int64).astype(np.int32)
    r = oracle.b200_repack(qw)
    assert r.size == oracle.b200_words(K, N) == ((N + 127) // 128) * (K // 128) * 2048
    assert np.array_equal(r, oracle.np_b200_repack(qw))
    assert np.array_equal(oracle.b200_unrepack(r, K, N), qw)


def test_b200_layout_definition(oracle):
    # word j*512 + c*4 + w of tile (nt, kb) holds channel nt*128+c, rows kb*128+(j*4+w)*8+e
    K, N = 256, 256
    q = np.zeros((K, N), dtype=np.uint8)
    k, n = 128 + 77, 128 + 5
    q[k, n] = 9
    r = oracle.b200_repack(oracle.pack(q)).view(np.uint32)
    nt, kb, c, o, e = n // 128, k // 128, n % 128, (k % 128) // 8, k % 8
    idx = (nt * (K // 128) + kb) * 2048 + (o // 4) * 512 + c * 4 + (o % 4)
    assert r[idx] == 9 << (4 * oracle.AWQ_ORDER[e])
    assert np.count_nonzero(r) == 1


def test_gemm_oracle_matches_numpy(oracle):
    x, qw, s, qz = oracle.make_inputs(5, 256, 128, 128, seed=3)
    W = oracle.dequantize(qw, s, qz)
    Yf, Yh = oracle.gemm_f32(x, W)
    ref = x.astype(np.float64) @ W.astype(np.float64)
    assert np.allclose(Yf, ref, rtol=1e-5, atol=1e-4)
    assert np.array_equal(Yh, Yf.astype(np.float16))
    assert np.allclose(oracle.awq_gemm(x, qw, s, qz), Yh)


def test_reference_splitk_model_is_close_to_fp32_oracle(oracle):
    x, qw, s, qz = oracle.make_inputs(4, 1024, 64, 128, seed=4)
    W = oracle.dequantize(qw, s, qz)
    Yf, _ = oracle.gemm_f32(x, W)
    Ys = oracle.gemm_ref_splitk(x, W, 8).astype(np.float32)
    assert np.max(np.abs(Ys - Yf)) < 2e-2


def test_empty_and_ragged(oracle):
    x, qw, s, qz = oracle.make_inputs(0, 128, 64, 128, seed=1)
    assert oracle.awq_gemm(x, qw, s, qz).shape == (0, 64)
    x, qw, s, qz = oracle.make_inputs(3, 128, 64, 32, seed=2)
    assert oracle.awq_gemm(x, qw, s, qz).shape == (3, 64)
