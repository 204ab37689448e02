"""CPU tests of the RMSNorm oracle (oracle/oracle.py): the restatement of the
reference's CUDA kernels (csrc/layernorm_kernels.cu:12-76) against the restatement of the
reference's own PyTorch statement ``RMSNorm._forward`` (layers/layernorm.py:26-44), with the
reference's tolerance and input recipe (tests/kernels/test_layernorm.py:7-59), plus the host
logic of the RMSNorm mirror."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import oracle  # noqa: E402


def _inputs(tokens, hidden, seed, scale=None):
    rng = np.random.default_rng(seed)
    scale = 1.0 / (2 * hidden) if scale is None else scale      # test_layernorm.py:36-38
    x = (rng.standard_normal((tokens, hidden)) * scale).astype(np.float16)
    r = (rng.standard_normal((tokens, hidden)) * scale).astype(np.float16)
    w = rng.normal(1.0, 0.1, hidden).astype(np.float16)         # test_layernorm.py:35
    return x, r, w


@pytest.mark.parametrize("tokens", [7, 83])
@pytest.mark.parametrize("hidden", [768, 5120, 8192])
@pytest.mark.parametrize("add_residual", [False, True])
def test_kernel_restatement_matches_forward_restatement(tokens, hidden, add_residual):
    x, r, w = _inputs(tokens, hidden, tokens + hidden)
    res = r if add_residual else None
    a = oracle.np_rms_norm_kernel(x, w, 1e-6, res)
    b = oracle.np_rms_norm_forward(x, w, 1e-6, res)
    if add_residual:
        assert np.array_equal(a[1].view(np.uint16), b[1].view(np.uint16))   # residual: exact
        a, b = a[0], b[0]
    assert np.allclose(a.astype(np.float32), b.astype(np.float32), atol=1e-2, rtol=1e-2)
    # the two differ only by where x is rounded: a few fp16 ulps
    ulp = np.spacing(np.abs(b).astype(np.float16)).astype(np.float32)
    assert np.all(np.abs(a.astype(np.float32) - b.astype(np.float32)) <= 4 * ulp)


def test_forward_restatement_matches_torch_formula():
    """Independent check of the NumPy restatement against the formula written with torch ops
    exactly as layernorm.py:31-41 writes it."""
    x, r, w = _inputs(33, 4096, 5, scale=1.0)
    tx, tr, tw = (torch.from_numpy(a.copy()) for a in (x, r, w))
    xf = tx.to(torch.float32) + tr.to(torch.float32)
    want_res = xf.to(torch.float16)
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    want = (xf * torch.rsqrt(var + 1e-5)).to(torch.float16) * tw
    got, got_res = oracle.np_rms_norm_forward(x, w, 1e-5, r)
    assert np.array_equal(got_res.view(np.uint16), want_res.numpy().view(np.uint16))
    d = np.abs(got.astype(np.float32) - want.float().numpy())
    assert np.all(d <= 2 * np.spacing(np.abs(want.numpy())).astype(np.float32))


def test_known_answers():
    # a constant row: mean(x^2) = c^2, out = sign * weight (eps = 0), residual doubles it
    w = np.arange(1, 9, dtype=np.float16)
    x = np.full((2, 8), 3.0, dtype=np.float16)
    assert np.array_equal(oracle.np_rms_norm_kernel(x, w, 0.0), np.tile(w, (2, 1)))
    out, res = oracle.np_rms_norm_kernel(x, w, 0.0, -2 * x)
    assert np.array_equal(res, -x) and np.array_equal(out, -np.tile(w, (2, 1)))
    # zero row with eps: rsqrt(eps) * 0 = 0
    z = np.zeros((1, 8), dtype=np.float16)
    assert not oracle.np_rms_norm_kernel(z, w, 1e-6).any()
    # statistics come from the UNROUNDED sum: 2048 + 1 is not an fp16 number
    x = np.full((1, 8), 2048.0, dtype=np.float16)
    r = np.ones((1, 8), dtype=np.float16)
    out, res = oracle.np_rms_norm_kernel(x, np.ones(8, np.float16), 0.0, r)
    assert np.all(res == np.float16(2048.0))
    assert np.all(out == np.float16(np.float32(2048.0) * (np.float32(1) / np.sqrt(np.float32(2049.0 ** 2)))))


def test_allreduce_restatement_rank_order():
    parts = [np.full(8, v, dtype=np.float16) for v in (2048.0, 1.0, 1.0)]
    # fp32 accumulate, one rounding: 2050 is representable; pairwise fp16 sums would give 2048
    assert np.all(oracle.np_allreduce_f16(parts) == np.float16(2050.0))


def test_rmsnorm_layer_host_logic():
    pkg = importlib.import_module("vllm-quick_b200")
    layer = pkg.RMSNorm(64, eps=1e-5)
    assert tuple(layer.weight.shape) == (64,) and layer.variance_epsilon == 1e-5
    assert torch.equal(layer.weight.data, torch.ones(64))
    with pytest.raises(RuntimeError):                     # CPU tensors: no fallback path
        layer(torch.zeros(2, 64, dtype=torch.float16))
    with pytest.raises(RuntimeError):
        pkg.ops.fused_add_rms_norm(torch.zeros(2, 64, dtype=torch.float16),
                                   torch.zeros(2, 64, dtype=torch.float16),
                                   torch.ones(64, dtype=torch.float16), 1e-6)


# ---- SiLU-and-multiply oracle ---------------------------------------------------

@pytest.mark.parametrize("tokens,d", [(7, 512), (83, 5120)])
def test_silu_restatement_matches_torch_forward(tokens, d):
    """oracle.np_silu_and_mul (the kernel's arithmetic) against SiluAndMul._forward written
    with torch ops (activation.py:25-28), at the reference's tolerance for fp16."""
    rng = np.random.default_rng(d)
    x = rng.standard_normal((tokens, 2 * d)).astype(np.float16)
    tx = torch.from_numpy(x.copy())
    ref = torch.nn.functional.silu(tx[..., :d].float()).half() * tx[..., d:]
    got = oracle.np_silu_and_mul(x)
    assert np.allclose(got.astype(np.float32), ref.float().numpy(), atol=1e-3, rtol=1e-3)


def test_silu_known_answers_and_pairing():
    x = np.array([[0.0, 65504.0, -65504.0, 1.0, 2.0, 3.0, 4.0, -5.0]], dtype=np.float16)
    out = oracle.np_silu_and_mul(x)
    assert out[0, 0] == 0 and np.isposinf(out[0, 1]) and out[0, 2] == 0
    assert np.isclose(float(out[0, 3]), -5.0 * 1.0 / (1.0 + np.exp(-1.0)), rtol=2e-3)
    # pairing is a permutation that alternates gate / up blocks and keeps each half's order
    a = np.arange(32, dtype=np.int32).reshape(1, 32)
    p = oracle.np_pair_gate_up(a, 8)
    assert p[0].tolist() == list(range(0, 8)) + list(range(16, 24)) + list(range(8, 16)) + list(range(24, 32))
    w = oracle.np_pair_gate_up(np.arange(6, dtype=np.int32).reshape(1, 6), 1)
    assert w[0].tolist() == [0, 3, 1, 4, 2, 5]
    # dequantize commutes with the pairing: channel-paired packed tensors are the paired matrix
    _, qw, s, qz = oracle.make_inputs(1, 256, 128, 128, seed=3)
    W = oracle.np_dequantize(qw, s, qz)
    Wp = oracle.np_dequantize(oracle.np_pair_gate_up(qw, 1), oracle.np_pair_gate_up(s, 8),
                              oracle.np_pair_gate_up(qz, 1))
    assert np.array_equal(Wp.view(np.uint16), oracle.np_pair_gate_up(W, 8).view(np.uint16))
