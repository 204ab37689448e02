"""-m gpu parity tests of SiLU-and-multiply: the standalone kernel and the
gate_up GEMM with the activation in its epilogue (csrc/w4a16_norm.cu, w4a16_decode.cu).

Tolerances: the activation is floating point with a library exponential -- two fp16 ulps
against the oracle (oracle.np_silu_and_mul) and the reference's own bar for this kernel,
atol = rtol = 1e-3 against F.silu(x[:d]) * x[d:] (tests/kernels/test_activation.py:22-40,
allclose_default.py).  The fused GEMM must equal GEMM-then-activation on the device bit for
bit, and the fp32 oracle within the GEMM bar (1e-2 of max(1, |ref|))."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tokens", [1, 7, 83, 2048])
@pytest.mark.parametrize("d", [512, 4096, 5120, 13824])
def test_silu_and_mul_matches_oracle(pkg, oracle, tokens, d):
    from gpu_util import to_dev
    rng = np.random.default_rng(tokens + d)
    x = rng.standard_normal((tokens, 2 * d)).astype(np.float16)
    x[0, :8] = [0.0, -0.0, 65504.0, -65504.0, 1e-7, -20.0, 20.0, 11.0]      # edge values
    out = pkg.SiluAndMul()(to_dev(x))
    assert out.shape == (tokens, d)
    want = oracle.np_silu_and_mul(x)
    got = out.float().cpu().numpy()
    fin = np.isfinite(want)
    assert np.array_equal(got[~fin], want[~fin].astype(np.float32))          # overflow -> the same inf
    ulp = np.spacing(np.maximum(np.abs(want[fin]), np.float16(6.1e-5))).astype(np.float32)
    # expf is a library function: an fp32 ulp there can move silu(x) by one fp16 ulp, which
    # the product's rounding can turn into two
    assert np.all(np.abs(got[fin] - want[fin].astype(np.float32)) <= 2 * ulp)
    tx = torch.from_numpy(x.copy()).cuda()
    ref = torch.nn.functional.silu(tx[..., :d]) * tx[..., d:]                # SiluAndMul._forward
    assert torch.allclose(out, ref, atol=1e-3, rtol=1e-3, equal_nan=True)
    # paired order: the same numbers after the column permutation
    xp = oracle.np_pair_gate_up(x, 8)
    outp = torch.empty_like(out)
    pkg.ops.silu_and_mul_paired(outp, to_dev(xp))
    assert torch.equal(outp, out)


def test_silu_and_mul_rejects_bad_arguments(pkg, cuda_lib):
    s = torch.cuda.current_stream().cuda_stream
    x = torch.zeros(4, 64, dtype=torch.float16, device="cuda")
    o = torch.zeros(4, 32, dtype=torch.float16, device="cuda")
    L, e = cuda_lib, pkg._lib
    assert L.b200_w4a16_silu_and_mul(x.data_ptr(), o.data_ptr(), 4, 12, 0, s) == e.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_silu_and_mul(x.data_ptr(), o.data_ptr(), 4, 32, 2, s) == e.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_silu_and_mul(None, o.data_ptr(), 4, 32, 0, s) == e.B200_ERR_BAD_ARG
    assert L.b200_w4a16_silu_and_mul(x.data_ptr(), o.data_ptr(), 0, 32, 0, s) == 0
    with pytest.raises(ValueError):
        pkg.ops.silu_and_mul(torch.zeros(4, 16, dtype=torch.float16, device="cuda"), x)
    with pytest.raises(ValueError):
        pkg.ops.silu_and_mul(o.float(), x.float())


def _paired(pkg, oracle, qw, s, qz):
    from gpu_util import to_dev
    return (pkg.ops.quick_repack(to_dev(oracle.np_pair_gate_up(qw, 1))),
            to_dev(oracle.np_pair_gate_up(s, 8)), to_dev(oracle.np_pair_gate_up(qz, 1)))


@pytest.mark.parametrize("M", [1, 5, 8, 16])
@pytest.mark.parametrize("K,N", [(512, 256), (4096, 2 * 11008), (1024, 2 * 1376), (11008, 1024)])
def test_gemm_silu_mul_fused_equals_two_step_and_oracle(pkg, oracle, cuda_lib, M, K, N):
    """N = 2 * intermediate.  (1024, 2752): intermediate 1376 = 11008 / 8, N % 128 = 64 --
    the padded last tile column; (4096, 22016): Llama-2-7B gate_up at full size."""
    from gpu_util import to_dev
    x, qw, s, qz = oracle.make_inputs(M, K, N, 128, seed=M + N)
    dx = to_dev(x)
    r, ds, dqz = _paired(pkg, oracle, qw, s, qz)
    assert cuda_lib.b200_w4a16_gemm_variant(M, K, N, 128) == 1
    fused = pkg.ops.quick_gemm_silu_mul(dx, r, ds, dqz)
    assert fused.shape == (M, N // 2)
    full = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)                 # paired channel order
    two = torch.empty_like(fused)
    pkg.ops.silu_and_mul_paired(two, full)
    assert torch.equal(fused, two)
    assert torch.equal(fused, pkg.ops.quick_gemm_silu_mul(dx, r, ds, dqz))      # deterministic
    # against the oracle: fp32 GEMM of the ORIGINAL weights, rounded like the reference's
    # GEMM output, then the activation restatement
    W = oracle.np_dequantize(qw, s, qz)
    Y = (x.astype(np.float32) @ W.astype(np.float32))
    want = oracle.np_silu_and_mul(Y.astype(np.float16)).astype(np.float32)
    d = np.abs(fused.float().cpu().numpy() - want)
    # the GEMM bar on both factors: |d(silu(g) u)| <= 1e-2 max(1,|g|) |u| + 1e-2 max(1,|u|) |silu(g)|
    g, u = np.abs(Y[:, : N // 2]), np.abs(Y[:, N // 2:])
    bound = 1.1e-2 * (np.maximum(1, g) * np.maximum(u, 1) + np.maximum(1, u) * np.maximum(g, 1))
    assert np.all(d <= bound)
    # and tight in the typical case
    assert np.median(d / np.maximum(np.abs(want), 1.0)) < 2e-3


@pytest.mark.parametrize("M", [17, 64, 300])
def test_gemm_silu_mul_above_decode_batches(pkg, oracle, M):
    from gpu_util import to_dev
    K, N = 1024, 2 * 1024
    x, qw, s, qz = oracle.make_inputs(M, K, N, 128, seed=M)
    r, ds, dqz = _paired(pkg, oracle, qw, s, qz)
    out = pkg.ops.quick_gemm_silu_mul(to_dev(x), r, ds, dqz)
    W = oracle.np_dequantize(qw, s, qz)
    Y = (x.astype(np.float32) @ W.astype(np.float32)).astype(np.float16)
    want = oracle.np_silu_and_mul(Y).astype(np.float32)
    d = np.abs(out.float().cpu().numpy() - want)
    assert np.median(d / np.maximum(np.abs(want), 1.0)) < 2e-3
    assert np.all(d <= 5e-2 * np.maximum(np.abs(want), 1.0) + 5e-2)


def test_llama_mlp_fused_matches_unfused(pkg, oracle):
    """LlamaMLP mirror (llama.py:48-80): weights loaded through the layers' weight_loader, the
    fused-activation block against the unfused block on the same checkpoint tensors."""
    from gpu_util import to_dev
    hidden, inter = 1024, 2816
    method = pkg.QuickConfig(4, 128, True).get_linear_method()
    _, gqw, gs, gqz = oracle.make_inputs(1, hidden, inter, 128, seed=1)
    _, uqw, us, uqz = oracle.make_inputs(1, hidden, inter, 128, seed=2)
    _, dqw, dsc, dqz = oracle.make_inputs(1, inter, hidden, 128, seed=3)
    blocks = []
    for fuse in (True, False):
        mlp = pkg.LlamaMLP(hidden, inter, "silu", method, torch.float16, fuse_activation=fuse).cuda()
        for name, (a, b) in {"qweight": (gqw, uqw), "scales": (gs, us), "qzeros": (gqz, uqz)}.items():
            prm = getattr(mlp.gate_up_proj, name)
            prm.weight_loader(prm, to_dev(a), 0)
            prm.weight_loader(prm, to_dev(b), 1)
        for name, a in {"qweight": dqw, "scales": dsc, "qzeros": dqz}.items():
            prm = getattr(mlp.down_proj, name)
            prm.weight_loader(prm, to_dev(a))
        blocks.append(mlp)
    assert blocks[0].fuse_activation and not blocks[1].fuse_activation
    for M in (1, 16, 40):
        x = (torch.randn(M, hidden, device="cuda") * 0.5).half()
        a, b = blocks[0](x), blocks[1](x)
        assert a.shape == (M, hidden)
        err = (a.float() - b.float()).abs()
        assert float((err / b.float().abs().clamp(min=1.0)).max()) <= 1e-2
    with pytest.raises(RuntimeError):                       # a paired layer serves the fused call only
        blocks[0].gate_up_proj(x)
    # CUDA-graph capture of the fused block
    x = (torch.randn(16, hidden, device="cuda") * 0.5).half()
    want = blocks[0](x)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            y = blocks[0](x)
        g.replay(); g.replay()
        st.synchronize()
    assert torch.equal(y, want)
