"""Parity against outputs of the REFERENCE'S OWN kernels (tests/golden/, see its README).

CPU (-m "not gpu"): the oracle restatements reproduce the reference's dequantize bit for bit
and its GEMM within the reference's own split-K fp16 rounding.
GPU (-m gpu): the CUDA path reproduces the reference's dequantize bit for bit (both layouts)
and its GEMM within the north-star bar (1e-2, absolute below |y| = 1, relative above)."""
import hashlib
import os

import numpy as np
import pytest
import torch

from gpu_util import to_dev

HERE = os.path.dirname(os.path.abspath(__file__))
SMALL = np.load(os.path.join(HERE, "golden", "ref_small.npz"))
FULL = np.load(os.path.join(HERE, "golden", "ref_full.npz"))
N_SMALL = sum(1 for k in SMALL.files if k.endswith("_shape"))
N_FULL = sum(1 for k in FULL.files if k.endswith("_shape"))
TOL = 1e-2


def _has_gemm_pin(i, N):
    # The reference's 64-column kernel variant (taken when N % 128 != 0,
    # csrc/quantization/awq/gemm_kernels.cu:424-444) disagrees with the reference's OWN
    # dequantize on the same inputs (cases 2 and 3: max error 5.6 / 2.8), while its 128-column
    # variant agrees to 3e-3.  The dequantize + matmul semantics -- also the reference's branch
    # for >= 256 tokens -- are the contract; GEMM outputs are pinned where N % 128 == 0.
    return f"case{i}_Y" in SMALL.files and N % 128 == 0


def _f16(bits):
    return bits.view(np.float16)


def _close(y, ref, tol):
    y, ref = y.astype(np.float32), ref.astype(np.float32)
    return np.max(np.abs(y - ref) / np.maximum(1.0, np.abs(ref))) <= tol


@pytest.mark.parametrize("i", range(N_SMALL))
def test_oracle_matches_reference_kernels_small(oracle, i):
    M, K, N, G, seed = (int(v) for v in SMALL[f"case{i}_shape"])
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=seed)
    W = oracle.dequantize(qw, s, qz)
    assert np.array_equal(W.view(np.uint16), SMALL[f"case{i}_W"])            # C restatement
    assert np.array_equal(oracle.np_dequantize(qw, s, qz).view(np.uint16), SMALL[f"case{i}_W"])  # NumPy
    if _has_gemm_pin(i, N):
        Yref = _f16(SMALL[f"case{i}_Y"])
        Yf, _ = oracle.gemm_f32(x, W)
        assert _close(Yf, Yref, TOL)
        # the model of the reference's own arithmetic (fp16 split-K partials, then sum) is closer
        Ys = oracle.gemm_ref_splitk(x, W, 8)
        assert _close(Ys, Yref, 4e-3)


@pytest.mark.parametrize("i", range(N_FULL))
def test_oracle_matches_reference_kernels_full_size(oracle, i):
    M, K, N, G, seed = (int(v) for v in FULL[f"case{i}_shape"])
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=seed)
    W = oracle.dequantize(qw, s, qz).view(np.uint16)
    assert hashlib.sha256(W.tobytes()).digest() == FULL[f"case{i}_W_sha256"].tobytes()
    assert np.array_equal(W[::97, ::89], FULL[f"case{i}_W_sample"])
    cols = np.arange(0, N, 61)
    Yf = x.astype(np.float32) @ W.view(np.float16)[:, cols].astype(np.float32)
    assert _close(Yf, _f16(FULL[f"case{i}_Y_sample"]), TOL)


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(N_SMALL))
def test_cuda_path_matches_reference_kernels_small(pkg, oracle, cuda_lib, i):
    M, K, N, G, seed = (int(v) for v in SMALL[f"case{i}_shape"])
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=seed)
    dqw, ds, dqz = to_dev(qw), to_dev(s), to_dev(qz)
    W = pkg.ops.awq_dequantize(dqw, ds, dqz, 0, 0, 0)
    assert np.array_equal(W.cpu().view(torch.int16).numpy().view(np.uint16), SMALL[f"case{i}_W"])
    if K % 128 == 0 and N % 64 == 0:
        Wq = pkg.ops.quick_dequantize(pkg.ops.quick_repack(dqw), ds, dqz)
        assert torch.equal(W, Wq)
    if _has_gemm_pin(i, N):
        y = pkg.ops.awq_gemm(to_dev(x), dqw, ds, dqz, 8)
        assert _close(y.cpu().numpy(), _f16(SMALL[f"case{i}_Y"]), TOL)


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(N_FULL))
def test_cuda_path_matches_reference_kernels_full_size(pkg, oracle, cuda_lib, i):
    M, K, N, G, seed = (int(v) for v in FULL[f"case{i}_shape"])
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=seed)
    dqw, ds, dqz = to_dev(qw), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(dqw)
    for W in (pkg.ops.awq_dequantize(dqw, ds, dqz, 0, 0, 0), pkg.ops.quick_dequantize(r, ds, dqz)):
        Wb = W.cpu().view(torch.int16).numpy().view(np.uint16)
        assert hashlib.sha256(Wb.tobytes()).digest() == FULL[f"case{i}_W_sha256"].tobytes()
    y = pkg.ops.quick_gemm(to_dev(x), r, ds, dqz, 8)
    assert _close(y.cpu().numpy()[:, ::61], _f16(FULL[f"case{i}_Y_sample"]), TOL)
