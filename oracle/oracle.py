"""CPU oracle for the W4A16 (AWQ / QUICK) hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product package
(``vllm-quick_b200/``) never does and fails loudly without its CUDA library.

Two independent restatements of the reference arithmetic live here:

* ``liboracle.so`` (``awq_oracle.c``) -- plain C, including a bit-level emulation of the
  reference's PTX sequence (``csrc/quantization/awq/dequantize.cuh:17-84``);
* the NumPy functions below (``np_*``) -- shift/mask unpack and
  ``fp16((q - z) * s)`` (``csrc/quantization/awq/gemm_kernels.cu:153-160``).

Parity pinning: the reference has no golden vectors for this path.
``tests/golden/`` holds outputs of the reference's own kernels compiled unmodified for
sm_100a (``oracle/build_ref.sh``) once a B200 run has produced them; see
``tests/golden/README.md`` for the current state.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

# Logical element j of a packed word sits in nibble AWQ_ORDER[j]
# (dequantize.cuh:42-56: lop3 pairs (0,4) (1,5) (2,6) (3,7)).
AWQ_ORDER = (0, 4, 1, 5, 2, 6, 3, 7)


def build(force: bool = False) -> str:
    """Compile ``liboracle.so`` with the committed Makefile (gcc only)."""
    if force or not os.path.exists(_LIB_PATH) or (
        os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "awq_oracle.c"))
    ):
        subprocess.run(["make", "-C", _HERE, "-s"] + (["-B"] if force else []), check=True)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        vp, i32 = ctypes.c_void_p, ctypes.c_int
        L.oracle_awq_dequantize.argtypes = [vp, vp, vp, vp, i32, i32, i32]
        L.oracle_awq_dequantize_lop3.argtypes = [vp, vp, vp, vp, i32, i32, i32]
        L.oracle_gemm_f32.argtypes = [vp, vp, vp, vp, i32, i32, i32]
        L.oracle_gemm_ref_splitk.argtypes = [vp, vp, vp, i32, i32, i32, i32]
        L.oracle_awq_pack.argtypes = [vp, vp, i32, i32]
        L.oracle_awq_unpack.argtypes = [vp, vp, i32, i32]
        L.oracle_b200_repack.argtypes = [vp, vp, i32, i32]
        L.oracle_b200_unrepack.argtypes = [vp, vp, i32, i32]
        L.oracle_b200_words.argtypes = [i32, i32]
        L.oracle_b200_words.restype = ctypes.c_size_t
        L.oracle_dequant_word_lop3.argtypes = [ctypes.c_uint32, vp]
        for f in ("oracle_awq_dequantize", "oracle_awq_dequantize_lop3", "oracle_gemm_f32",
                  "oracle_gemm_ref_splitk", "oracle_awq_pack", "oracle_awq_unpack",
                  "oracle_b200_repack", "oracle_b200_unrepack", "oracle_dequant_word_lop3"):
            getattr(L, f).restype = None
        _lib = L
    return _lib


def _p(a: np.ndarray) -> ctypes.c_void_p:
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


# ----------------------------------------------------------------------------- C oracle

def dequantize(qweight: np.ndarray, scales: np.ndarray, qzeros: np.ndarray,
               lop3: bool = False) -> np.ndarray:
    """``W[K,N]`` fp16 from AWQ tensors (gemm_kernels.cu:274-327 semantics)."""
    K, NW = qweight.shape
    N = NW * 8
    G = K // scales.shape[0]
    qweight = np.ascontiguousarray(qweight, dtype=np.int32)
    qzeros = np.ascontiguousarray(qzeros, dtype=np.int32)
    s16 = np.ascontiguousarray(scales, dtype=np.float16).view(np.uint16)
    W = np.empty((K, N), dtype=np.uint16)
    fn = lib().oracle_awq_dequantize_lop3 if lop3 else lib().oracle_awq_dequantize
    fn(_p(qweight), _p(s16), _p(qzeros), _p(W), K, N, G)
    return W.view(np.float16)


def gemm_f32(x: np.ndarray, W: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """fp32 dequant-then-matmul oracle: returns (Y fp32, Y rounded to fp16)."""
    M, K = x.shape
    N = W.shape[1]
    x16 = np.ascontiguousarray(x, dtype=np.float16).view(np.uint16)
    w16 = np.ascontiguousarray(W, dtype=np.float16).view(np.uint16)
    Yf = np.empty((M, N), dtype=np.float32)
    Yh = np.empty((M, N), dtype=np.uint16)
    lib().oracle_gemm_f32(_p(x16), _p(w16), _p(Yf), _p(Yh), M, K, N)
    return Yf, Yh.view(np.float16)


def gemm_ref_splitk(x: np.ndarray, W: np.ndarray, split_k: int = 8) -> np.ndarray:
    """Model of the reference's fp16-partial split-K numerics (gemm_kernels.cu:112-115,262-270,445)."""
    M, K = x.shape
    N = W.shape[1]
    x16 = np.ascontiguousarray(x, dtype=np.float16).view(np.uint16)
    w16 = np.ascontiguousarray(W, dtype=np.float16).view(np.uint16)
    Yh = np.empty((M, N), dtype=np.uint16)
    lib().oracle_gemm_ref_splitk(_p(x16), _p(w16), _p(Yh), M, K, N, split_k)
    return Yh.view(np.float16)


def awq_gemm(x, qweight, scales, qzeros) -> np.ndarray:
    """Oracle for ``ops.awq_gemm`` / ``ops.quick_gemm``: fp16 of the fp32 matmul."""
    return gemm_f32(x, dequantize(qweight, scales, qzeros))[1]


def pack(q: np.ndarray) -> np.ndarray:
    """``q[K,N]`` uint8 nibble values -> AWQ ``qweight[K,N/8]`` int32."""
    K, N = q.shape
    q = np.ascontiguousarray(q, dtype=np.uint8)
    out = np.empty((K, N // 8), dtype=np.int32)
    lib().oracle_awq_pack(_p(q), _p(out), K, N)
    return out


def unpack(qweight: np.ndarray) -> np.ndarray:
    K, NW = qweight.shape
    qweight = np.ascontiguousarray(qweight, dtype=np.int32)
    out = np.empty((K, NW * 8), dtype=np.uint8)
    lib().oracle_awq_unpack(_p(qweight), _p(out), K, NW * 8)
    return out


def b200_words(K: int, N: int) -> int:
    return int(lib().oracle_b200_words(K, N))


def b200_repack(qweight: np.ndarray) -> np.ndarray:
    """AWQ ``qweight[K,N/8]`` -> tile-major K-packed B200 layout (flat int32)."""
    K, NW = qweight.shape
    qweight = np.ascontiguousarray(qweight, dtype=np.int32)
    out = np.empty(b200_words(K, NW * 8), dtype=np.int32)
    lib().oracle_b200_repack(_p(qweight), _p(out), K, NW * 8)
    return out


def b200_unrepack(packed: np.ndarray, K: int, N: int) -> np.ndarray:
    packed = np.ascontiguousarray(packed, dtype=np.int32)
    out = np.empty((K, N // 8), dtype=np.int32)
    lib().oracle_b200_unrepack(_p(packed), _p(out), K, N)
    return out


def dequant_word_lop3(w: int) -> np.ndarray:
    out = np.empty(8, dtype=np.uint16)
    lib().oracle_dequant_word_lop3(ctypes.c_uint32(w & 0xFFFFFFFF), _p(out))
    return out.view(np.float16)


# ----------------------------------------------------------------------------- NumPy restatement

def np_unpack(qweight: np.ndarray) -> np.ndarray:
    """Shift/mask unpack in logical order (dequantize.cuh:17-84 restated)."""
    w = qweight.astype(np.int64) & 0xFFFFFFFF
    shifts = np.array([4 * o for o in AWQ_ORDER], dtype=np.int64)
    nib = (w[..., None] >> shifts) & 0xF
    return nib.reshape(*qweight.shape[:-1], qweight.shape[-1] * 8).astype(np.uint8)


def np_pack(q: np.ndarray) -> np.ndarray:
    q = q.astype(np.int64).reshape(*q.shape[:-1], q.shape[-1] // 8, 8)
    shifts = np.array([4 * o for o in AWQ_ORDER], dtype=np.int64)
    w = ((q & 0xF) << shifts).sum(axis=-1) & 0xFFFFFFFF
    return w.astype(np.uint32).view(np.int32)


def np_dequantize(qweight: np.ndarray, scales: np.ndarray, qzeros: np.ndarray) -> np.ndarray:
    """``fp16((q - z) * s + 0)`` with the reference's pairing (gemm_kernels.cu:128-160)."""
    K = qweight.shape[0]
    G = K // scales.shape[0]
    q = np_unpack(qweight).astype(np.float32)
    z = np.repeat(np_unpack(qzeros).astype(np.float32), G, axis=0)
    s = np.repeat(scales.astype(np.float32), G, axis=0)
    return ((q - z) * s + np.float32(0.0)).astype(np.float16)


def np_gemm(x: np.ndarray, W: np.ndarray) -> np.ndarray:
    return x.astype(np.float32) @ W.astype(np.float32)


def np_b200_repack(qweight: np.ndarray) -> np.ndarray:
    """Vectorised restatement of the B200 layout, N % 128 == 0 or 64."""
    K, NW = qweight.shape
    N = NW * 8
    NT, KB = (N + 127) // 128, K // 128
    q = np.zeros((K, NT * 128), dtype=np.int64)
    q[:, :N] = np_unpack(qweight)
    # [kb, j, w, e, nt, c] -> [nt, kb, j, c, w, e]
    q = q.reshape(KB, 4, 4, 8, NT, 128).transpose(4, 0, 1, 5, 2, 3)
    shifts = np.array([4 * o for o in AWQ_ORDER], dtype=np.int64)
    words = ((q & 0xF) << shifts).sum(axis=-1) & 0xFFFFFFFF
    return words.astype(np.uint32).view(np.int32).reshape(-1)


# ----------------------------------------------------------------------------- synthetic inputs

def make_inputs(M: int, K: int, N: int, G: int = 128, seed: int = 0, signed_scales: bool = False):
    """Seeded synthetic AWQ tensors: x ~ N(0,1) fp16, uniform nibbles,
    scales ~ U(1e-3, 2e-2) fp16.  NumPy generator so CPU and GPU tests share inputs."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((M, K), dtype=np.float32).astype(np.float16)
    qweight = rng.integers(-2**31, 2**31, size=(K, N // 8), dtype=np.int64).astype(np.int32)
    qzeros = rng.integers(-2**31, 2**31, size=(K // G, N // 8), dtype=np.int64).astype(np.int32)
    scales = rng.uniform(1e-3, 2e-2, size=(K // G, N)).astype(np.float16)
    if signed_scales:
        scales = (scales * rng.choice([-1.0, 1.0], size=scales.shape)).astype(np.float16)
    return x, qweight, scales, qzeros


# ---- residual add + RMSNorm ---------------------------------------------------
# The reference has its own CPU statement of this op -- RMSNorm._forward,
# vllm/model_executor/layers/layernorm.py:26-44 -- and tests its kernel against it with
# atol = rtol = 1e-2 (tests/kernels/test_layernorm.py:30-59).  `np_rms_norm_forward` restates
# `_forward`; `np_rms_norm_kernel` restates the CUDA kernels' arithmetic
# (csrc/layernorm_kernels.cu:12-76): statistics from the UNROUNDED fp32 sum, the scaling pass
# from the ROUNDED residual, half x half product rounded once.  The sum of squares is taken in
# float64 here (the kernels' shuffle-tree order is not part of the contract): tolerance parity.
# Parity pin: the reference's kernel cannot run in this container and its test holds no
# vectors; the two restatements are checked against each other and against the reference's
# tolerance ("parity unpinned" beyond that).

def make_norm_inputs(tokens: int, hidden: int, seed: int, scale=None):
    """Seeded (x, residual, weight) for the RMSNorm parity cases; recipe of the reference's own
    test (tests/kernels/test_layernorm.py:35-38: weight ~ N(1, 0.1), x scaled by 1/(2 hidden))."""
    rng = np.random.default_rng(seed)
    scale = 1.0 / (2 * hidden) if scale is None else scale
    x = (rng.standard_normal((tokens, hidden)) * scale).astype(np.float16)
    r = (rng.standard_normal((tokens, hidden)) * scale).astype(np.float16)
    w = rng.normal(1.0, 0.1, hidden).astype(np.float16)
    return x, r, w


def make_act_inputs(tokens: int, d: int, seed: int, scale: float = 1.0) -> np.ndarray:
    """Seeded [tokens, 2 d] input of SiLU-and-multiply (tests/kernels/test_activation.py:33)."""
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((tokens, 2 * d)) * scale).astype(np.float16)


def np_rms_norm_forward(x: np.ndarray, weight: np.ndarray, eps: float, residual=None):
    """RMSNorm._forward (layernorm.py:26-44) for fp16 arrays."""
    x32 = x.astype(np.float32)
    res_out = None
    if residual is not None:
        x32 = x32 + residual.astype(np.float32)
        res_out = x32.astype(np.float16)
    var = np.mean(x32 * x32, axis=-1, keepdims=True, dtype=np.float32)
    y = x32 * (np.float32(1.0) / np.sqrt(var + np.float32(eps)))
    out = (y.astype(np.float16).astype(np.float32) * weight.astype(np.float32)).astype(np.float16)
    return out if residual is None else (out, res_out)


def np_rms_norm_kernel(x: np.ndarray, weight: np.ndarray, eps: float, residual=None):
    """rms_norm_kernel / fused_add_rms_norm_kernel (csrc/layernorm_kernels.cu:12-76)."""
    x32 = x.astype(np.float32)
    if residual is not None:
        x32 = x32 + residual.astype(np.float32)            # :49-50
    res_out = x32.astype(np.float16)                        # :52 (== x without a residual)
    hidden = x.shape[-1]
    var = (x32.astype(np.float64) ** 2).sum(axis=-1, keepdims=True)          # :51,54
    inv = (np.float32(1.0) / np.sqrt((var / hidden).astype(np.float32) + np.float32(eps))).astype(np.float32)  # :56
    y = (res_out.astype(np.float32) * inv).astype(np.float16)               # :61-62
    out = (y.astype(np.float32) * weight.astype(np.float32)).astype(np.float16)
    return out if residual is None else (out, res_out)


def np_allreduce_f16(parts) -> np.ndarray:
    """fp16 all-reduce as the peer-memory kernel defines it: fp32 sum in rank order, one
    rounding (custom_all_reduce.cuh:204-232 accumulates 16-bit types in fp32 likewise)."""
    acc = np.zeros(parts[0].shape, dtype=np.float32)
    for p in parts:
        acc = acc + p.astype(np.float32)
    return acc.astype(np.float16)


# ---- SiLU-and-multiply --------------------------------------------------------
# np_silu_and_mul restates silu_and_mul_kernel (csrc/activation_kernels.cu:10-28):
# half(float(x) / (1 + expf(-float(x)))) * y as a half multiply; the reference's own CPU
# statement is SiluAndMul._forward = F.silu(x[..., :d]) * x[..., d:] (activation.py:25-28) and
# its test accepts atol = rtol = 1e-3 for fp16 (tests/kernels/allclose_default.py).  expf is a
# library function (device expf vs NumPy exp differ by an fp32 ulp at most): tolerance parity,
# one fp16 ulp.  No vectors in the reference: "parity unpinned" beyond its tolerance.

def np_silu_and_mul(x: np.ndarray) -> np.ndarray:
    d = x.shape[-1] // 2
    g = x[..., :d].astype(np.float32)
    s = (g / (np.float32(1.0) + np.exp(-g, dtype=np.float32))).astype(np.float16)     # :10-14
    return (s.astype(np.float32) * x[..., d:].astype(np.float32)).astype(np.float16)  # :26


def np_pair_gate_up(a: np.ndarray, block: int) -> np.ndarray:
    """Column permutation of a merged [gate | up] tensor into the paired order: blocks of
    `block` columns (1 for packed int32 words, 8 for per-channel tensors) alternate gate / up."""
    n = a.shape[-1] // block
    half = np.arange(n // 2)
    order = np.stack([half, half + n // 2], axis=1).reshape(-1)
    return np.ascontiguousarray(a.reshape(a.shape[:-1] + (n, block))[..., order, :].reshape(a.shape))
