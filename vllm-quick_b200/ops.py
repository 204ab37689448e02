"""Operator surface of the QUICK / AWQ hot path, same names and argument order as the
reference's ``vllm._C.ops`` (csrc/pybind.cpp:51-55, csrc/ops.h:68-83):

* ``awq_gemm(in_feats, kernel, scaling_factors, zeros, split_k_iters)``
* ``awq_dequantize(kernel, scaling_factors, zeros, split_k_iters, thx, thy)``
* ``quick_gemm(...)`` -- same five arguments, ``kernel`` in the B200 interleave
* ``quick_repack`` / ``quick_unrepack`` / ``quick_dequantize`` -- layout helpers

PyTorch is used for device memory and streams only; all arithmetic is in
``libb200_w4a16.so``.  Nothing here falls back to torch math.
"""
from __future__ import annotations

import contextlib
import ctypes
import threading
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import LAYOUT_AWQ, LAYOUT_B200

PACK_FACTOR = 8
# fp32 partial tiles of one full wave (148 CTAs x 2 slots x 2 channel tiles x 128 x 128 x 4 B):
# covers every plan
_BASE_WORKSPACE_BYTES = 148 * 2 * 2 * 128 * 128 * 4

# Split-K scratch (include/b200_w4a16.h: b200_w4a16_workspace_init).  Eager launches use one
# workspace per (device, stream) -- GEMMs on different streams may overlap.  Launches recorded
# into a CUDA graph use ONE workspace per device that is created and initialised eagerly,
# before any capture: the reference's CUDAGraphRunner captures ~35 graphs on torch's own
# capture stream (worker/model_runner.py:781), and scratch first created inside capture #1
# would be initialised by a node of that graph only -- replaying any other graph first would
# read uninitialised partial slots.  Captured graphs of one device therefore share scratch and
# must not be replayed concurrently with each other (the reference replays one at a time).
_workspaces: Dict[Tuple[int, int], torch.Tensor] = {}
_capture_workspaces: Dict[int, torch.Tensor] = {}


def _stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _device_index(device: torch.device) -> int:
    return device.index if device.index is not None else torch.cuda.current_device()


def _new_workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    ws = torch.empty(max(nbytes, _BASE_WORKSPACE_BYTES), dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        _lib.check(_lib.lib().b200_w4a16_workspace_init(ws.data_ptr(), ws.numel(), _stream_ptr(device)),
                   "workspace init")
    return ws


def ensure_workspace(device: torch.device, nbytes: int = 0) -> None:
    """Create (eagerly, never inside a capture) the scratch that CUDA-graph captures on
    ``device`` will use.  Called by the linear methods when weights are created / loaded;
    call it yourself with a larger ``nbytes`` before capturing ``awq_gemm`` on checkpoint-layout
    weights (their on-the-fly interleave needs room for the whole matrix)."""
    device = torch.device(device)
    idx = _device_index(device)
    ws = _capture_workspaces.get(idx)
    if ws is not None and ws.numel() >= nbytes:
        return
    if torch.cuda.is_current_stream_capturing():
        raise RuntimeError("b200 w4a16: the split-K workspace must exist before CUDA-graph capture "
                           "starts -- call ops.ensure_workspace(device[, nbytes]) first")
    _capture_workspaces[idx] = _new_workspace(device, nbytes)
    torch.cuda.current_stream(device).synchronize()      # graphs may replay on any stream


def _workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    idx = _device_index(device)
    if torch.cuda.is_current_stream_capturing():
        ensure_workspace(device, nbytes)                  # raises if it would have to allocate
        return _capture_workspaces[idx]
    if idx not in _capture_workspaces:
        ensure_workspace(device)
    key = (idx, _stream_ptr(device))
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = _new_workspace(device, nbytes)
        _workspaces[key] = ws
    return ws


def _check_cuda(*tensors: torch.Tensor) -> torch.device:
    dev = tensors[0].device
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError("b200 w4a16 ops need CUDA tensors: there is no CPU path")
        if t.device != dev:
            raise RuntimeError("all tensors must be on the same device")
        if not t.is_contiguous():
            raise RuntimeError("tensors must be contiguous")
    return dev


# Routed-expert gate: inside ``with expert_gate(flag):`` every GEMM issued by this thread
# carries ``flag`` (int32 device scalar: tokens of the batch routed to the expert) down to
# b200_w4a16_gemm_gated.  A context instead of an argument because the plugin interface
# (LinearMethodBase.apply_weights(weights, x, bias), linear.py:30-37) has no room for one.
_gate = threading.local()


@contextlib.contextmanager
def expert_gate(flag: Optional[torch.Tensor]):
    if flag is not None and (flag.dtype != torch.int32 or not flag.is_cuda or flag.numel() != 1):
        raise ValueError("expert_gate needs an int32 CUDA scalar")
    prev = getattr(_gate, "flag", None)
    _gate.flag = flag
    try:
        yield
    finally:
        _gate.flag = prev


def _gemm(in_feats: torch.Tensor, kernel: torch.Tensor, scaling_factors: torch.Tensor,
          zeros: torch.Tensor, split_k_iters: int, layout: int) -> torch.Tensor:
    dev = _check_cuda(in_feats, kernel, scaling_factors, zeros)
    if in_feats.dtype != torch.float16 or scaling_factors.dtype != torch.float16:
        raise ValueError("in_feats and scaling_factors must be float16")
    if kernel.dtype != torch.int32 or zeros.dtype != torch.int32:
        raise ValueError("kernel and zeros must be int32")
    if in_feats.dim() != 2:
        raise ValueError("in_feats must be [num_tokens, in_features]")
    M, K = in_feats.shape
    N = scaling_factors.shape[1]
    if scaling_factors.shape[0] == 0 or K % scaling_factors.shape[0] != 0:
        raise ValueError("in_features must be a multiple of the number of groups")
    G = K // scaling_factors.shape[0]
    if zeros.shape != (K // G, N // PACK_FACTOR):
        raise ValueError("zeros must be [in_features / group_size, out_features / 8]")
    L = _lib.lib()
    if layout == LAYOUT_AWQ:
        if kernel.shape != (K, N // PACK_FACTOR):
            raise ValueError("kernel must be [in_features, out_features / 8]")
    elif kernel.numel() != L.b200_w4a16_repacked_words(K, N):
        raise ValueError("kernel does not have the B200 layout size for these shapes")
    out = torch.empty((M, N), dtype=torch.float16, device=dev)
    with torch.cuda.device(dev):
        need = L.b200_w4a16_workspace_bytes(M, K, N, layout)
        ws = _workspace(dev, need)
        flag = getattr(_gate, "flag", None)
        st = L.b200_w4a16_gemm_gated(in_feats.data_ptr(), kernel.data_ptr(),
                                     scaling_factors.data_ptr(), zeros.data_ptr(), out.data_ptr(),
                                     M, K, N, G, layout, int(split_k_iters),
                                     flag.data_ptr() if flag is not None else None,
                                     ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    _lib.check(st, "w4a16 gemm")
    return out


def awq_gemm(in_feats: torch.Tensor, kernel: torch.Tensor, scaling_factors: torch.Tensor,
             zeros: torch.Tensor, split_k_iters: int) -> torch.Tensor:
    """``in_feats[M,K] @ dequant(kernel[K,N/8], scaling_factors[K/G,N], zeros[K/G,N/8])``.

    Same contract as the reference op (csrc/quantization/awq/gemm_kernels.cu:389-446);
    ``kernel`` is in the AWQ checkpoint layout and is interleaved on the fly (use
    ``quick_gemm`` with a pre-interleaved weight on the hot path).
    """
    return _gemm(in_feats, kernel, scaling_factors, zeros, split_k_iters, LAYOUT_AWQ)


def quick_gemm(in_feats: torch.Tensor, kernel: torch.Tensor, scaling_factors: torch.Tensor,
               zeros: torch.Tensor, split_k_iters: int) -> torch.Tensor:
    """Same five arguments as ``awq_gemm`` with ``kernel`` produced by ``quick_repack``."""
    return _gemm(in_feats, kernel, scaling_factors, zeros, split_k_iters, LAYOUT_B200)


def quick_gemm_all_reduce(in_feats: torch.Tensor, kernel: torch.Tensor,
                          scaling_factors: torch.Tensor, zeros: torch.Tensor, handle) -> torch.Tensor:
    """Row-parallel ``quick_gemm`` fused with the tensor-parallel all-reduce (one kernel per
    rank, tiles pushed over NVLink as they finish): ``handle`` is the CustomAllreduce whose
    ``can_fuse(M, N, G)`` accepted this shape.  Collective over the TP group."""
    dev = _check_cuda(in_feats, kernel, scaling_factors, zeros)
    if in_feats.dtype != torch.float16 or scaling_factors.dtype != torch.float16:
        raise ValueError("in_feats and scaling_factors must be float16")
    M, K = in_feats.shape
    N = scaling_factors.shape[1]
    G = K // scaling_factors.shape[0]
    L = _lib.lib()
    if kernel.numel() != L.b200_w4a16_repacked_words(K, N):
        raise ValueError("kernel does not have the B200 layout size for these shapes")
    with torch.cuda.device(dev):
        ws = _workspace(dev, L.b200_w4a16_workspace_bytes(M, K, N, LAYOUT_B200))
        return handle.gemm_all_reduce(in_feats, kernel, scaling_factors, zeros, ws, G)


def _dequant(kernel, scaling_factors, zeros, layout) -> torch.Tensor:
    dev = _check_cuda(kernel, scaling_factors, zeros)
    if kernel.dtype != torch.int32 or zeros.dtype != torch.int32:
        raise ValueError("kernel and zeros must be int32")
    if scaling_factors.dtype != torch.float16:
        raise ValueError("scaling_factors must be float16")
    N = scaling_factors.shape[1]
    L = _lib.lib()
    if layout == LAYOUT_AWQ:
        K = kernel.shape[0]
    else:
        if N == 0 or (kernel.numel() * PACK_FACTOR) % (((N + 127) // 128) * 128) != 0:
            raise ValueError("kernel does not have the B200 layout size for these shapes")
        K = kernel.numel() * PACK_FACTOR // (((N + 127) // 128) * 128)
    if scaling_factors.shape[0] == 0 or K % scaling_factors.shape[0] != 0:
        raise ValueError("in_features must be a multiple of the number of groups")
    G = K // scaling_factors.shape[0]  # gemm_kernels.cu:343
    out = torch.empty((K, N), dtype=scaling_factors.dtype, device=dev)
    with torch.cuda.device(dev):
        st = L.b200_w4a16_dequant(kernel.data_ptr(), scaling_factors.data_ptr(), zeros.data_ptr(),
                                  out.data_ptr(), K, N, G, layout, _stream_ptr(dev))
    _lib.check(st, "w4a16 dequantize")
    return out


def awq_dequantize(kernel: torch.Tensor, scaling_factors: torch.Tensor, zeros: torch.Tensor,
                   split_k_iters: int = 0, thx: int = 0, thy: int = 0) -> torch.Tensor:
    """``W[K,N] = fp16((q - z) * s)``, bit-exact with the reference
    (csrc/quantization/awq/gemm_kernels.cu:274-381).  The three trailing launch-shape
    arguments of the reference are accepted and ignored."""
    return _dequant(kernel, scaling_factors, zeros, LAYOUT_AWQ)


def quick_dequantize(kernel: torch.Tensor, scaling_factors: torch.Tensor,
                     zeros: torch.Tensor) -> torch.Tensor:
    return _dequant(kernel, scaling_factors, zeros, LAYOUT_B200)


def _repack(kernel: torch.Tensor, K: int, N: int, direction: int) -> torch.Tensor:
    dev = _check_cuda(kernel)
    if kernel.dtype != torch.int32:
        raise ValueError("kernel must be int32")
    L = _lib.lib()
    words = L.b200_w4a16_repacked_words(K, N)
    if direction == _lib.REPACK_AWQ_TO_B200:
        out = torch.empty((K, N // PACK_FACTOR) if words == K * N // PACK_FACTOR else (words,),
                          dtype=torch.int32, device=dev)
    else:
        if kernel.numel() != words:
            raise ValueError("kernel does not have the B200 layout size for these shapes")
        out = torch.empty((K, N // PACK_FACTOR), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        st = L.b200_w4a16_repack(kernel.data_ptr(), out.data_ptr(), K, N, direction,
                                 _stream_ptr(dev))
    _lib.check(st, "w4a16 repack")
    return out


def quick_repack(kernel: torch.Tensor) -> torch.Tensor:
    """AWQ ``qweight[K, N/8]`` -> B200 interleave (same number of words when N % 128 == 0)."""
    if kernel.dim() != 2:
        raise ValueError("kernel must be [in_features, out_features / 8]")
    K, NW = kernel.shape
    return _repack(kernel, K, NW * PACK_FACTOR, _lib.REPACK_AWQ_TO_B200)


def quick_unrepack(kernel: torch.Tensor, in_features: int, out_features: int) -> torch.Tensor:
    """Inverse of ``quick_repack``."""
    return _repack(kernel, in_features, out_features, _lib.REPACK_B200_TO_AWQ)


def _norm_args(x: torch.Tensor, weight: torch.Tensor, *others: Optional[torch.Tensor]) -> Tuple[torch.device, int, int]:
    dev = _check_cuda(x, weight, *[t for t in others if t is not None])
    for t in (x, weight) + tuple(o for o in others if o is not None):
        if t.dtype != torch.float16:
            raise ValueError("the B200 RMSNorm kernels take float16 tensors only")
        if not t.is_contiguous():
            raise ValueError("RMSNorm tensors must be contiguous")
    hidden = x.shape[-1]
    if weight.numel() != hidden:
        raise ValueError("weight must have hidden_size elements")
    for o in others:
        if o is not None and o.shape != x.shape:
            raise ValueError("tensors of an RMSNorm call must have the same shape")
    return dev, (x.numel() // hidden if hidden else 0), hidden


def rms_norm(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    """``out = rms_norm(input) * weight`` -- the reference op ``ops.rms_norm(out, input,
    weight, epsilon)`` (csrc/layernorm_kernels.cu:78-102)."""
    dev, rows, hidden = _norm_args(input, weight, out)
    with torch.cuda.device(dev):
        st = _lib.lib().b200_w4a16_add_rms_norm(input.data_ptr(), None, weight.data_ptr(), out.data_ptr(),
                                          None, float(epsilon), rows, hidden, _stream_ptr(dev))
    _lib.check(st, "rms_norm")


def fused_add_rms_norm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                       epsilon: float) -> None:
    """In place: ``residual += input; input = rms_norm(residual) * weight`` -- the reference op
    ``ops.fused_add_rms_norm(input, residual, weight, epsilon)``
    (csrc/layernorm_kernels.cu:104-129)."""
    dev, rows, hidden = _norm_args(input, weight, residual)
    with torch.cuda.device(dev):
        st = _lib.lib().b200_w4a16_add_rms_norm(input.data_ptr(), residual.data_ptr(), weight.data_ptr(),
                                          input.data_ptr(), residual.data_ptr(), float(epsilon),
                                          rows, hidden, _stream_ptr(dev))
    _lib.check(st, "fused_add_rms_norm")


def _rows_ptr(rows: Optional[torch.Tensor], dev: torch.device) -> int:
    """Device int32 scalar holding a live row count (routed experts: known on the device only)."""
    if not rows.is_cuda or rows.device != dev or rows.dtype != torch.int32 or rows.numel() != 1:
        raise ValueError("rows must be a one-element int32 tensor on the same device")
    return rows.data_ptr()


def _silu(out: torch.Tensor, input: torch.Tensor, paired: int, rows: Optional[torch.Tensor] = None) -> None:
    dev = _check_cuda(out, input)
    if out.dtype != torch.float16 or input.dtype != torch.float16:
        raise ValueError("the B200 activation kernels take float16 tensors only")
    d = input.shape[-1] // 2
    if input.shape[-1] != 2 * d or out.shape != input.shape[:-1] + (d,):
        raise ValueError("silu_and_mul: out must be [..., d] for input [..., 2 * d]")
    n_rows = input.numel() // (2 * d) if d else 0
    with torch.cuda.device(dev):
        if rows is None:
            st = _lib.lib().b200_w4a16_silu_and_mul(input.data_ptr(), out.data_ptr(), n_rows, d, paired,
                                                    _stream_ptr(dev))
        else:
            st = _lib.lib().b200_w4a16_silu_and_mul_rows(input.data_ptr(), out.data_ptr(), n_rows,
                                                         _rows_ptr(rows, dev), d, paired, _stream_ptr(dev))
    _lib.check(st, "silu_and_mul")


def silu_and_mul(out: torch.Tensor, input: torch.Tensor) -> None:
    """``out = silu(input[..., :d]) * input[..., d:]`` -- the reference op
    ``ops.silu_and_mul(out, input)`` (csrc/activation_kernels.cu:17-52)."""
    _silu(out, input, 0)


def silu_and_mul_paired(out: torch.Tensor, input: torch.Tensor, rows: Optional[torch.Tensor] = None) -> None:
    """The same on the channel-paired order a paired gate_up weight produces (every 16
    channels = 8 gate + the 8 up channels of the same indices).  ``rows``: device int32 scalar,
    only that many leading rows are processed."""
    _silu(out, input, 1, rows)


def quick_gemm_silu_mul(in_feats: torch.Tensor, kernel: torch.Tensor, scaling_factors: torch.Tensor,
                        zeros: torch.Tensor) -> torch.Tensor:
    """``silu_and_mul(in_feats @ dequant(W))`` for a channel-paired, interleaved gate_up weight:
    one kernel for decode batches (``b200_w4a16_gemm_silu_mul``), GEMM + paired activation
    kernel above them.  Returns ``[M, N / 2]``."""
    M = in_feats.shape[0]
    N = scaling_factors.shape[1]
    K = in_feats.shape[1]
    G = K // scaling_factors.shape[0] if scaling_factors.shape[0] else 0
    L = _lib.lib()
    if M == 0 or G == 0 or G % 128 != 0 or L.b200_w4a16_gemm_variant(M, K, N, G) != 1:
        full = quick_gemm(in_feats, kernel, scaling_factors, zeros, PACK_FACTOR)
        out = torch.empty((M, N // 2), dtype=torch.float16, device=in_feats.device)
        silu_and_mul_paired(out, full)
        return out
    dev = _check_cuda(in_feats, kernel, scaling_factors, zeros)
    if in_feats.dtype != torch.float16 or scaling_factors.dtype != torch.float16:
        raise ValueError("in_feats and scaling_factors must be float16")
    if kernel.dtype != torch.int32 or zeros.dtype != torch.int32:
        raise ValueError("kernel and zeros must be int32")
    if kernel.numel() != L.b200_w4a16_repacked_words(K, N) or zeros.shape != (K // G, N // PACK_FACTOR):
        raise ValueError("kernel / zeros do not match the B200 layout for these shapes")
    out = torch.empty((M, N // 2), dtype=torch.float16, device=dev)
    with torch.cuda.device(dev):
        ws = _workspace(dev, L.b200_w4a16_workspace_bytes(M, K, N, LAYOUT_B200))
        flag = getattr(_gate, "flag", None)                   # routed-expert gate, see expert_gate()
        st = L.b200_w4a16_gemm_silu_mul_gated(in_feats.data_ptr(), kernel.data_ptr(),
                                              scaling_factors.data_ptr(), zeros.data_ptr(), out.data_ptr(),
                                              M, K, N, G, 0, flag.data_ptr() if flag is not None else None,
                                              ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    _lib.check(st, "w4a16 gemm + silu_and_mul")
    return out


def moe_gather(x: torch.Tensor, idx: torch.Tensor, rows: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``x[idx]`` for an int32 index vector: the tokens routed to one expert as a dense batch.
    ``rows`` (device int32 scalar): only ``idx[:rows]`` is valid -- the other rows of the result
    are left unwritten (``b200_w4a16_moe_gather_rows``)."""
    dev = _check_cuda(x, idx)
    if x.dtype != torch.float16 or idx.dtype != torch.int32 or x.dim() != 2 or idx.dim() != 1:
        raise ValueError("moe_gather: x [tokens, hidden] float16, idx [rows] int32")
    out = torch.empty((idx.numel(), x.shape[1]), dtype=torch.float16, device=dev)
    with torch.cuda.device(dev):
        if rows is None:
            st = _lib.lib().b200_w4a16_moe_gather(x.data_ptr(), idx.data_ptr(), out.data_ptr(),
                                                  idx.numel(), x.shape[1], _stream_ptr(dev))
        else:
            st = _lib.lib().b200_w4a16_moe_gather_rows(x.data_ptr(), idx.data_ptr(), _rows_ptr(rows, dev),
                                                       out.data_ptr(), idx.numel(), x.shape[1], _stream_ptr(dev))
    _lib.check(st, "moe_gather")
    return out


def moe_scatter_add(acc: torch.Tensor, y: torch.Tensor, idx: torch.Tensor, w: torch.Tensor,
                    rows: Optional[torch.Tensor] = None) -> None:
    """In place: ``acc[idx[i]] += half(y[i] * w[i])`` (w fp32; rows of idx distinct), for
    ``i < rows`` when a device-side row count is given."""
    dev = _check_cuda(acc, y, idx, w)
    if (acc.dtype != torch.float16 or y.dtype != torch.float16 or idx.dtype != torch.int32
            or w.dtype != torch.float32):
        raise ValueError("moe_scatter_add: acc / y float16, idx int32, w float32")
    if y.dim() != 2 or acc.dim() != 2 or y.shape[1] != acc.shape[1] or idx.numel() != y.shape[0] \
            or w.numel() != y.shape[0]:
        raise ValueError("moe_scatter_add: y [rows, hidden], idx [rows], w [rows], acc [tokens, hidden]")
    with torch.cuda.device(dev):
        if rows is None:
            st = _lib.lib().b200_w4a16_moe_scatter_add(y.data_ptr(), idx.data_ptr(), w.data_ptr(),
                                                       acc.data_ptr(), y.shape[0], y.shape[1],
                                                       _stream_ptr(dev))
        else:
            st = _lib.lib().b200_w4a16_moe_scatter_add_rows(y.data_ptr(), idx.data_ptr(), w.data_ptr(),
                                                            _rows_ptr(rows, dev), acc.data_ptr(), y.shape[0],
                                                            y.shape[1], _stream_ptr(dev))
    _lib.check(st, "moe_scatter_add")


def moe_grouped_gemm(x: torch.Tensor, kernels, scaling_factors, zeros, counts: torch.Tensor,
                     act: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All local experts' GEMMs of a mixture-of-experts block in ONE launch (decode batches,
    ``b200_w4a16_moe_grouped_gemm``).  ``kernels`` / ``scaling_factors`` / ``zeros``: one tensor
    per local expert (B200 layout, identical shapes).  ``x``: ``[M, K]`` (every expert reads it)
    or ``[E, M, K]`` (expert e reads ``x[e]``).  ``counts``: int32 ``[E]`` on the device, tokens
    routed to each local expert; experts with a zero count are skipped and their slice of the
    result is left unwritten.  Returns ``[E, M, N]`` (``[E, M, N / 2]`` with ``act``: SiLU-and-
    multiply over channel-paired gate / up weights in the epilogue)."""
    E = len(kernels)
    if E == 0 or len(scaling_factors) != E or len(zeros) != E:
        raise ValueError("moe_grouped_gemm: one kernel / scales / zeros tensor per expert")
    dev = _check_cuda(x, counts, *kernels, *scaling_factors, *zeros)
    per_expert_x = x.dim() == 3
    if (per_expert_x and x.shape[0] != E) or x.dim() not in (2, 3) or not x.is_contiguous():
        raise ValueError("moe_grouped_gemm: x must be contiguous [M, K] or [E, M, K]")
    M, K = x.shape[-2], x.shape[-1]
    N = scaling_factors[0].shape[1]
    G = K // scaling_factors[0].shape[0] if scaling_factors[0].shape[0] else 0
    if x.dtype != torch.float16 or counts.dtype != torch.int32 or counts.numel() != E:
        raise ValueError("moe_grouped_gemm: x float16, counts int32 [E]")
    L = _lib.lib()
    for k, s, z in zip(kernels, scaling_factors, zeros):
        if k.dtype != torch.int32 or z.dtype != torch.int32 or s.dtype != torch.float16:
            raise ValueError("kernel and zeros must be int32, scaling_factors float16")
        if s.shape != scaling_factors[0].shape or k.numel() != L.b200_w4a16_repacked_words(K, N) \
                or z.shape != (K // G if G else 0, N // PACK_FACTOR):
            raise ValueError("moe_grouped_gemm: the experts must share one shape (B200 layout)")
    NO = N // 2 if act else N
    if out is None:
        out = torch.empty((E, M, NO), dtype=torch.float16, device=dev)
    elif out.shape != (E, M, NO) or out.dtype != torch.float16 or not out.is_contiguous():
        raise ValueError("moe_grouped_gemm: out must be contiguous float16 [E, M, N']")
    vp = ctypes.c_void_p
    ptrs = [(vp * E)(*[t.data_ptr() for t in ts]) for ts in (kernels, scaling_factors, zeros)]
    with torch.cuda.device(dev):
        ws = _workspace(dev, L.b200_w4a16_workspace_bytes(M, K, N, LAYOUT_B200))
        st = L.b200_w4a16_moe_grouped_gemm(x.data_ptr(), ptrs[0], ptrs[1], ptrs[2], out.data_ptr(),
                                           M, K, N, G, E, counts.data_ptr(),
                                           M * K if per_expert_x else 0, 1 if act else 0,
                                           ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    _lib.check(st, "moe_grouped_gemm")
    return out


def moe_combine(y: torch.Tensor, w: torch.Tensor, counts: torch.Tensor) -> torch.Tensor:
    """``sum_e half(y[e] * w[:, e])`` over the experts with ``counts[e] > 0``, in index order with
    the reference's roundings (``mul_`` then ``add_`` per expert, mixtral_quant.py:152-157).
    ``y``: ``[E, T, hidden]`` float16, ``w``: fp32 ``[T, >= E]`` (column e = expert e's routing
    weight per token; may be a column slice of a wider matrix), ``counts``: int32 ``[E]``."""
    dev = _check_cuda(y, counts)
    if not w.is_cuda or w.device != dev:
        raise RuntimeError("all tensors must be on the same CUDA device")
    if y.dtype != torch.float16 or w.dtype != torch.float32 or counts.dtype != torch.int32:
        raise ValueError("moe_combine: y float16, w float32, counts int32")
    if y.dim() != 3 or w.dim() != 2 or w.shape[0] != y.shape[1] or w.shape[1] != y.shape[0] \
            or w.stride(1) != 1 or counts.numel() != y.shape[0] or not y.is_contiguous():
        raise ValueError("moe_combine: y [E, T, hidden], w [T, E] (unit column stride), counts [E]")
    out = torch.empty(y.shape[1:], dtype=torch.float16, device=dev)
    with torch.cuda.device(dev):
        st = _lib.lib().b200_w4a16_moe_combine(y.data_ptr(), w.data_ptr(), counts.data_ptr(),
                                               out.data_ptr(), y.shape[0], max(w.stride(0), y.shape[0]),
                                               y.shape[1], y.shape[2], _stream_ptr(dev))
    _lib.check(st, "moe_combine")
    return out


def moe_route(x: torch.Tensor, gate_weight: torch.Tensor, top_k: int):
    """The router in one launch (``b200_w4a16_moe_route``): fp16 gate GEMM, fp32 softmax, top-k,
    renormalise.  Returns ``(topk_weights [T, k] fp32, topk_experts [T, k] int64, counts [E]
    int32, dense_weights [T, E] fp32)`` -- the first two are what the reference computes
    (mixtral_quant.py:133-142), ``counts`` feeds the expert gates, ``dense_weights`` holds each
    token's weight for every expert (zero where not chosen)."""
    dev = _check_cuda(x, gate_weight)
    if x.dtype != torch.float16 or gate_weight.dtype != torch.float16 or x.dim() != 2 \
            or gate_weight.dim() != 2 or gate_weight.shape[1] != x.shape[1]:
        raise ValueError("moe_route: x [tokens, hidden] and gate_weight [experts, hidden] float16")
    T, E = x.shape[0], gate_weight.shape[0]
    topk_w = torch.empty((T, top_k), dtype=torch.float32, device=dev)
    topk_idx = torch.empty((T, top_k), dtype=torch.int64, device=dev)
    dense = torch.empty((T, E), dtype=torch.float32, device=dev)
    counts = torch.empty((E,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        st = _lib.lib().b200_w4a16_moe_route(x.data_ptr(), gate_weight.data_ptr(), topk_w.data_ptr(),
                                             topk_idx.data_ptr(), dense.data_ptr(), counts.data_ptr(),
                                             T, x.shape[1], E, top_k, _stream_ptr(dev))
    _lib.check(st, "moe_route")
    return topk_w, topk_idx, counts, dense


def quick_gemm_rows(in_feats: torch.Tensor, kernel: torch.Tensor, scaling_factors: torch.Tensor,
                    zeros: torch.Tensor, rows: torch.Tensor) -> torch.Tensor:
    """``quick_gemm`` on the first ``rows`` rows of ``in_feats`` where ``rows`` is a device int32
    scalar (``b200_w4a16_gemm_rows``): the work follows the live count without the host ever
    reading it.  Returns ``[M, N]``; rows beyond the count are unspecified."""
    dev = _check_cuda(in_feats, kernel, scaling_factors, zeros)
    if in_feats.dtype != torch.float16 or scaling_factors.dtype != torch.float16:
        raise ValueError("in_feats and scaling_factors must be float16")
    if kernel.dtype != torch.int32 or zeros.dtype != torch.int32 or in_feats.dim() != 2:
        raise ValueError("kernel and zeros must be int32, in_feats [M, K]")
    M, K = in_feats.shape
    N = scaling_factors.shape[1]
    G = K // scaling_factors.shape[0] if scaling_factors.shape[0] else 0
    L = _lib.lib()
    if G == 0 or kernel.numel() != L.b200_w4a16_repacked_words(K, N) or zeros.shape != (K // G, N // PACK_FACTOR):
        raise ValueError("kernel / zeros do not match the B200 layout for these shapes")
    out = torch.empty((M, N), dtype=torch.float16, device=dev)
    with torch.cuda.device(dev):
        ws = _workspace(dev, L.b200_w4a16_workspace_bytes(M, K, N, LAYOUT_B200))
        st = L.b200_w4a16_gemm_rows(in_feats.data_ptr(), kernel.data_ptr(), scaling_factors.data_ptr(),
                                    zeros.data_ptr(), out.data_ptr(), M, _rows_ptr(rows, dev), K, N, G,
                                    ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    _lib.check(st, "w4a16 gemm (device-side row count)")
    return out


def moe_compact(topk_experts: torch.Tensor, topk_weights: torch.Tensor, first_expert: int, n_local: int):
    """Per local expert, the tokens that chose it (ascending) and their routing weights, from
    the router's ``[T, k]`` lists: ``(tok [n_local, T] int32, w [n_local, T] fp32)``, row e valid
    up to the expert's token count (``b200_w4a16_moe_compact``; no host synchronisation)."""
    dev = _check_cuda(topk_experts, topk_weights)
    if topk_experts.dtype != torch.int64 or topk_weights.dtype != torch.float32 \
            or topk_experts.dim() != 2 or topk_experts.shape != topk_weights.shape:
        raise ValueError("moe_compact: topk_experts int64 [T, k], topk_weights float32 [T, k]")
    T, k = topk_experts.shape
    tok = torch.empty((n_local, T), dtype=torch.int32, device=dev)
    w = torch.empty((n_local, T), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        st = _lib.lib().b200_w4a16_moe_compact(topk_experts.data_ptr(), topk_weights.data_ptr(), tok.data_ptr(),
                                               w.data_ptr(), T, k, first_expert, n_local, _stream_ptr(dev))
    _lib.check(st, "moe_compact")
    return tok, w
