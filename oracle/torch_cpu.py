"""The reference's CPU path for this operator, as BASELINE.json's north_star defines it:
"PyTorch dequantize + matmul" of the same packed tensors, on the host cores.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/oracle.py): used by bench.py's ``cpu_baseline``
leg and ``--impl reference`` arm, never by the product package.  The arithmetic restates
csrc/quantization/awq/dequantize.cuh:17-84 (nibble order) and gemm_kernels.cu:153-160
(``fp16((q - z) * s)``) with torch CPU ops, then ``x.float() @ W.float()`` as in
vllm/model_executor/layers/quantization/awq.py:158-159 (dequantize, then matmul).
"""
from __future__ import annotations

import torch

AWQ_ORDER = (0, 4, 1, 5, 2, 6, 3, 7)


def unpack(packed: torch.Tensor) -> torch.Tensor:
    """int32 [R, C] -> uint8-valued int32 [R, C*8] in logical order."""
    shifts = torch.tensor([4 * o for o in AWQ_ORDER], dtype=torch.int32)
    return ((packed.unsqueeze(-1) >> shifts) & 0xF).reshape(packed.shape[0], -1)


def dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    K = qweight.shape[0]
    G = K // scales.shape[0]
    q = unpack(qweight).view(K // G, G, -1)
    z = unpack(qzeros).unsqueeze(1)
    w = (q - z).to(torch.float32) * scales.to(torch.float32).unsqueeze(1)
    return w.to(torch.float16).view(K, -1)


def awq_gemm(x: torch.Tensor, qweight: torch.Tensor, scales: torch.Tensor,
             qzeros: torch.Tensor) -> torch.Tensor:
    W = dequantize(qweight, scales, qzeros)
    return (x.to(torch.float32) @ W.to(torch.float32)).to(torch.float16)
