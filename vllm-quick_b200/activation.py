"""SwiGLU activation and the Llama MLP block on the B200 library -- mirrors of
vllm/model_executor/layers/activation.py:14-38 (``SiluAndMul``) and
vllm/model_executor/models/llama.py:48-80 (``LlamaMLP``).

``LlamaMLP(..., fuse_activation=True)`` (default) runs ``gate_up_proj`` and ``act_fn`` as one
call (``MergedColumnParallelLinear.forward_silu_and_mul``): for decode batches the activation
sits in the W4A16 GEMM's epilogue and the ``[tokens, 2 * intermediate]`` tensor never exists.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import ops
from .linear import LinearMethodBase
from .parallel_linear import MergedColumnParallelLinear, RowParallelLinear


class SiluAndMul(nn.Module):
    """x -> silu(x[..., :d]) * x[..., d:], d = x.shape[-1] // 2 (activation.py:14-38)."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        d = x.shape[-1] // 2
        out = torch.empty(x.shape[:-1] + (d, ), dtype=x.dtype, device=x.device)
        ops.silu_and_mul(out, x)
        return out


class LlamaMLP(nn.Module):

    def __init__(self, hidden_size: int, intermediate_size: int, hidden_act: str = "silu",
                 linear_method: Optional[LinearMethodBase] = None,
                 params_dtype: Optional[torch.dtype] = None, fuse_activation: bool = True) -> None:
        super().__init__()
        self.gate_up_proj = MergedColumnParallelLinear(hidden_size, [intermediate_size] * 2, bias=False,
                                                       linear_method=linear_method,
                                                       params_dtype=params_dtype)
        self.down_proj = RowParallelLinear(intermediate_size, hidden_size, bias=False,
                                           linear_method=linear_method, params_dtype=params_dtype)
        if hidden_act != "silu":
            raise ValueError(f"Unsupported activation: {hidden_act}. "
                             "Only silu is supported for now.")
        self.act_fn = SiluAndMul()
        self.fuse_activation = fuse_activation and hasattr(linear_method, "apply_weights_silu_and_mul")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.fuse_activation:
            x = self.gate_up_proj.forward_silu_and_mul(x)
        else:
            gate_up, _ = self.gate_up_proj(x)
            x = self.act_fn(gate_up)
        x, _ = self.down_proj(x)
        return x
