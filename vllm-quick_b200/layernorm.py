"""RMSNorm on the B200 library -- mirror of the reference's
vllm/model_executor/layers/layernorm.py:9-66 (same constructor, same ``forward(x, residual)``
contract: with a residual the op works IN PLACE on both tensors and returns them).

``forward_reduced`` is the addition: called with the un-reduced output of a row-parallel
linear (``RowParallelLinear(..., reduce_results=False)``) it runs the tensor-parallel
all-reduce, the residual add and the norm as ONE kernel over NVLink peer memory
(``b200_w4a16_allreduce_add_rms_norm``), replacing the pair
``tensor_model_parallel_all_reduce`` (layers/linear.py:541-542) + ``fused_add_rms_norm``
(models/llama.py:205-222).  No PyTorch fallback for the norm itself: without the CUDA library
every call raises.
"""
from __future__ import annotations

from typing import Optional, Tuple, Union

import torch
import torch.nn as nn

from . import ops
from .parallel_utils import custom_all_reduce
from .parallel_utils.communication_op import tensor_model_parallel_all_reduce
from .parallel_utils.parallel_state import get_tensor_model_parallel_world_size


class RMSNorm(nn.Module):
    """x -> w * x / sqrt(E[x^2] + eps) (layernorm.py:9-24)."""

    def __init__(self, hidden_size: int, eps: float = 1e-6) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.variance_epsilon = eps

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None
                ) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
        if residual is not None:
            ops.fused_add_rms_norm(x, residual, self.weight.data, self.variance_epsilon)
            return x, residual
        out = torch.empty_like(x)
        ops.rms_norm(out, x, self.weight.data, self.variance_epsilon)
        return out

    def forward_reduced(self, partial: torch.Tensor, residual: torch.Tensor
                        ) -> Tuple[torch.Tensor, torch.Tensor]:
        """``forward(all_reduce(partial), residual)`` in one kernel when the peer-memory
        all-reduce is up and the message fits; otherwise the two steps one after the other."""
        if get_tensor_model_parallel_world_size() == 1:
            return self.forward(partial, residual)
        h = custom_all_reduce.norm_handle(partial)
        if h is None:
            return self.forward(tensor_model_parallel_all_reduce(partial), residual)
        out = h.all_reduce_add_rms_norm(partial, residual, self.weight.data, self.variance_epsilon)
        return out, residual
