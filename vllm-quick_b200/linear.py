"""Plugin interface of the linear layers (mirror of the reference's
vllm/model_executor/layers/linear.py:20-37 and vllm/model_executor/utils.py:set_weight_attrs).
The tensor-parallel layer classes that call it live in ``parallel_linear.py``."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Dict, Optional

import torch


def set_weight_attrs(weight: torch.Tensor, weight_attrs: Optional[Dict[str, Any]]) -> None:
    """Attach loader metadata (input_dim / output_dim / packed_dim / pack_factor /
    weight_loader) to a parameter; refuses to overwrite, like the reference."""
    if not weight_attrs:
        return
    for key, value in weight_attrs.items():
        if hasattr(weight, key):
            raise AssertionError(f"Overwriting existing tensor attribute: {key}")
        setattr(weight, key, value)


class LinearMethodBase(ABC):
    """How a linear layer creates and applies its (maybe quantized) weights."""

    @abstractmethod
    def create_weights(self, input_size_per_partition: int, output_size_per_partition: int,
                       input_size: int, output_size: int,
                       params_dtype: torch.dtype) -> Dict[str, Any]:
        raise NotImplementedError

    @abstractmethod
    def apply_weights(self, weights: Dict[str, Any], x: torch.Tensor,
                      bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError
