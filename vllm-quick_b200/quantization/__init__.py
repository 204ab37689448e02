"""Quantization registry (reference: vllm/model_executor/layers/quantization/__init__.py:8-18),
with the ``"quick"`` entry the north star adds."""
from typing import Type

from .awq import AWQConfig, AWQLinearMethod
from .base_config import QuantizationConfig
from .quick import QuickConfig, QuickLinearMethod

_QUANTIZATION_CONFIG_REGISTRY = {
    "awq": AWQConfig,
    "quick": QuickConfig,
}


def get_quantization_config(quantization: str) -> Type[QuantizationConfig]:
    if quantization not in _QUANTIZATION_CONFIG_REGISTRY:
        raise ValueError(f"Invalid quantization method: {quantization}")
    return _QUANTIZATION_CONFIG_REGISTRY[quantization]


__all__ = ["QuantizationConfig", "get_quantization_config", "AWQConfig", "AWQLinearMethod",
           "QuickConfig", "QuickLinearMethod"]
