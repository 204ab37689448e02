"""QuantizationConfig ABC (mirror of the reference's
vllm/model_executor/layers/quantization/base_config.py:9-64)."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Dict, List

import torch

from ..linear import LinearMethodBase


class QuantizationConfig(ABC):

    @abstractmethod
    def get_name(self) -> str: ...

    @abstractmethod
    def get_supported_act_dtypes(self) -> List[torch.dtype]: ...

    @abstractmethod
    def get_min_capability(self) -> int: ...

    @staticmethod
    @abstractmethod
    def get_config_filenames() -> List[str]: ...

    @classmethod
    @abstractmethod
    def from_config(cls, config: Dict[str, Any]) -> "QuantizationConfig": ...

    @staticmethod
    def get_from_keys(config: Dict[str, Any], keys: List[str]) -> Any:
        for key in keys:
            if key in config:
                return config[key]
        raise ValueError(f"Cannot find any of {keys} in the model's quantization config.")

    @abstractmethod
    def get_linear_method(self) -> LinearMethodBase: ...

    @abstractmethod
    def get_scaled_act_names(self) -> List[str]: ...
