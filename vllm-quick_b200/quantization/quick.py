"""``quantization="quick"``: the QUICK flavour of the AWQ plugin.

The QUICK repository's own offline interleave is not part of the mounted reference
(SURVEY.md F1), so this method consumes AWQ-format checkpoints
(qweight / qzeros / scales, awq.py:95-141 of the reference) and produces the B200 operand
interleave itself after loading -- either eagerly with ``process_weights_after_loading`` or
lazily on the first ``apply_weights``.  Everything else is inherited from the AWQ plugin.
"""
from __future__ import annotations

from typing import Any, Dict, List

from .awq import AWQConfig, AWQLinearMethod


class QuickConfig(AWQConfig):

    def get_name(self) -> str:
        return "quick"

    @staticmethod
    def get_config_filenames() -> List[str]:
        return ["quant_config.json", "quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "QuickConfig":
        weight_bits = cls.get_from_keys(config, ["w_bit", "bits"])
        group_size = cls.get_from_keys(config, ["q_group_size", "group_size"])
        zero_point = cls.get_from_keys(config, ["zero_point"])
        return cls(weight_bits, group_size, zero_point)

    def get_linear_method(self) -> "QuickLinearMethod":
        return QuickLinearMethod(self)


class QuickLinearMethod(AWQLinearMethod):

    def process_weights_after_loading(self, weights: Dict[str, Any]) -> None:
        """Interleave now instead of on the first forward (e.g. before CUDA-graph capture)."""
        self._interleave_once(weights)
