"""``quantization="quick"``: the QUICK flavour of the AWQ plugin.

The QUICK repository's own offline interleave is not part of the mounted reference
(SURVEY.md F1), so this method consumes AWQ-format checkpoints
(qweight / qzeros / scales, awq.py:95-141 of the reference) and produces the B200 operand
interleave itself after loading -- either eagerly with ``process_weights_after_loading`` or
lazily on the first ``apply_weights``.  Everything else is inherited from the AWQ plugin.
"""
from __future__ import annotations

from typing import Any, Dict, List

import torch
from torch.nn.parameter import Parameter

from .. import ops
from ..linear import set_weight_attrs
from .awq import LAYOUT_INTERLEAVED, AWQConfig, AWQLinearMethod

# qweight of a pre-interleaved checkpoint: [N / 128, K / 128, 2048] int32, tile (nt, kb) of the
# B200 layout (include/b200_w4a16.h) at [nt, kb, :].  A tile holds a 128 x 128 block of the
# matrix and nothing else, so tensor-parallel shards are plain narrows: output features along
# dim 0, input features along dim 1, both in units of 128.
TILE = 128
TILE_WORDS = 2048


class QuickConfig(AWQConfig):
    """``version``: "gemm" (default) = AWQ checkpoint tensors, interleaved after loading;
    "quick" = the checkpoint's qweight tensors are already interleaved (written by
    ``convert_awq_state_dict``), loaded shard by shard with no repack at start-up."""

    def __init__(self, weight_bits: int, group_size: int, zero_point: bool,
                 version: str = "gemm") -> None:
        super().__init__(weight_bits, group_size, zero_point)
        self.version = str(version).lower()
        if self.version not in ("gemm", "quick"):
            raise ValueError(f"unknown QUICK checkpoint version {version!r} (gemm | quick)")

    @property
    def pre_interleaved(self) -> bool:
        return self.version == "quick"

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
        version = config.get("version", "gemm")
        return cls(weight_bits, group_size, zero_point, version)

    def get_linear_method(self) -> "QuickLinearMethod":
        return QuickLinearMethod(self)


class QuickLinearMethod(AWQLinearMethod):

    def create_weights(self, input_size_per_partition: int, output_size_per_partition: int,
                       input_size: int, output_size: int,
                       params_dtype: torch.dtype) -> Dict[str, Any]:
        weights = super().create_weights(input_size_per_partition, output_size_per_partition,
                                         input_size, output_size, params_dtype)
        if not getattr(self.quant_config, "pre_interleaved", False):
            return weights
        if input_size_per_partition % TILE != 0 or output_size_per_partition % TILE != 0:
            raise ValueError("A pre-interleaved QUICK checkpoint needs per-partition sizes that "
                             "are multiples of 128. This can be caused by too large tensor "
                             "parallel size.")
        qweight = Parameter(torch.empty(output_size_per_partition // TILE,
                                        input_size_per_partition // TILE, TILE_WORDS,
                                        dtype=torch.int32), requires_grad=False)
        # loaders narrow dim 0 by output features // 128 and dim 1 by whole K blocks
        set_weight_attrs(qweight, {"input_dim": 1, "output_dim": 0, "packed_dim": 0,
                                   "pack_factor": TILE})
        weights["qweight"] = qweight
        weights["qweight_layout"] = LAYOUT_INTERLEAVED
        return weights

    def process_weights_after_loading(self, weights: Dict[str, Any]) -> None:
        """Interleave now instead of on the first forward (e.g. before CUDA-graph capture)."""
        self._interleave_once(weights)


def convert_awq_state_dict(state_dict: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """AWQ checkpoint tensors -> the pre-interleaved format (``"version": "quick"``): every
    ``*.qweight [K, N/8]`` becomes ``[N/128, K/128, 2048]`` (interleaved on the GPU by
    ``ops.quick_repack``); ``qzeros`` / ``scales`` and everything else pass through.  Tensors
    whose shapes are not multiples of 128 are left in AWQ form (and such a model must be served
    with version "gemm")."""
    out: Dict[str, torch.Tensor] = {}
    for name, t in state_dict.items():
        if name.endswith("qweight") and t.dim() == 2 and t.dtype == torch.int32:
            K, N = t.shape[0], t.shape[1] * 8
            if K % TILE == 0 and N % TILE == 0:
                dev = t.device
                packed = ops.quick_repack(t.cuda().contiguous())
                out[name] = packed.reshape(N // TILE, K // TILE, TILE_WORDS).to(dev)
                continue
        out[name] = t
    return out
