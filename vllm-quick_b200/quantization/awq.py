"""AWQ config and linear method on the B200 W4A16 kernels.

Mirrors the reference's plugin (vllm/model_executor/layers/quantization/awq.py:12-165):
same config keys, same parameter names / shapes / loader attributes, same ``ValueError``s,
so vLLM's tensor-parallel linear layers and checkpoint loaders work unmodified.  What
differs is below the operator boundary: one GEMM kernel serves every token count (the
reference switches to dequantize + cuBLAS at >= 256 tokens, awq.py:154-159), and the packed
weight is interleaved into the B200 operand layout once, on the first ``apply_weights``
(the reference's GPTQ method sets the precedent for a post-load shuffle,
vllm/model_executor/layers/quantization/gptq.py:195-204).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn.parameter import Parameter

from .. import ops
from ..linear import LinearMethodBase, set_weight_attrs
from .base_config import QuantizationConfig

# state of the packed weight held in the weights dict (a non-tensor entry, ignored by
# parameter registration -- linear.py:110-112 of the reference)
LAYOUT_CHECKPOINT = "awq"
LAYOUT_INTERLEAVED = "b200"


class AWQConfig(QuantizationConfig):
    """``weight_bits`` must be 4; ``pack_factor`` = 8 nibbles per int32."""

    def __init__(self, weight_bits: int, group_size: int, zero_point: bool) -> None:
        self.weight_bits = weight_bits
        self.group_size = group_size
        self.zero_point = zero_point
        if self.weight_bits != 4:
            raise ValueError("Currently, only 4-bit weight quantization is supported for "
                             f"AWQ, but got {self.weight_bits} bits.")
        self.pack_factor = 32 // self.weight_bits

    def __repr__(self) -> str:
        return (f"{type(self).__name__}(weight_bits={self.weight_bits}, "
                f"group_size={self.group_size}, zero_point={self.zero_point})")

    def get_name(self) -> str:
        return "awq"

    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        return [torch.half]

    def get_min_capability(self) -> int:
        return 100  # sm_100a kernels only (the reference's kernel needs 75)

    @staticmethod
    def get_config_filenames() -> List[str]:
        return ["quant_config.json", "quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "AWQConfig":
        weight_bits = cls.get_from_keys(config, ["w_bit", "bits"])
        group_size = cls.get_from_keys(config, ["q_group_size", "group_size"])
        zero_point = cls.get_from_keys(config, ["zero_point"])
        return cls(weight_bits, group_size, zero_point)

    def get_linear_method(self) -> "AWQLinearMethod":
        return AWQLinearMethod(self)

    def get_scaled_act_names(self) -> List[str]:
        return ["gelu", "gelu_fast", "gelu_new", "gelu_pytorch_tanh"]


class AWQLinearMethod(LinearMethodBase):
    """create_weights / apply_weights for W4 g-quantized linears."""

    def __init__(self, quant_config: AWQConfig):
        self.quant_config = quant_config

    def create_weights(self, input_size_per_partition: int, output_size_per_partition: int,
                       input_size: int, output_size: int,
                       params_dtype: torch.dtype) -> Dict[str, Any]:
        cfg = self.quant_config
        if input_size_per_partition % cfg.group_size != 0:
            raise ValueError("The input size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        if output_size_per_partition % cfg.pack_factor != 0:
            raise ValueError("The output size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        packed = {"input_dim": 0, "output_dim": 1, "packed_dim": 1,
                  "pack_factor": cfg.pack_factor}
        groups = input_size_per_partition // cfg.group_size
        qweight = Parameter(torch.empty(input_size_per_partition,
                                        output_size_per_partition // cfg.pack_factor,
                                        dtype=torch.int32), requires_grad=False)
        set_weight_attrs(qweight, dict(packed))
        qzeros = Parameter(torch.empty(groups, output_size_per_partition // cfg.pack_factor,
                                       dtype=torch.int32), requires_grad=False)
        set_weight_attrs(qzeros, dict(packed))
        scales = Parameter(torch.empty(groups, output_size_per_partition, dtype=params_dtype),
                           requires_grad=False)
        set_weight_attrs(scales, {"input_dim": 0, "output_dim": 1})
        return {"qweight": qweight, "qzeros": qzeros, "scales": scales,
                "qweight_layout": LAYOUT_CHECKPOINT}

    @staticmethod
    def _interleave_once(weights: Dict[str, Any]) -> torch.Tensor:
        """Swap the checkpoint-layout words for the B200 interleave, in place (1x memory)."""
        qweight = weights["qweight"]
        if weights.get("qweight_layout", LAYOUT_CHECKPOINT) == LAYOUT_CHECKPOINT:
            qweight.data = ops.quick_repack(qweight.data.contiguous())
            weights["qweight_layout"] = LAYOUT_INTERLEAVED
            # load time, never inside a capture: the scratch later CUDA-graph captures will use
            ops.ensure_workspace(qweight.device)
        return qweight

    def pair_gate_up(self, weights: Dict[str, Any]) -> None:
        """Channel-pair a merged [gate | up] projection for the fused SiLU-and-multiply
        epilogue: packed word column 2 j = gate word j, 2 j + 1 = up word j (a permutation of
        whole int32 columns of qweight / qzeros and of 8-column blocks of scales, done once on
        the checkpoint layout, before the interleave).  Idempotent."""
        if weights.get("gate_up_paired", False):
            return
        qweight, qzeros, scales = weights["qweight"], weights["qzeros"], weights["scales"]
        if weights.get("qweight_layout", LAYOUT_CHECKPOINT) != LAYOUT_CHECKPOINT:
            qweight.data = ops.quick_unrepack(qweight.data, qweight.shape[0] if qweight.dim() == 2
                                              else scales.shape[0] * self.quant_config.group_size,
                                              scales.shape[1])
            weights["qweight_layout"] = LAYOUT_CHECKPOINT
        words = scales.shape[1] // self.quant_config.pack_factor
        if words % 2 != 0:
            raise ValueError("gate and up projections must have the same, 8-aligned, width")
        half = torch.arange(words // 2, device=qweight.device)
        order = torch.stack([half, half + words // 2], dim=1).reshape(-1)
        qweight.data = qweight.data.reshape(-1, words)[:, order].contiguous()
        qzeros.data = qzeros.data[:, order].contiguous()
        groups = scales.shape[0]
        scales.data = scales.data.reshape(groups, words, -1)[:, order].reshape(groups, -1).contiguous()
        weights["gate_up_paired"] = True

    def apply_weights_silu_and_mul(self, weights: Dict[str, Any], x: torch.Tensor) -> torch.Tensor:
        """``SiluAndMul(apply_weights(weights, x))`` for a merged gate_up projection
        (llama.py:76-80), the activation in the GEMM's epilogue for decode batches."""
        self.pair_gate_up(weights)
        qweight = self._interleave_once(weights)
        scales = weights["scales"]
        reshaped_x = x.reshape(-1, x.shape[-1])
        out = ops.quick_gemm_silu_mul(reshaped_x, qweight, scales, weights["qzeros"])
        return out.reshape(x.shape[:-1] + (scales.shape[-1] // 2, ))

    def apply_weights_all_reduce(self, weights: Dict[str, Any],
                                 x: torch.Tensor) -> Optional[torch.Tensor]:
        """Row-parallel hook: GEMM on this rank's K shard fused with the tensor-parallel
        all-reduce (one kernel; decode batches).  Returns None when the fused kernel does not
        apply -- the caller then runs ``apply_weights`` + ``tensor_model_parallel_all_reduce``
        as the reference does (linear.py:538-542)."""
        from ..parallel_utils import custom_all_reduce
        scales = weights["scales"]
        reshaped_x = x.reshape(-1, x.shape[-1])
        if scales.shape[0] == 0:
            return None
        handle = custom_all_reduce.fused_handle(reshaped_x.shape[0], scales.shape[-1],
                                                reshaped_x.shape[1] // scales.shape[0])
        if handle is None:
            return None
        qweight = self._interleave_once(weights)
        out = ops.quick_gemm_all_reduce(reshaped_x, qweight, scales, weights["qzeros"], handle)
        return out.reshape(x.shape[:-1] + (scales.shape[-1], ))

    def apply_weights(self, weights: Dict[str, Any], x: torch.Tensor,
                      bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        scales = weights["scales"]
        qzeros = weights["qzeros"]
        if weights.get("gate_up_paired", False):
            raise RuntimeError("this gate_up weight was channel-paired for the fused activation: "
                               "call apply_weights_silu_and_mul / forward_silu_and_mul")
        qweight = self._interleave_once(weights)
        pack_factor = self.quant_config.pack_factor
        out_shape = x.shape[:-1] + (scales.shape[-1], )
        reshaped_x = x.reshape(-1, x.shape[-1])
        out = ops.quick_gemm(reshaped_x, qweight, scales, qzeros, pack_factor)
        if bias is not None:
            out = out + bias
        return out.reshape(out_shape)
