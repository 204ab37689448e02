"""Tensor-parallel linear layers that call the (quantized) linear method.

Same classes, constructor arguments, parameter attributes and ``weight_loader`` behaviour as
the reference (vllm/model_executor/layers/linear.py:74-552), so a model definition and a
checkpoint loader written against the reference work unchanged:

* ``ReplicatedLinear``            :74-124   no sharding
* ``ColumnParallelLinear``        :127-215  output features split over the TP ranks
* ``MergedColumnParallelLinear``  :218-309  several column-parallel matrices fused (gate_up)
* ``QKVParallelLinear``           :312-437  q / k / v fused, k/v heads replicated when
                                            tp_size >= num_kv_heads
* ``RowParallelLinear``           :440-552  input features split; all-reduce after the GEMM

How the loaders are written here: every layer describes a checkpoint tensor as a list of
*pieces* ``(offset in the parameter, piece size, which slice of the checkpoint piece this
rank takes)`` along the sharded dimension, and one helper copies them.  Sizes and offsets
along a packed dimension (AWQ ``qweight`` / ``qzeros``: eight output features per int32,
``packed_dim == output_dim``) are divided by ``pack_factor``, exactly as the reference
does (:276-278, :293-295, :393-395, :417-419).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
from torch.nn.parameter import Parameter

from .linear import LinearMethodBase, set_weight_attrs
from .parallel_utils.communication_op import (split_tensor_along_last_dim,
                                              tensor_model_parallel_all_gather,
                                              tensor_model_parallel_all_reduce)
from .parallel_utils.parallel_state import (get_tensor_model_parallel_rank,
                                            get_tensor_model_parallel_world_size)


def divide(numerator: int, denominator: int) -> int:
    if numerator % denominator != 0:
        raise AssertionError(f"{numerator} is not divisible by {denominator}")
    return numerator // denominator


def _packed(param: torch.Tensor, dim: Optional[int], value: int) -> int:
    """``value`` output features expressed in elements of ``param`` along ``dim``."""
    if dim is not None and getattr(param, "packed_dim", None) == dim:
        if value % param.pack_factor != 0:
            raise AssertionError(f"{value} output features do not fall on a boundary of the "
                                 f"packed dimension (pack factor {param.pack_factor})")
        return value // param.pack_factor
    return value


def _copy_slice(param: torch.Tensor, loaded: torch.Tensor, dim: Optional[int],
                dst_offset: int, size: int, src_index: int) -> None:
    """param[dst_offset : dst_offset+size] <- loaded[src_index*size : (src_index+1)*size]
    along ``dim`` (whole-tensor copy when the parameter has no such dimension)."""
    dst = param.data
    if dim is not None:
        if dst_offset + size > dst.shape[dim] or (src_index + 1) * size > loaded.shape[dim]:
            raise AssertionError(f"shard [{src_index * size}, {(src_index + 1) * size}) does not fit "
                                 f"a checkpoint tensor of size {loaded.shape[dim]} along dim {dim}")
        dst = dst.narrow(dim, dst_offset, size)
        loaded = loaded.narrow(dim, src_index * size, size)
    if dst.shape != loaded.shape:
        raise AssertionError(f"shape mismatch loading a shard: parameter slice {tuple(dst.shape)} "
                             f"vs checkpoint slice {tuple(loaded.shape)}")
    dst.copy_(loaded)


class _LinearBase(torch.nn.Module):
    """Owns the weights dict of the linear method and registers its tensors as parameters."""

    def _setup(self, linear_method: LinearMethodBase, in_pp: int, out_pp: int, in_full: int,
               out_full: int, params_dtype: Optional[torch.dtype], loader) -> None:
        if linear_method is None:
            raise ValueError("a linear method is required: this package ships the quantized "
                             "W4A16 path only (the reference's UnquantizedLinearMethod, "
                             "linear.py:40-71, is outside the hot path)")
        self.params_dtype = params_dtype if params_dtype is not None else torch.get_default_dtype()
        self.linear_method = linear_method
        self.linear_weights: Dict[str, Any] = linear_method.create_weights(
            in_pp, out_pp, in_full, out_full, self.params_dtype)
        for name, weight in self.linear_weights.items():
            if isinstance(weight, torch.Tensor):      # non-tensor entries are method state
                self.register_parameter(name, weight)
                if loader is not None:
                    set_weight_attrs(weight, {"weight_loader": loader})

    def _make_bias(self, size: int, loader) -> None:
        self.bias = Parameter(torch.empty(size, dtype=self.params_dtype), requires_grad=False)
        attrs = {"output_dim": 0}
        if loader is not None:
            attrs["weight_loader"] = loader
        set_weight_attrs(self.bias, attrs)


class ReplicatedLinear(_LinearBase):
    """Every rank holds the whole matrix (Mixtral experts, linear.py:74-124 of the reference)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True,
                 skip_bias_add: bool = False, params_dtype: Optional[torch.dtype] = None,
                 linear_method: Optional[LinearMethodBase] = None):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.skip_bias_add = skip_bias_add
        self._setup(linear_method, input_size, output_size, input_size, output_size,
                    params_dtype, loader=None)
        if bias:
            self._make_bias(output_size, loader=None)
        else:
            self.register_parameter("bias", None)

    def forward(self, x: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        fused_bias = None if self.skip_bias_add else self.bias
        out = self.linear_method.apply_weights(self.linear_weights, x, fused_bias)
        return out, (self.bias if self.skip_bias_add else None)


class ColumnParallelLinear(_LinearBase):
    """Y = X [A_1 .. A_p]: rank r holds A_r; no communication unless ``gather_output``."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True,
                 gather_output: bool = False, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None,
                 linear_method: Optional[LinearMethodBase] = None):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.gather_output = gather_output
        self.skip_bias_add = skip_bias_add
        tp_size = get_tensor_model_parallel_world_size()
        self.output_size_per_partition = divide(output_size, tp_size)
        self._setup(linear_method, input_size, self.output_size_per_partition, input_size,
                    output_size, params_dtype, loader=self.weight_loader)
        if bias:
            self._make_bias(self.output_size_per_partition, loader=self.weight_loader)
        else:
            self.register_parameter("bias", None)

    # A fused layer is a sequence of (piece id, full-model size in output features); the plain
    # column-parallel layer is the one-piece case.
    def _pieces(self) -> Sequence[Tuple[Any, int]]:
        return [(None, self.output_size)]

    def _piece_plan(self, piece_id) -> Tuple[int, int, int]:
        """(offset in this rank's parameter, size per rank, index of the checkpoint slice),
        all in output features."""
        return 0, self.output_size_per_partition, get_tensor_model_parallel_rank()

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor,
                      loaded_shard_id=None) -> None:
        output_dim = getattr(param, "output_dim", None)
        pieces = self._pieces()
        if loaded_shard_id is None and len(pieces) > 1:
            # an already-fused checkpoint tensor: cut it into its pieces and load each
            if output_dim is None:
                _copy_slice(param, loaded_weight, None, 0, 0, 0)
                return
            start = 0
            for piece_id, full in pieces:
                piece = loaded_weight.narrow(output_dim, _packed(param, output_dim, start),
                                             _packed(param, output_dim, full))
                self.weight_loader(param, piece, piece_id)
                start += full
            return
        if output_dim is None:
            # no output dimension (e.g. a per-tensor scalar): identical on every rank
            _copy_slice(param, loaded_weight, None, 0, 0, 0)
            return
        offset, size, index = self._piece_plan(loaded_shard_id)
        _copy_slice(param, loaded_weight, output_dim, _packed(param, output_dim, offset),
                    _packed(param, output_dim, size), index)

    def forward(self, input_: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        fused_bias = None if self.skip_bias_add else self.bias
        out = self.linear_method.apply_weights(self.linear_weights, input_, fused_bias)
        if self.gather_output:
            out = tensor_model_parallel_all_gather(out)
        return out, (self.bias if self.skip_bias_add else None)


class MergedColumnParallelLinear(ColumnParallelLinear):
    """Several column-parallel matrices sharing one input, fused along the output dimension
    (Llama ``gate_up_proj``); each is sharded on its own, then the shards are concatenated."""

    def __init__(self, input_size: int, output_sizes: List[int], bias: bool = True,
                 gather_output: bool = False, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None,
                 linear_method: Optional[LinearMethodBase] = None):
        self.output_sizes = list(output_sizes)
        tp_size = get_tensor_model_parallel_world_size()
        if any(size % tp_size != 0 for size in self.output_sizes):
            raise AssertionError("every merged output size must be divisible by the TP size")
        super().__init__(input_size, sum(self.output_sizes), bias, gather_output, skip_bias_add,
                         params_dtype, linear_method)

    def _pieces(self):
        return list(enumerate(self.output_sizes))

    def forward_silu_and_mul(self, input_: torch.Tensor) -> torch.Tensor:
        """``SiluAndMul()(self(input_)[0])`` -- the two lines of LlamaMLP.forward
        (vllm/model_executor/models/llama.py:77-78) as one call, so that the activation can run
        in the GEMM's epilogue.  The first call channel-pairs the loaded weight; from then on
        the layer serves this method only."""
        if len(self.output_sizes) != 2 or self.output_sizes[0] != self.output_sizes[1]:
            raise ValueError("forward_silu_and_mul needs a [gate, up] pair of equal widths")
        if self.bias is not None or self.gather_output:
            raise ValueError("forward_silu_and_mul: no bias, no gathered output")
        method = self.linear_method
        if not hasattr(method, "apply_weights_silu_and_mul"):
            raise ValueError("the linear method has no fused SiLU-and-multiply path")
        return method.apply_weights_silu_and_mul(self.linear_weights, input_)

    def _piece_plan(self, piece_id):
        if not 0 <= piece_id < len(self.output_sizes):
            raise AssertionError(f"shard id {piece_id} out of range")
        tp_size = get_tensor_model_parallel_world_size()
        offset = sum(self.output_sizes[:piece_id]) // tp_size
        return offset, self.output_sizes[piece_id] // tp_size, get_tensor_model_parallel_rank()


class QKVParallelLinear(ColumnParallelLinear):
    """Fused q/k/v projection, sharded by attention head.  With fewer k/v heads than ranks
    each k/v head is replicated on ``tp_size / num_kv_heads`` consecutive ranks."""

    def __init__(self, hidden_size: int, head_size: int, total_num_heads: int,
                 total_num_kv_heads: Optional[int] = None, bias: bool = True,
                 skip_bias_add: bool = False, params_dtype: Optional[torch.dtype] = None,
                 linear_method: Optional[LinearMethodBase] = None):
        self.hidden_size, self.head_size = hidden_size, head_size
        self.total_num_heads = total_num_heads
        self.total_num_kv_heads = total_num_heads if total_num_kv_heads is None else total_num_kv_heads
        tp_size = get_tensor_model_parallel_world_size()
        self.num_heads = divide(self.total_num_heads, tp_size)
        if tp_size >= self.total_num_kv_heads:
            self.num_kv_heads = 1
            self.num_kv_head_replicas = divide(tp_size, self.total_num_kv_heads)
        else:
            self.num_kv_heads = divide(self.total_num_kv_heads, tp_size)
            self.num_kv_head_replicas = 1
        output_size = (self.num_heads + 2 * self.num_kv_heads) * tp_size * head_size
        super().__init__(hidden_size, output_size, bias, False, skip_bias_add, params_dtype,
                         linear_method)

    def _pieces(self):
        q = self.total_num_heads * self.head_size
        kv = self.total_num_kv_heads * self.head_size
        return [("q", q), ("k", kv), ("v", kv)]

    def _piece_plan(self, piece_id):
        if piece_id not in ("q", "k", "v"):
            raise AssertionError(f"unknown shard id {piece_id!r}")
        rank = get_tensor_model_parallel_rank()
        q, kv = self.num_heads * self.head_size, self.num_kv_heads * self.head_size
        if piece_id == "q":
            return 0, q, rank
        offset = q if piece_id == "k" else q + kv
        return offset, kv, rank // self.num_kv_head_replicas


class RowParallelLinear(_LinearBase):
    """Y = sum_r X_r A_r: rank r holds the K rows A_r and the matching slice of the input;
    the partial products are summed over the tensor-parallel group after the GEMM."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True,
                 input_is_parallel: bool = True, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None, reduce_results: bool = True,
                 linear_method: Optional[LinearMethodBase] = None):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.input_is_parallel = input_is_parallel
        self.reduce_results = reduce_results
        self.skip_bias_add = skip_bias_add
        self.tp_size = get_tensor_model_parallel_world_size()
        self.input_size_per_partition = divide(input_size, self.tp_size)
        if not reduce_results and bias and not skip_bias_add:
            raise ValueError("When not reduce the results, adding bias to the results can lead "
                             "to incorrect results")
        self._setup(linear_method, self.input_size_per_partition, output_size, input_size,
                    output_size, params_dtype, loader=self.weight_loader)
        if bias:
            self._make_bias(output_size, loader=self.weight_loader)
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor) -> None:
        input_dim = getattr(param, "input_dim", None)      # the bias has none: replicated
        size = param.data.shape[input_dim] if input_dim is not None else 0
        _copy_slice(param, loaded_weight, input_dim, 0, size, get_tensor_model_parallel_rank())

    def forward(self, input_: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        if not self.input_is_parallel:
            input_ = split_tensor_along_last_dim(input_, self.tp_size)[
                get_tensor_model_parallel_rank()].contiguous()
        out = None
        if self.reduce_results and self.tp_size > 1:
            fused = getattr(self.linear_method, "apply_weights_all_reduce", None)
            if fused is not None:                 # GEMM + all-reduce in one kernel where it applies
                out = fused(self.linear_weights, input_)
            if out is None:
                out = tensor_model_parallel_all_reduce(
                    self.linear_method.apply_weights(self.linear_weights, input_))
        else:
            out = self.linear_method.apply_weights(self.linear_weights, input_)
        if self.skip_bias_add:
            return out, self.bias
        return (out + self.bias if self.bias is not None else out), None
