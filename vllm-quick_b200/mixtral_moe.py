"""Quantized Mixtral expert path: ``MixtralMLP`` and ``MixtralMoE`` with the reference's names,
constructor arguments, parameter names and result
(vllm/model_executor/models/mixtral_quant.py:57-160), over the W4A16 linear method.

What is the same: experts are split over the tensor-parallel ranks with ``np.array_split``
(:113-118), every local expert is three ``ReplicatedLinear`` (w1, w2, w3), the router is an
unquantized ``gate`` matrix, routing is softmax -> top-k -> renormalise (:138-142), an expert's
output is scaled by the routing weight of the tokens that chose it and zero for the others
(:147-151), one all-reduce per block (:159).

What is B200-native: the reference runs every local expert on every token whether or not a
token chose it.  At decode batch sizes the three GEMMs of an expert are pure weight streaming,
so an expert nobody chose costs as much as one everybody chose.  Here the number of tokens
routed to each expert is computed on the device and handed to the GEMM launches
(``ops.expert_gate`` -> ``b200_w4a16_gemm_gated``): a launch whose count is zero streams no
weights and writes zeros.  No host synchronisation, so the block stays CUDA-graph capturable.
The result is identical: such an expert's output is multiplied by a zero routing weight.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn
from torch.nn.parameter import Parameter

from . import ops
from .linear import LinearMethodBase, set_weight_attrs
from .parallel_linear import ReplicatedLinear
from .parallel_utils.communication_op import tensor_model_parallel_all_reduce
from .parallel_utils.parallel_state import (get_tensor_model_parallel_rank,
                                            get_tensor_model_parallel_world_size)


class MixtralMLP(nn.Module):
    """One expert: ``w2(silu(w1 x) * w3 x)`` (mixtral_quant.py:57-95)."""

    def __init__(self, num_experts: int, hidden_size: int, intermediate_size: int,
                 linear_method: Optional[LinearMethodBase] = None) -> None:
        super().__init__()
        self.num_experts = num_experts
        self.ffn_dim = intermediate_size
        self.hidden_dim = hidden_size
        self.w1 = ReplicatedLinear(self.hidden_dim, self.ffn_dim, bias=False,
                                   linear_method=linear_method)
        self.w2 = ReplicatedLinear(self.ffn_dim, self.hidden_dim, bias=False,
                                   linear_method=linear_method)
        self.w3 = ReplicatedLinear(self.hidden_dim, self.ffn_dim, bias=False,
                                   linear_method=linear_method)
        self.act_fn = nn.SiLU()

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        w1_out, _ = self.w1(hidden_states)
        w1_out = self.act_fn(w1_out)
        w3_out, _ = self.w3(hidden_states)
        current_hidden_states = w1_out * w3_out
        current_hidden_states, _ = self.w2(current_hidden_states)
        return current_hidden_states


class _RouterGate(nn.Module):
    """The router: an unquantized ``[num_experts, hidden]`` matrix (the reference builds it as
    ``ReplicatedLinear(..., linear_method=None)``, mixtral_quant.py:128-131).  A plain library
    GEMM of a few KFLOP; same parameter name (``weight``) so checkpoints load unchanged."""

    def __init__(self, hidden_size: int, num_experts: int,
                 params_dtype: Optional[torch.dtype] = None) -> None:
        super().__init__()
        dtype = params_dtype if params_dtype is not None else torch.get_default_dtype()
        self.weight = Parameter(torch.empty(num_experts, hidden_size, dtype=dtype),
                                requires_grad=False)
        set_weight_attrs(self.weight, {"input_dim": 1, "output_dim": 0})
        self.register_parameter("bias", None)

    def forward(self, x: torch.Tensor):
        return F.linear(x, self.weight), None


class MixtralMoE(nn.Module):
    """Mixture-of-experts block (mixtral_quant.py:98-160).  ``config`` needs
    ``num_local_experts``, ``num_experts_per_tok``, ``hidden_size``, ``intermediate_size``
    (transformers' MixtralConfig has them).  ``skip_idle_experts=False`` gives the reference's
    dense behaviour (every local expert streams its weights for every batch)."""

    def __init__(self, config, linear_method: Optional[LinearMethodBase] = None,
                 skip_idle_experts: bool = True):
        super().__init__()
        self.config = config
        self.rank = get_tensor_model_parallel_rank()
        self.tp_size = get_tensor_model_parallel_world_size()
        self.num_total_experts = config.num_local_experts
        self.top_k = config.num_experts_per_tok
        self.skip_idle_experts = skip_idle_experts
        if self.tp_size > self.num_total_experts:
            raise ValueError(
                f"Tensor parallel size {self.tp_size} is greater than "
                f"the number of experts {self.num_total_experts}.")
        # Split experts equally between ranks
        self.expert_indicies = np.array_split(range(
            self.num_total_experts), self.tp_size)[self.rank].tolist()
        if not self.expert_indicies:
            raise ValueError(
                f"Rank {self.rank} has no experts assigned to it.")
        self.experts = nn.ModuleList([
            MixtralMLP(self.num_total_experts, config.hidden_size, config.intermediate_size,
                       linear_method=linear_method)
            if idx in self.expert_indicies else None
            for idx in range(self.num_total_experts)
        ])
        self.gate = _RouterGate(config.hidden_size, self.num_total_experts)

    def route(self, hidden_states: torch.Tensor):
        """softmax -> top-k -> renormalise, as the reference (:136-142).  Returns the routing
        weights ``[T, top_k]`` fp32, the chosen experts ``[T, top_k]`` and, per expert, how
        many tokens chose it ``[num_experts]`` int32 (device; feeds the GEMM gates)."""
        router_logits, _ = self.gate(hidden_states)
        routing_weights = F.softmax(router_logits, dim=1, dtype=torch.float)
        routing_weights, selected_experts = torch.topk(routing_weights, self.top_k, dim=-1)
        routing_weights /= routing_weights.sum(dim=-1, keepdim=True)
        counts = torch.zeros(self.num_total_experts, dtype=torch.int32,
                             device=hidden_states.device)
        counts.scatter_add_(0, selected_experts.reshape(-1),
                            torch.ones_like(selected_experts.reshape(-1), dtype=torch.int32))
        return routing_weights, selected_experts, counts

    # Large batches outside CUDA-graph capture (prefill): the tokens of each local expert are
    # gathered, the expert runs on M_e rows instead of T, the weighted result is scattered back
    # (SURVEY 8f-4).  Needs the per-expert counts on the host (ONE sync per block), which a
    # captured decode step cannot have -- that one keeps the dense loop with device-side gates
    # below.  Threshold: where the routed path overtakes the dense loop on B200
    # (profiles/r1e_moe_prefill_bench.jsonl).
    routed_min_tokens = 512

    def _forward_routed(self, hidden_states, routing_weights, selected_experts, counts):
        final = torch.zeros_like(hidden_states)
        order = torch.argsort(selected_experts.reshape(-1), stable=True)   # by expert, token order kept
        tokens = (order // self.top_k).to(torch.int32)
        weights = routing_weights.reshape(-1)[order].contiguous()          # fp32, one per (token, choice)
        host_counts = counts.tolist()                                      # the one host sync
        start = 0
        for expert_idx in range(self.num_total_experts):
            n = host_counts[expert_idx]
            if n and expert_idx in self.expert_indicies:
                idx = tokens[start:start + n]
                out = self.experts[expert_idx](ops.moe_gather(hidden_states, idx))
                ops.moe_scatter_add(final, out, idx, weights[start:start + n])
            start += n
        return final

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        batch_size, sequence_length, hidden_dim = hidden_states.shape
        hidden_states = hidden_states.view(-1, hidden_dim)
        routing_weights, selected_experts, counts = self.route(hidden_states)

        if (self.skip_idle_experts and hidden_states.is_cuda
                and hidden_states.shape[0] >= self.routed_min_tokens
                and not torch.cuda.is_current_stream_capturing()):
            final_hidden_states = self._forward_routed(hidden_states.contiguous(), routing_weights,
                                                       selected_experts, counts)
            return tensor_model_parallel_all_reduce(final_hidden_states).view(
                batch_size, sequence_length, hidden_dim)

        final_hidden_states = None
        for expert_idx in self.expert_indicies:
            expert_layer = self.experts[expert_idx]
            expert_mask = (selected_experts == expert_idx)
            expert_weights = (routing_weights * expert_mask).sum(dim=-1, keepdim=True)
            gate = counts[expert_idx:expert_idx + 1] if (
                self.skip_idle_experts and hidden_states.is_cuda) else None
            with ops.expert_gate(gate):
                current_hidden_states = expert_layer(hidden_states)
            current_hidden_states = current_hidden_states.mul_(expert_weights)
            if final_hidden_states is None:
                final_hidden_states = current_hidden_states
            else:
                final_hidden_states.add_(current_hidden_states)

        return tensor_model_parallel_all_reduce(final_hidden_states).view(
            batch_size, sequence_length, hidden_dim)
