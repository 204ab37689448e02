"""Quantized Mixtral expert path over the W4A16 linear method -- the caller of the hot path in
BASELINE config 5 (reference: vllm/model_executor/models/mixtral_quant.py:57-160).

Kept from the reference, because checkpoints and model code depend on it: the class names and
constructor arguments, the parameter names (``experts.<i>.w1|w2|w3.{qweight,qzeros,scales}``,
``gate.weight``), which experts a tensor-parallel rank owns (contiguous blocks, the remainder
to the first ranks -- what ``np.array_split`` gives, :113-118), the routing arithmetic
(softmax in fp32 -> top-k -> renormalise, :136-142) and the result: every local expert's output
scaled by the routing weight of the tokens that chose it, summed, all-reduced (:143-159).

What runs instead of the reference's three GEMMs + four eager tensor ops per expert:

* w1 and w3 read the same activations (:87-93): after loading they are ONE merged, channel-paired
  ``[hidden, 2 * ffn]`` weight; ``silu(w1 x) * w3 x`` comes out of that GEMM's epilogue at decode
  batches (``b200_w4a16_gemm_silu_mul_gated``), GEMM + one activation kernel above them.
* the routing-weight multiply and the accumulation over experts (``mul_`` / ``add_``, :152-157)
  are one kernel (``b200_w4a16_moe_scatter_add``: ``acc[t] = half(acc[t] + half(y[t] * w[t]))``,
  the same two roundings).
* the reference streams every local expert's weights for every batch.  At decode sizes that is
  the whole cost, so the number of tokens routed to each expert is computed on the device and
  handed to the GEMM launches (``ops.expert_gate``): an expert nobody chose streams nothing
  and contributes zeros.  No host synchronisation: the block stays CUDA-graph capturable.
* large prefill batches (>= 8192 tokens) outside graph capture gather each expert's tokens, run
  the expert on M_e rows instead of T, and scatter-add the weighted result (one host
  read of the counts).
* from 65 tokens up to there, and under graph capture at any size, the same routed computation
  runs with the per-expert row counts read ON THE DEVICE by every kernel (``ops.quick_gemm_rows``
  and friends): no host synchronisation.
* decode batches: the router is one kernel, all local experts with tokens run as ONE grouped
  launch per projection, one pass accumulates their weighted outputs.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

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


def local_expert_ids(num_experts: int, tp_size: int, rank: int) -> List[int]:
    """Experts owned by ``rank``: contiguous blocks, the first ``num_experts % tp_size`` ranks
    one longer (identical to ``np.array_split(range(num_experts), tp_size)[rank]``)."""
    base, extra = divmod(num_experts, tp_size)
    first = rank * base + min(rank, extra)
    return list(range(first, first + base + (1 if rank < extra else 0)))


class MixtralMLP(nn.Module):
    """One expert: ``w2(silu(w1 x) * w3 x)`` (mixtral_quant.py:57-95).  w1 / w2 / w3 are the
    checkpoint's parameters; on CUDA the first forward merges w1 and w3 (and frees them)."""

    def __init__(self, num_experts: int, hidden_size: int, intermediate_size: int,
                 linear_method: Optional[LinearMethodBase] = None) -> None:
        super().__init__()
        self.num_experts = num_experts
        self.ffn_dim = intermediate_size
        self.hidden_dim = hidden_size
        self.w1 = ReplicatedLinear(hidden_size, intermediate_size, bias=False, linear_method=linear_method)
        self.w2 = ReplicatedLinear(intermediate_size, hidden_size, bias=False, linear_method=linear_method)
        self.w3 = ReplicatedLinear(hidden_size, intermediate_size, bias=False, linear_method=linear_method)
        self.linear_method = linear_method
        self._w13: Optional[Dict[str, Any]] = None      # merged [w1 | w3], built after loading

    def process_weights_after_loading(self) -> None:
        """Build the merged gate/up weight from the loaded w1 and w3 (checkpoint layout, columns
        concatenated), hand it to the linear method for pairing + interleave, release the two
        originals.  Idempotent; called by the first CUDA forward if the loader did not."""
        if self._w13 is not None:
            return
        a, b = self.w1.linear_weights, self.w3.linear_weights
        if a.get("qweight_layout", "awq") != "awq" or b.get("qweight_layout", "awq") != "awq":
            raise RuntimeError("w1 / w3 were already interleaved separately: merge them right after loading")
        merged = {k: torch.cat([a[k].data, b[k].data], dim=1).contiguous() for k in ("qweight", "qzeros", "scales")}
        self._w13 = {k: Parameter(v, requires_grad=False) for k, v in merged.items()}
        self._w13["qweight_layout"] = "awq"
        self.linear_method.pair_gate_up(self._w13)
        self.linear_method._interleave_once(self._w13)
        for w in (a, b):                                    # the merged copy is the live one
            for k in ("qweight", "qzeros", "scales"):
                w[k].data = torch.empty(0, dtype=w[k].dtype, device=w[k].device)
        self.linear_method._interleave_once(self.w2.linear_weights)

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        fused = hidden_states.is_cuda and hasattr(self.linear_method, "apply_weights_silu_and_mul")
        if fused:
            self.process_weights_after_loading()
            act = self.linear_method.apply_weights_silu_and_mul(self._w13, hidden_states)
        else:
            # host-side tests drive this block with a CPU stand-in for the linear method
            gate, _ = self.w1(hidden_states)
            up, _ = self.w3(hidden_states)
            act = F.silu(gate) * up
        out, _ = self.w2(act)
        return out


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

    # Batches of at least this many tokens, outside CUDA-graph capture, take the HOST-routed path
    # (one read of the per-expert counts; every expert GEMM then gets the plan of its exact row
    # count, CTA pairs included).  Below, the device-routed path is ahead -- no synchronisation:
    # 1.1-1.8x at 512-4096 tokens, level at 8192 (profiles/r3y_moe_prefill_epr{1,8}.jsonl).
    routed_min_tokens = 8192
    # Up to here (the decode kernel's batch sizes) the local experts run as ONE grouped launch
    # per projection instead of one launch per expert.
    grouped_max_tokens = 16
    # From here up (below routed_min_tokens, or at any size under CUDA-graph capture) the routed
    # path runs with the row counts read on the device: even at 32-64 tokens, 1.37x ahead from 96
    # (profiles/r3p_moe_mid_epr{1,8}.jsonl, r3q_moe_mid_epr1.jsonl).
    routed_on_device_min_tokens = 65

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
        # attribute name (spelling included) as in the reference: model code reads it
        self.expert_indicies = local_expert_ids(self.num_total_experts, self.tp_size, self.rank)
        if not self.expert_indicies:
            raise ValueError(f"Rank {self.rank} has no experts assigned to it.")
        # a slot per global expert id so that parameter names match the checkpoint's
        slots: List[Optional[nn.Module]] = [None] * self.num_total_experts
        for e in self.expert_indicies:
            slots[e] = MixtralMLP(self.num_total_experts, config.hidden_size,
                                  config.intermediate_size, linear_method=linear_method)
        self.experts = nn.ModuleList(slots)
        self.gate = _RouterGate(config.hidden_size, self.num_total_experts)
        self._token_ids: Dict[Any, torch.Tensor] = {}       # arange(T) int32 per (device, T)

    def route(self, hidden_states: torch.Tensor):
        """softmax -> top-k -> renormalise, as the reference (:136-142).  Returns the routing
        weights ``[T, top_k]`` fp32, the chosen experts ``[T, top_k]`` and, per expert, how
        many tokens chose it ``[num_experts]`` int32 (device; feeds the GEMM gates)."""
        return self._route(hidden_states)[:3]

    def _route(self, hidden_states: torch.Tensor):
        """``route`` plus the dense ``[T, num_experts]`` weight matrix where the router kernel
        produced it (None otherwise).  On the GPU the whole router is one launch
        (``ops.moe_route``) instead of a library GEMM and ten elementwise kernels -- at decode
        batches those cost as much as streaming two experts."""
        if (hidden_states.is_cuda and hidden_states.dtype == torch.float16
                and self.gate.weight.dtype == torch.float16 and self.num_total_experts <= 32
                and hidden_states.shape[1] % 8 == 0 and hidden_states.is_contiguous()):
            return ops.moe_route(hidden_states, self.gate.weight.data, self.top_k)
        router_logits, _ = self.gate(hidden_states)
        routing_weights = F.softmax(router_logits, dim=1, dtype=torch.float)
        routing_weights, selected_experts = torch.topk(routing_weights, self.top_k, dim=-1)
        routing_weights /= routing_weights.sum(dim=-1, keepdim=True)
        counts = torch.zeros(self.num_total_experts, dtype=torch.int32,
                             device=hidden_states.device)
        counts.scatter_add_(0, selected_experts.reshape(-1),
                            torch.ones_like(selected_experts.reshape(-1), dtype=torch.int32))
        return routing_weights, selected_experts, counts, None

    def _all_tokens(self, device: torch.device, tokens: int) -> torch.Tensor:
        key = (device, tokens)
        ids = self._token_ids.get(key)
        if ids is None:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("run the MoE block once eagerly at this batch size before capturing it")
            ids = torch.arange(tokens, dtype=torch.int32, device=device)
            self._token_ids[key] = ids
        return ids

    def _forward_routed(self, hidden_states, routing_weights, selected_experts, counts, per_token=None):
        """Gather each local expert's tokens, run it on M_e rows, scatter-add the weighted rows."""
        final = torch.zeros_like(hidden_states)
        order = torch.argsort(selected_experts.reshape(-1), stable=True)   # by expert, token order kept
        tokens = (order // self.top_k).to(torch.int32)
        weights = routing_weights.reshape(-1)[order].contiguous()          # fp32, one per (token, choice)
        host_counts = counts.tolist()                                      # the one host sync
        start = 0
        for expert_idx in range(self.num_total_experts):
            n = host_counts[expert_idx]
            if n and self.experts[expert_idx] is not None:
                idx = tokens[start:start + n]
                out = self.experts[expert_idx](ops.moe_gather(hidden_states, idx))
                ops.moe_scatter_add(final, out, idx, weights[start:start + n])
            start += n
        return final

    def _can_group(self, hidden_states: torch.Tensor) -> bool:
        """Decode batches on the GPU, at most eight local experts, the decode kernel's group sizes."""
        if not (hidden_states.is_cuda and self.skip_idle_experts and len(self.expert_indicies) <= 8
                and 0 < hidden_states.shape[0] <= self.grouped_max_tokens):
            return False
        mlp = self.experts[self.expert_indicies[0]]
        if not hasattr(mlp.linear_method, "apply_weights_silu_and_mul"):
            return False
        mlp.process_weights_after_loading()
        # the decode kernel takes quantization groups of whole 128-row tiles
        return all(k % 128 == 0 and (k // w["scales"].shape[0]) % 128 == 0
                   for w, k in ((mlp._w13, mlp.hidden_dim), (mlp.w2.linear_weights, mlp.ffn_dim)))

    def _forward_grouped(self, hidden_states, routing_weights, selected_experts, counts, per_token=None):
        """All local experts as two grouped launches + one accumulation pass: the
        tile walk of each launch covers the experts WITH tokens only (counts read on the
        device), so the SMs share the live weights evenly and idle experts cost nothing."""
        mlps = [self.experts[e] for e in self.expert_indicies]
        for m in mlps:
            m.process_weights_after_loading()
        first, n_local = self.expert_indicies[0], len(mlps)
        local_counts = counts[first:first + n_local]
        gate_up = [m._w13 for m in mlps]
        down = [m.w2.linear_weights for m in mlps]
        act = ops.moe_grouped_gemm(hidden_states, [w["qweight"].data for w in gate_up],
                                   [w["scales"].data for w in gate_up], [w["qzeros"].data for w in gate_up],
                                   local_counts, act=True)                       # [E, T, ffn]
        out = ops.moe_grouped_gemm(act, [w["qweight"].data for w in down], [w["scales"].data for w in down],
                                   [w["qzeros"].data for w in down], local_counts)  # [E, T, hidden]
        # routing weight of every (token, expert): each expert is chosen at most once per token
        if per_token is None:
            per_token = torch.zeros((hidden_states.shape[0], self.num_total_experts), dtype=torch.float32,
                                    device=hidden_states.device).scatter_(1, selected_experts, routing_weights)
        return ops.moe_combine(out, per_token[:, first:first + n_local], local_counts)

    def _forward_routed_on_device(self, hidden_states, routing_weights, selected_experts, counts, dense_w=None):
        """The routed path without the host ever seeing the routing (CUDA-graph capturable): the
        tokens of every local expert are compacted on the device, and gather, the expert's GEMMs,
        the activation and the scatter-add all take the live row count from ``counts`` on the
        device -- the GEMM derives its token tiles from it (``ops.quick_gemm_rows``), so an expert
        streams its weights once per LIVE token tile instead of once per tile of the batch."""
        tokens = hidden_states.shape[0]
        first, n_local = self.expert_indicies[0], len(self.expert_indicies)
        tok, w = ops.moe_compact(selected_experts, routing_weights, first, n_local)
        final = torch.zeros_like(hidden_states)
        for j, expert_idx in enumerate(self.expert_indicies):
            mlp = self.experts[expert_idx]
            mlp.process_weights_after_loading()
            rows = counts[expert_idx:expert_idx + 1]
            gate_up, down = mlp._w13, mlp.w2.linear_weights
            x_e = ops.moe_gather(hidden_states, tok[j], rows)
            h = ops.quick_gemm_rows(x_e, gate_up["qweight"].data, gate_up["scales"].data,
                                    gate_up["qzeros"].data, rows)
            act = torch.empty((tokens, h.shape[1] // 2), dtype=h.dtype, device=h.device)
            ops.silu_and_mul_paired(act, h, rows)
            out = ops.quick_gemm_rows(act, down["qweight"].data, down["scales"].data, down["qzeros"].data, rows)
            ops.moe_scatter_add(final, out, tok[j], w[j], rows)
        return final

    def _forward_dense(self, hidden_states, routing_weights, selected_experts, counts, dense_w=None):
        """Every local expert on every token (the reference's structure); experts without tokens
        are skipped on the device, the weighted accumulation is one kernel per expert."""
        on_gpu = hidden_states.is_cuda
        final = None
        for expert_idx in self.expert_indicies:
            per_token = (dense_w[:, expert_idx] if dense_w is not None else
                         (routing_weights * (selected_experts == expert_idx)).sum(dim=-1))   # fp32 [T]
            gate = counts[expert_idx:expert_idx + 1] if (self.skip_idle_experts and on_gpu) else None
            with ops.expert_gate(gate):
                out = self.experts[expert_idx](hidden_states)
            if on_gpu:
                if final is None:
                    final = torch.zeros_like(out)
                ops.moe_scatter_add(final, out, self._all_tokens(out.device, out.shape[0]),
                                    per_token.contiguous())
            else:                                   # host-side tests: the reference's two tensor ops
                out = out.mul_(per_token.unsqueeze(-1))
                final = out if final is None else final.add_(out)
        return final

    def select_path(self, tokens: int, on_gpu: bool, capturing: bool, can_group, fused_method: bool) -> str:
        """Which body runs a batch of ``tokens``: ``"grouped"`` (decode batches, one launch per
        projection over the local experts), ``"dense"`` (the reference's loop, idle experts gated
        on the device), ``"routed_on_device"`` (per-expert row counts read by the kernels) or
        ``"routed"`` (counts read once by the host; never under graph capture).  ``can_group`` is
        called only where it matters (it may interleave the weights)."""
        if not (on_gpu and self.skip_idle_experts):
            return "dense"
        if tokens >= self.routed_min_tokens and not capturing:
            return "routed"
        if can_group():
            return "grouped"
        if tokens >= self.routed_on_device_min_tokens and fused_method:
            return "routed_on_device"
        return "dense"

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        batch_size, sequence_length, hidden_dim = hidden_states.shape
        hidden_states = hidden_states.view(-1, hidden_dim)
        hidden_states = hidden_states.contiguous()
        routing_weights, selected_experts, counts, dense_w = self._route(hidden_states)
        on_gpu = hidden_states.is_cuda
        path = self.select_path(
            hidden_states.shape[0], on_gpu, on_gpu and torch.cuda.is_current_stream_capturing(),
            lambda: self._can_group(hidden_states),
            hasattr(self.experts[self.expert_indicies[0]].linear_method, "apply_weights_silu_and_mul"))
        body = getattr(self, "_forward_" + path)
        final = body(hidden_states, routing_weights, selected_experts, counts, dense_w)
        return tensor_model_parallel_all_reduce(final).view(batch_size, sequence_length, hidden_dim)
