"""Tensor-parallel group bookkeeping.

Same entry points as the reference (vllm/model_executor/parallel_utils/parallel_state.py:22-86,
223-245) for the part the linear path uses: one process per GPU, ranks of a node grouped into
tensor-parallel groups of consecutive ranks.  Pipeline parallelism is outside the hot path
(SURVEY.md section 8): ``pipeline_model_parallel_size`` is accepted for signature parity and
must divide the world, but only the tensor group is built.

Without ``torch.distributed`` (single GPU) every query answers for a world of one.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

_TP_GROUP: Optional[dist.ProcessGroup] = None
_TP_RANKS: Optional[list] = None


def initialize_model_parallel(tensor_model_parallel_size: int = 1,
                              pipeline_model_parallel_size: int = 1) -> None:
    """Build the tensor-parallel group this rank belongs to (collective: every rank calls it)."""
    global _TP_GROUP, _TP_RANKS
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed must be initialised before model parallelism")
    world = dist.get_world_size()
    if world != tensor_model_parallel_size * pipeline_model_parallel_size:
        raise RuntimeError(f"world_size ({world}) is not equal to tensor_model_parallel_size "
                           f"({tensor_model_parallel_size}) x pipeline_model_parallel_size "
                           f"({pipeline_model_parallel_size})")
    if _TP_GROUP is not None:
        raise AssertionError("tensor model parallel group is already initialized")
    rank = dist.get_rank()
    for first in range(0, world, tensor_model_parallel_size):
        ranks = list(range(first, first + tensor_model_parallel_size))
        group = dist.new_group(ranks)          # every rank creates every group, in the same order
        if rank in ranks:
            _TP_GROUP, _TP_RANKS = group, ranks


def model_parallel_is_initialized() -> bool:
    return _TP_GROUP is not None


def get_tensor_model_parallel_group() -> dist.ProcessGroup:
    if _TP_GROUP is None:
        raise AssertionError("tensor model parallel group is not initialized")
    return _TP_GROUP


def get_tensor_model_parallel_world_size() -> int:
    if _TP_GROUP is None:
        return 1
    return dist.get_world_size(group=_TP_GROUP)


def get_tensor_model_parallel_rank() -> int:
    if _TP_GROUP is None:
        return 0
    return dist.get_rank(group=_TP_GROUP)


def get_tensor_model_parallel_src_rank() -> int:
    """Global rank of the first member of this rank's tensor-parallel group."""
    return _TP_RANKS[0] if _TP_RANKS else 0


def get_tensor_model_parallel_global_ranks() -> list:
    return list(_TP_RANKS) if _TP_RANKS else [0]


def destroy_model_parallel() -> None:
    global _TP_GROUP, _TP_RANKS
    if _TP_GROUP is not None and dist.is_initialized():
        dist.destroy_process_group(_TP_GROUP)
    _TP_GROUP, _TP_RANKS = None, None
