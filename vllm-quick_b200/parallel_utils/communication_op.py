"""Collectives of the tensor-parallel linear path.

Reference: vllm/model_executor/parallel_utils/communication_op.py:17-41 (all-reduce after a
row-parallel GEMM, backend picked per call) and :44-70 (all-gather of a column-parallel
output).  Here the all-reduce has two backends:

* the NVLink peer-memory kernel of ``custom_all_reduce.py`` (one-shot, fixed rank order, runs
  on the compute stream and is CUDA-graph capturable) when it has been set up and the message
  fits -- the B200 counterpart of the reference's ``custom_all_reduce`` (which the reference
  disables for world_size > 1, vllm/config.py:403-410), and
* ``torch.distributed.all_reduce`` on the group's backend (NCCL over NVLink / NVSwitch on
  GPUs; gloo in the CPU tests) otherwise.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .parallel_state import (get_tensor_model_parallel_group, get_tensor_model_parallel_rank,
                             get_tensor_model_parallel_world_size)


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """Sum ``input_`` over the tensor-parallel group.  World of one: returned untouched.
    The result may or may not alias ``input_`` (as in the reference, callers use the return)."""
    if get_tensor_model_parallel_world_size() == 1:
        return input_
    from . import custom_all_reduce
    out = custom_all_reduce.maybe_all_reduce(input_)
    if out is not None:
        return out
    dist.all_reduce(input_, group=get_tensor_model_parallel_group())
    return input_


def tensor_model_parallel_all_gather(input_: torch.Tensor, dim: int = -1) -> torch.Tensor:
    """Concatenate ``input_`` of every rank along ``dim`` (rank order)."""
    world_size = get_tensor_model_parallel_world_size()
    if world_size == 1:
        return input_
    if not -input_.dim() <= dim < input_.dim():
        raise AssertionError(f"Invalid dim ({dim}) for input tensor with shape {tuple(input_.size())}")
    if dim < 0:
        dim += input_.dim()
    gathered = torch.empty((world_size,) + tuple(input_.shape), dtype=input_.dtype,
                           device=input_.device)
    # list-of-views form: supported by NCCL and by gloo (the CPU tests) alike
    dist.all_gather(list(gathered.unbind(0)), input_.contiguous(),
                    group=get_tensor_model_parallel_group())
    # [world, ..., d, ...] -> [..., world * d, ...]
    gathered = gathered.movedim(0, dim)
    shape = list(input_.shape)
    shape[dim] *= world_size
    return gathered.reshape(shape)


def split_tensor_along_last_dim(tensor: torch.Tensor, num_partitions: int):
    """Equal chunks of the last dimension (views), as RowParallelLinear needs when its input
    is not already sharded (reference: parallel_utils/utils.py:26-48)."""
    last = tensor.dim() - 1
    if tensor.size(last) % num_partitions != 0:
        raise AssertionError(f"{tensor.size(last)} is not divisible by {num_partitions}")
    return torch.split(tensor, tensor.size(last) // num_partitions, dim=last)


__all__ = ["tensor_model_parallel_all_reduce", "tensor_model_parallel_all_gather",
           "split_tensor_along_last_dim", "get_tensor_model_parallel_rank"]
