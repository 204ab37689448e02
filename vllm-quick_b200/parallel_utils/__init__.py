"""Tensor-parallel plumbing of the linear path (reference: vllm/model_executor/parallel_utils/)."""
from . import custom_all_reduce  # noqa: F401
from .communication_op import (tensor_model_parallel_all_gather,  # noqa: F401
                               tensor_model_parallel_all_reduce)
from .parallel_state import (destroy_model_parallel, get_tensor_model_parallel_group,  # noqa: F401
                             get_tensor_model_parallel_rank,
                             get_tensor_model_parallel_world_size, initialize_model_parallel,
                             model_parallel_is_initialized)
