from .communication_op import tensor_model_parallel_all_gather, tensor_model_parallel_all_reduce
from .parallel_state import (
    GroupCoordinator,
    destroy_distributed_environment,
    destroy_model_parallel,
    get_tensor_model_parallel_rank,
    get_tensor_model_parallel_world_size,
    get_tp_group,
    init_distributed_environment,
    initialize_model_parallel,
    model_parallel_is_initialized,
)
from .utils import divide, ensure_divisibility, split_tensor_along_last_dim

__all__ = [
    "GroupCoordinator", "destroy_distributed_environment", "destroy_model_parallel", "get_tensor_model_parallel_rank",
    "get_tensor_model_parallel_world_size", "get_tp_group", "init_distributed_environment",
    "initialize_model_parallel", "model_parallel_is_initialized", "divide", "ensure_divisibility",
    "split_tensor_along_last_dim", "tensor_model_parallel_all_gather", "tensor_model_parallel_all_reduce",
]
