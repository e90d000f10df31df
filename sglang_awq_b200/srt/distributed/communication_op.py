"""reference: python/sglang/srt/distributed/communication_op.py:11-20."""
import torch

from .parallel_state import get_tp_group


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """Sum over the TP group (closes RowParallelLinear, linear.py:1407-1408)."""
    return get_tp_group().all_reduce(input_)


def tensor_model_parallel_all_gather(input_: torch.Tensor, dim: int = -1) -> torch.Tensor:
    return get_tp_group().all_gather(input_, dim)
